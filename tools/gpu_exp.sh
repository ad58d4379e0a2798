P=reinforcement-learning_b200
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python tools/phase_profile.py 160 30 2>&1 | head -1
python tools/phase_profile.py 160 20 3 16384 2>&1 | head -1
python tools/phase_profile.py 160 20 1 256 2>&1 | head -1
python tools/phase_profile.py 160 20 2 8192 2>&1 | head -1
GGL_LIB_PATH=$P/libgigaenv_cuda_prof.so python tools/phase_profile.py 160 20 2>&1 | tail -38 | head -19
