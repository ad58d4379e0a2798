P=reinforcement-learning_b200
python -m pytest tests/test_env_gpu.py -m gpu -x -q 2>&1 | tail -3
python tools/phase_profile.py 160 20 2>&1 | head -1
GGL_LIB_PATH=$P/libgigaenv_cuda_prof.so python tools/phase_profile.py 160 20 2>&1 | tail -38 | head -18
