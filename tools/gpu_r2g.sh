# round 2, run G: env timing + phase profile after the inline culls / free box / car-centric pads; env parity tests
mkdir -p gpurun_out/r2g
P=reinforcement-learning_b200
python tools/phase_profile.py 160 60 > gpurun_out/r2g/time_base.txt 2>&1; tail -1 gpurun_out/r2g/time_base.txt
python tools/phase_profile.py 160 30 3 16384 > gpurun_out/r2g/time_3v3.txt 2>&1; tail -1 gpurun_out/r2g/time_3v3.txt
python tools/phase_profile.py 160 30 1 256 > gpurun_out/r2g/time_1v1.txt 2>&1; tail -1 gpurun_out/r2g/time_1v1.txt
GGL_LIB_PATH=$P/libgigaenv_cuda_prof.so python tools/phase_profile.py 160 20 > gpurun_out/r2g/phase.txt 2>&1; head -18 gpurun_out/r2g/phase.txt
python -m pytest tests/test_env_gpu.py -q -x > gpurun_out/r2g/pytest_env.log 2>&1; tail -2 gpurun_out/r2g/pytest_env.log
