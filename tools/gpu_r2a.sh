# round 2, run A: baseline timing, lane-group-16 variant, workload census, phase profiles
mkdir -p gpurun_out/r2a
P=reinforcement-learning_b200
python tools/phase_profile.py 160 60 > gpurun_out/r2a/time_base.txt 2>&1; tail -3 gpurun_out/r2a/time_base.txt
GGL_LIB_PATH=$P/libgigaenv_cuda_g16.so python tools/phase_profile.py 160 60 > gpurun_out/r2a/time_g16.txt 2>&1; tail -3 gpurun_out/r2a/time_g16.txt
GGL_LIB_PATH=$P/libgigaenv_cuda_g16.so python -m pytest tests/test_env_gpu.py -x -q > gpurun_out/r2a/pytest_g16.txt 2>&1; tail -3 gpurun_out/r2a/pytest_g16.txt
python tools/tick_stats.py 160 16 > gpurun_out/r2a/census.txt 2>&1; cat gpurun_out/r2a/census.txt
GGL_LIB_PATH=$P/libgigaenv_cuda_prof.so python tools/phase_profile.py 160 20 > gpurun_out/r2a/phase_base.txt 2>&1; head -30 gpurun_out/r2a/phase_base.txt
GGL_LIB_PATH=$P/libgigaenv_cuda_g16prof.so python tools/phase_profile.py 160 20 > gpurun_out/r2a/phase_g16.txt 2>&1; head -30 gpurun_out/r2a/phase_g16.txt
GGL_LIB_PATH=$P/libgigaenv_cuda_g16.so python tools/phase_profile.py 160 30 3 16384 > gpurun_out/r2a/time_g16_3v3.txt 2>&1; tail -2 gpurun_out/r2a/time_g16_3v3.txt
