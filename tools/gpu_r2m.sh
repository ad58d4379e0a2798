# round 2, run M: solver bodies in shared memory
mkdir -p gpurun_out/r2m
python tools/phase_profile.py 160 60 > gpurun_out/r2m/time_base.txt 2>&1; tail -1 gpurun_out/r2m/time_base.txt
python tools/phase_profile.py 160 30 3 16384 > gpurun_out/r2m/time_3v3.txt 2>&1; tail -1 gpurun_out/r2m/time_3v3.txt
python -m pytest tests/test_env_gpu.py -q -x > gpurun_out/r2m/pytest_env.log 2>&1; tail -2 gpurun_out/r2m/pytest_env.log
