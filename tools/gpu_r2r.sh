# round 2, run R: simplex scratch in shared memory; barrier placement after the solver got cheap
# masks: 2084 = rays + contacts + end, 2080 = contacts + end, 2212 = rays + contacts + solver + end (product), 2212|8 ...
mkdir -p gpurun_out/r2r
P=reinforcement-learning_b200
for v in "" _psA _psB _psC; do
  GGL_LIB_PATH=$P/libgigaenv_cuda$v.so python tools/phase_profile.py 160 60 > gpurun_out/r2r/time$v.txt 2>&1; echo "variant '$v'"; tail -1 gpurun_out/r2r/time$v.txt
done
python tools/phase_profile.py 160 30 3 16384 > gpurun_out/r2r/time_3v3.txt 2>&1; tail -1 gpurun_out/r2r/time_3v3.txt
python -m pytest tests/test_env_gpu.py -q -x > gpurun_out/r2r/pytest_env.log 2>&1; tail -2 gpurun_out/r2r/pytest_env.log
