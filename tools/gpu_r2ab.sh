# round 2, run AB: what -fmad=false costs (the parity TUs are built without FMA contraction): timing and parity of an FMA build
mkdir -p gpurun_out/r2ab
P=reinforcement-learning_b200
for v in "" _fma; do
  GGL_LIB_PATH=$P/libgigaenv_cuda$v.so timeout 300 python tools/phase_profile.py 160 60 > gpurun_out/r2ab/time$v.txt 2>&1; echo "variant '$v' $(tail -n 1 gpurun_out/r2ab/time$v.txt)"
done
GGL_LIB_PATH=$P/libgigaenv_cuda_fma.so timeout 900 python -m pytest tests/test_env_gpu.py -q -s -k "steady_state or env_steps_match or multi_warp" > gpurun_out/r2ab/pytest_fma.log 2>&1; tail -n 12 gpurun_out/r2ab/pytest_fma.log | cut -c1-600
python -m pytest tests/test_env_gpu.py -q -s -k "steady_state" > gpurun_out/r2ab/pytest_base.log 2>&1; grep "steady state" gpurun_out/r2ab/pytest_base.log | cut -c1-600
