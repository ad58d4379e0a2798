# round 2, run Z2: 4 lanes per arena (8 arenas per warp, 28 arenas per SM in 3.5 warps) against the product's 8
mkdir -p gpurun_out/r2z
P=reinforcement-learning_b200
for v in "" _g4; do
  GGL_LIB_PATH=$P/libgigaenv_cuda$v.so timeout 300 python tools/phase_profile.py 160 60 > gpurun_out/r2z/time2$v.txt 2>&1; echo "variant '$v' rc=$? $(tail -n 1 gpurun_out/r2z/time2$v.txt)"
done
GGL_LIB_PATH=$P/libgigaenv_cuda_g4.so timeout 600 python -m pytest tests/test_env_gpu.py -q -x -k "env_steps_match or steady_state_scripted or block_geometry" > gpurun_out/r2z/pytest_g4.log 2>&1; tail -n 3 gpurun_out/r2z/pytest_g4.log
