# round 2, run B: no-barrier build: per-phase cycles without barrier inflation, ncu stall breakdown
mkdir -p gpurun_out/r2b
P=reinforcement-learning_b200
GGL_LIB_PATH=$P/libgigaenv_cuda_ps0prof.so python tools/phase_profile.py 160 20 > gpurun_out/r2b/phase_ps0.txt 2>&1; head -20 gpurun_out/r2b/phase_ps0.txt
GGL_LIB_PATH=$P/libgigaenv_cuda_ps0.so python tools/phase_profile.py 160 40 > gpurun_out/r2b/time_ps0.txt 2>&1; tail -1 gpurun_out/r2b/time_ps0.txt
GGL_LIB_PATH=$P/libgigaenv_cuda_ps0.so ncu --set full --clock-control none --import-source on -k regex:env_step_kernel -s 166 -c 1 -f -o gpurun_out/r2b/env_step_ps0 python bench.py --steps 5 --warmup 3 > gpurun_out/r2b/ncu.log 2>&1; tail -2 gpurun_out/r2b/ncu.log
