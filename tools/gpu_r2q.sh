# round 2, run Q: phase profile + ncu capture of the current kernel
mkdir -p gpurun_out/r2q
P=reinforcement-learning_b200
GGL_LIB_PATH=$P/libgigaenv_cuda_prof.so python tools/phase_profile.py 160 20 > gpurun_out/r2q/phase.txt 2>&1; head -22 gpurun_out/r2q/phase.txt
ncu --set full --clock-control none --import-source on -k regex:env_step_kernel -s 166 -c 1 -f -o gpurun_out/r2q/env_step python bench.py --steps 5 --warmup 3 > gpurun_out/r2q/ncu.log 2>&1; tail -1 gpurun_out/r2q/ncu.log
