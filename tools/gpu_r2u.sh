# round 2, run U: weights of the work estimate behind the arena->block balancing
mkdir -p gpurun_out/r2u
P=reinforcement-learning_b200
for w in 4,6,1 8,6,1 4,12,1 4,6,0 4,6,4 1,2,4 0,0,1 1,0,0 16,24,1; do
  GGL_COST_W=$w GGL_LIB_PATH=$P/libgigaenv_cuda_cw.so python tools/phase_profile.py 160 60 > gpurun_out/r2u/time_$w.txt 2>&1; echo "w=$w $(tail -1 gpurun_out/r2u/time_$w.txt)"
done
GGL_BALANCE=0 GGL_LIB_PATH=$P/libgigaenv_cuda_cw.so python tools/phase_profile.py 160 60 > gpurun_out/r2u/time_nobal.txt 2>&1; echo "nobal $(tail -1 gpurun_out/r2u/time_nobal.txt)"
