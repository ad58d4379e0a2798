# round 2, run J: per-block spread with and without arena->block balancing (profiling build)
mkdir -p gpurun_out/r2j
P=reinforcement-learning_b200
GGL_LIB_PATH=$P/libgigaenv_cuda_prof.so python tools/phase_profile.py 160 20 > gpurun_out/r2j/phase_bal.txt 2>&1; grep -E "ms/step|per-BLOCK|per-arena" gpurun_out/r2j/phase_bal.txt
GGL_BALANCE=0 GGL_LIB_PATH=$P/libgigaenv_cuda_prof.so python tools/phase_profile.py 160 20 > gpurun_out/r2j/phase_nobal.txt 2>&1; grep -E "ms/step|per-BLOCK|per-arena" gpurun_out/r2j/phase_nobal.txt
