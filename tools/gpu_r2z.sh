# round 2, run Z: phase barriers per SET of warps (GGL_PSYNC_GROUPS = 2, 3, 4) instead of per block
mkdir -p gpurun_out/r2z
P=reinforcement-learning_b200
for v in "" _pg2 _pg3 _pg4; do
  GGL_LIB_PATH=$P/libgigaenv_cuda$v.so timeout 300 python tools/phase_profile.py 160 60 > gpurun_out/r2z/time$v.txt 2>&1; echo "variant '$v' rc=$? $(tail -n 1 gpurun_out/r2z/time$v.txt)"
done
for v in "" _pg2; do
  GGL_LIB_PATH=$P/libgigaenv_cuda$v.so timeout 300 python tools/phase_profile.py 160 30 3 16384 > gpurun_out/r2z/time_3v3$v.txt 2>&1; echo "3v3 '$v' $(tail -n 1 gpurun_out/r2z/time_3v3$v.txt)"
done
