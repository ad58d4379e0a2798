# what solver_order costs: step time of the shipped build against a build without it (-DGGL_NO_SOLVER_ORDER), same box, twice each,
# and how often it is called / runs (-DGGL_ORDER_COUNT). Build the variants first:
#   python reinforcement-learning_b200/build.py --variant noorder -DGGL_NO_SOLVER_ORDER
#   python reinforcement-learning_b200/build.py --variant count -DGGL_ORDER_COUNT
mkdir -p gpurun_out/r2ay
P=reinforcement-learning_b200
for rep in 1 2; do
for v in "" _noorder; do
  GGL_LIB_PATH=$P/libgigaenv_cuda$v.so timeout 300 python tools/phase_profile.py 160 60 > gpurun_out/r2ay/time${v}_$rep.txt 2>&1; echo "variant '$v' rep $rep: $(tail -n 1 gpurun_out/r2ay/time${v}_$rep.txt)"
done
done
for v in "" _noorder; do
  GGL_LIB_PATH=$P/libgigaenv_cuda$v.so timeout 300 python tools/phase_profile.py 160 30 3 16384 > gpurun_out/r2ay/time3v3${v}.txt 2>&1; echo "3v3 variant '$v': $(tail -n 1 gpurun_out/r2ay/time3v3${v}.txt)"
done
GGL_LIB_PATH=$P/libgigaenv_cuda_count.so timeout 300 python tools/order_call_rate.py 2 4096 2>&1 | tail -n 1 | tee gpurun_out/r2ay/call_rate_2v2.txt
GGL_LIB_PATH=$P/libgigaenv_cuda_count.so timeout 300 python tools/order_call_rate.py 3 16384 2>&1 | tail -n 1 | tee gpurun_out/r2ay/call_rate_3v3.txt
