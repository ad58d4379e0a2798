mkdir -p gpurun_out/r2ac
P=reinforcement-learning_b200
GGL_LIB_PATH=$P/libgigaenv_cuda_fma.so timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r2ac/pytest_fma.log 2>&1; tail -n 4 gpurun_out/r2ac/pytest_fma.log | cut -c1-300
for v in "" _fma; do
  GGL_LIB_PATH=$P/libgigaenv_cuda$v.so timeout 900 python tools/parity_rates.py 256 60 > gpurun_out/r2ac/rates$v.txt 2>&1; echo "== variant '$v'"; cat gpurun_out/r2ac/rates$v.txt | cut -c1-260
done
