P=reinforcement-learning_b200
timeout 600 python -m pytest tests/test_expbuf_gpu.py tests/test_traj_gpu.py -m gpu -x -q 2>&1 | tail -3
for v in t32s2 t32s3 t16s3 t16s4; do
GGL_LIB_PATH=$P/libgigaenv_var_$v.so timeout 300 python -m pytest tests/test_expbuf_gpu.py -m gpu -x -q 2>&1 | tail -1
for r in 1 2; do
GGL_LIB_PATH=$P/libgigaenv_var_$v.so timeout 300 python bench.py --workload expbuf 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v expbuf ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],4))"
done
done
