P=reinforcement-learning_b200
python -m pytest tests/test_expbuf_gpu.py tests/test_env_gpu.py -m gpu -x -q 2>&1 | tail -3
for v in u1 u2 u3; do
GGL_LIB_PATH=$P/libgigaenv_var_$v.so python bench.py --workload expbuf 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v expbuf ms', d['ms_per_step'], 'GB/s', d['value'], 'frac', d['roofline']['frac'])"
done
GGL_EXPBUF_IDENTITY=1 GGL_LIB_PATH=$P/libgigaenv_var_u2.so python bench.py --workload expbuf 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('u2 identity ms', d['ms_per_step'], 'GB/s', d['value'], 'frac', d['roofline']['frac'])"
