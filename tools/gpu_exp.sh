P=reinforcement-learning_b200
for v in u1 u2 u4; do
GGL_LIB_PATH=$P/libgigaenv_var_$v.so python bench.py --workload expbuf 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v expbuf ms', d['ms_per_step'], 'GB/s', d['value'], 'frac', d['roofline']['frac'])"
done
