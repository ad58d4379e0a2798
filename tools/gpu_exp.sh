P=reinforcement-learning_b200
for v in pn; do
echo "=== gae variant $v"
GGL_LIB_PATH=$P/libgigaenv_var_$v.so timeout 300 python bench.py --workload gae 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('GAE ms', d['ms_per_step'], 'frac', d['roofline']['frac'])"
done
