P=reinforcement-learning_b200
for v in m2048 m2080 m2052 m2084 m2176 m2180 m2230 m3052; do
GGL_LIB_PATH=$P/libgigaenv_var_$v.so timeout 200 python bench.py --steps 200 --warmup 20 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v ms', round(d['ms_per_step'],4), 'value', round(d['value']/1e6,3), 'e2e', round(d['e2e']['value']/1e6,3), 'kernel_ms', d['roofline'].get('kernel_ms'))"
done
timeout 200 python bench.py --steps 200 --warmup 20 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('default ms', round(d['ms_per_step'],4), 'value', round(d['value']/1e6,3), 'e2e', round(d['e2e']['value']/1e6,3))"
