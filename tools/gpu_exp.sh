timeout 600 python -m pytest tests/test_expbuf_gpu.py -m gpu -x -q 2>&1 | tail -5
timeout 300 python bench.py --workload expbuf 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('staged expbuf ms', d['ms_per_step'], 'GB/s', d['value'], 'frac', d['roofline']['frac'], 'e2e', d['e2e']['value'])"
GGL_EXPBUF_IDENTITY=1 timeout 300 python bench.py --workload expbuf 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('staged identity ms', d['ms_per_step'], 'GB/s', d['value'], 'frac', d['roofline']['frac'])"
