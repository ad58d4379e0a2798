python -m pytest tests/test_gae_gpu.py -m gpu -x -q 2>&1 | tail -2
python bench.py --workload gae 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('GAE ms', d['ms_per_step'], 'frac', d['roofline']['frac'])"
python tools/e2e_breakdown.py 2>&1 | tail -7
