# round 2, run H: GAE with pass 2 overlapped with the look-back
mkdir -p gpurun_out/r2h
P=reinforcement-learning_b200
python -m pytest tests/test_gae_gpu.py -q > gpurun_out/r2h/pytest_gae.log 2>&1; tail -2 gpurun_out/r2h/pytest_gae.log
for i in 1 2; do python bench.py --workload gae --steps 30 --warmup 5 > gpurun_out/r2h/gae_$i.json 2> gpurun_out/r2h/gae_$i.err; python -c "import json; d=json.load(open('gpurun_out/r2h/gae_$i.json')); print('gae', d['ms_per_step'], d['value'])"; done
GGL_LIB_PATH=$P/libgigaenv_cuda_gaeE0.so python bench.py --workload gae --steps 30 --warmup 5 > gpurun_out/r2h/gae_old.json 2>/dev/null; python -c "import json; d=json.load(open('gpurun_out/r2h/gae_old.json')); print('gae old', d['ms_per_step'], d['value'])"
ncu --set full --clock-control none --import-source on -k regex:gae_scan -s 6 -c 1 -f -o gpurun_out/r2h/gae_scan python bench.py --workload gae --steps 5 --warmup 3 > gpurun_out/r2h/ncu_gae.log 2>&1; tail -1 gpurun_out/r2h/ncu_gae.log
