# round 2, run F: GAE early carry-out publication (A/B against GAE_CFG_EARLY=0), env timing with car-centric pads
mkdir -p gpurun_out/r2f
P=reinforcement-learning_b200
python -m pytest tests/test_gae_gpu.py -q > gpurun_out/r2f/pytest_gae.log 2>&1; tail -2 gpurun_out/r2f/pytest_gae.log
for v in "" _gaeE0 _gae128x8; do
  GGL_LIB_PATH=$P/libgigaenv_cuda$v.so python bench.py --workload gae --steps 30 --warmup 5 > gpurun_out/r2f/gae$v.json 2> gpurun_out/r2f/gae$v.err
  python -c "import json,sys; d=json.load(open('gpurun_out/r2f/gae$v.json')); print('gae$v', d['ms_per_step'], d['value'])"
done
python tools/phase_profile.py 160 60 > gpurun_out/r2f/time_base.txt 2>&1; tail -1 gpurun_out/r2f/time_base.txt
python tools/phase_profile.py 160 30 3 16384 > gpurun_out/r2f/time_3v3.txt 2>&1; tail -1 gpurun_out/r2f/time_3v3.txt
python -m pytest tests/test_env_gpu.py -q -x > gpurun_out/r2f/pytest_env.log 2>&1; tail -2 gpurun_out/r2f/pytest_env.log
ncu --set full --clock-control none --import-source on -k regex:gae_scan -s 6 -c 1 -f -o gpurun_out/r2f/gae_scan python bench.py --workload gae --steps 5 --warmup 3 > gpurun_out/r2f/ncu_gae.log 2>&1; tail -1 gpurun_out/r2f/ncu_gae.log
