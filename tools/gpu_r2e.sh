# round 2, run E: env timing after the sink / gather changes; GAE tile-geometry variants
mkdir -p gpurun_out/r2e
P=reinforcement-learning_b200
python tools/phase_profile.py 160 60 > gpurun_out/r2e/time_base.txt 2>&1; tail -1 gpurun_out/r2e/time_base.txt
python tools/phase_profile.py 160 30 3 16384 > gpurun_out/r2e/time_3v3.txt 2>&1; tail -1 gpurun_out/r2e/time_3v3.txt
python -m pytest tests/test_env_gpu.py -q -x > gpurun_out/r2e/pytest_env.log 2>&1; tail -2 gpurun_out/r2e/pytest_env.log
for v in "" _gae128x8 _gae128x16i8 _gae512x2; do
  GGL_LIB_PATH=$P/libgigaenv_cuda$v.so python bench.py --workload gae --steps 30 --warmup 5 > gpurun_out/r2e/gae$v.json 2> gpurun_out/r2e/gae$v.err
  python -c "import json,sys; d=json.load(open('gpurun_out/r2e/gae$v.json')); print('gae$v', d['ms_per_step'], d['value'])"
done
GGL_LIB_PATH=$P/libgigaenv_cuda_gae128x8.so python -m pytest tests/test_gae_gpu.py -q > gpurun_out/r2e/pytest_gae128x8.log 2>&1; tail -1 gpurun_out/r2e/pytest_gae128x8.log
