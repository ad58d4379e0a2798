# round 2, run I: arena->block load balancing (A/B with GGL_BALANCE=0), GAE with the look-back after the zero-carry replay
mkdir -p gpurun_out/r2i
python tools/phase_profile.py 160 60 > gpurun_out/r2i/time_bal.txt 2>&1; tail -1 gpurun_out/r2i/time_bal.txt
GGL_BALANCE=0 python tools/phase_profile.py 160 60 > gpurun_out/r2i/time_nobal.txt 2>&1; tail -1 gpurun_out/r2i/time_nobal.txt
python tools/phase_profile.py 160 30 3 16384 > gpurun_out/r2i/time_3v3_bal.txt 2>&1; tail -1 gpurun_out/r2i/time_3v3_bal.txt
GGL_BALANCE=0 python tools/phase_profile.py 160 30 3 16384 > gpurun_out/r2i/time_3v3_nobal.txt 2>&1; tail -1 gpurun_out/r2i/time_3v3_nobal.txt
python tools/phase_profile.py 160 30 2 8192 > gpurun_out/r2i/time_8192_bal.txt 2>&1; tail -1 gpurun_out/r2i/time_8192_bal.txt
python -m pytest tests/test_env_gpu.py -q -x > gpurun_out/r2i/pytest_env.log 2>&1; tail -2 gpurun_out/r2i/pytest_env.log
python -m pytest tests/test_gae_gpu.py -q > gpurun_out/r2i/pytest_gae.log 2>&1; tail -1 gpurun_out/r2i/pytest_gae.log
for i in 1 2; do python bench.py --workload gae --steps 30 --warmup 5 > gpurun_out/r2i/gae_$i.json 2> gpurun_out/r2i/gae_$i.err; python -c "import json; d=json.load(open('gpurun_out/r2i/gae_$i.json')); print('gae', d['ms_per_step'], d['value'])"; done
