# round-end measurement pass: tests, bench lines, ncu launch list, ncu --set full of the two hot kernels (steady state)
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
python bench.py --steps 200 --warmup 20 > gpurun_out/bench_env.json 2> gpurun_out/bench_env.err; echo "bench rc=$?"
python bench.py --workload gae > gpurun_out/bench_gae.json 2> gpurun_out/bench_gae.err
python bench.py --workload expbuf > gpurun_out/bench_expbuf.json 2> gpurun_out/bench_expbuf.err
python bench.py --workload traj > gpurun_out/bench_traj.json 2> gpurun_out/bench_traj.err
python bench.py --workload obsstat > gpurun_out/bench_obsstat.json 2> gpurun_out/bench_obsstat.err
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref.json 2>/dev/null
python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches.csv python bench.py --steps 5 --warmup 3 > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:env_step_kernel -s 166 -c 1 -f -o gpurun_out/env_step python bench.py --steps 5 --warmup 3 > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gae_scan -s 6 -c 1 -f -o gpurun_out/gae_scan python bench.py --workload gae --steps 5 --warmup 3 > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:expbuf_gather -s 6 -c 1 -f -o gpurun_out/expbuf python bench.py --workload expbuf --steps 5 --warmup 3 > gpurun_out/ncu4.log 2>&1
tail -2 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/smoke.log; cat gpurun_out/bench_env.json
