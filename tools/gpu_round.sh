# round-end measurement pass: tests, bench lines, ncu launch list, ncu --set full of the hot kernels (steady state)
# usage: bash tools/gpu_round.sh [tag]   -> gpurun_out/<tag>/
T=${1:-round}
O=gpurun_out/$T
mkdir -p $O
python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"
python bench.py --steps 200 --warmup 20 > $O/bench_env.json 2> $O/bench_env.err; echo "bench rc=$?"
python bench.py --workload gae > $O/bench_gae.json 2> $O/bench_gae.err
python bench.py --workload expbuf > $O/bench_expbuf.json 2> $O/bench_expbuf.err
python bench.py --workload traj > $O/bench_traj.json 2> $O/bench_traj.err
python bench.py --workload obsstat > $O/bench_obsstat.json 2> $O/bench_obsstat.err
python bench.py --impl reference --steps 10 --warmup 3 > $O/bench_ref.json 2>/dev/null
python -c 'import __graft_entry__ as g; g.smoke()' > $O/smoke.log 2>&1; echo "smoke rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/launches.csv python bench.py --steps 5 --warmup 3 > $O/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:env_step_kernel -s 166 -c 1 -f -o $O/env_step python bench.py --steps 5 --warmup 3 > $O/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gae_scan -s 6 -c 1 -f -o $O/gae_scan python bench.py --workload gae --steps 5 --warmup 3 > $O/ncu3.log 2>&1
tail -2 $O/pytest_gpu.log; tail -3 $O/smoke.log; cat $O/bench_env.json
