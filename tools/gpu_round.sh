python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
python bench.py --steps 100 --warmup 10 > gpurun_out/bench_env.json 2> gpurun_out/bench_env.err; echo "bench rc=$?"
python bench.py --workload gae > gpurun_out/bench_gae.json 2> gpurun_out/bench_gae.err
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/bench_ref.json 2>/dev/null
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 5 --warmup 3 > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:env_step_kernel -s 4 -c 1 -f -o gpurun_out/env_step python bench.py --steps 5 --warmup 3 > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_env.json
