python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --steps 100 --warmup 10 > gpurun_out/bench_env2.json 2> gpurun_out/bench_env2.err; cat gpurun_out/bench_env2.json; tail -3 gpurun_out/bench_env2.err
python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref2.json 2>/dev/null; cat gpurun_out/bench_ref2.json
