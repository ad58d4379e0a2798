# solver_order (the reference's manifold order inside an island): GPU tests, parity rates on the r2av sample, step time
mkdir -p gpurun_out/r2aw
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2aw/pytest_gpu.txt 2>&1; tail -n 3 gpurun_out/r2aw/pytest_gpu.txt
timeout 1500 python tools/parity_rates.py 1024 60 > gpurun_out/r2aw/parity_rates_1024.txt 2>&1; cut -c1-200 gpurun_out/r2aw/parity_rates_1024.txt
timeout 600 python bench.py --steps 40 --warmup 5 > gpurun_out/r2aw/bench.json 2>gpurun_out/r2aw/bench.err; cut -c1-400 gpurun_out/r2aw/bench.json
