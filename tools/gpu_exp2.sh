python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/bench_env_n2.json 2> gpurun_out/bench_env_n2.err; echo "rc=$?"; tail -1 gpurun_out/bench_env_n2.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 5 --warmup 2 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err; echo "rc=$?"; tail -1 gpurun_out/bench_ref_n2.json
tail -5 gpurun_out/bench_env_n2.err
