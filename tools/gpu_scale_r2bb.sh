# weak-scaling pass on one 8-GPU box: bench.py at N = 1, 2, 4, 8 (4096 arenas per GPU), own arm only
mkdir -p gpurun_out/r2bb
python bench.py --gpus 1 --steps 100 --warmup 10 > gpurun_out/r2bb/scale_n1.json 2> gpurun_out/r2bb/scale_n1.err
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py --gpus $n --steps 100 --warmup 10 > gpurun_out/r2bb/scale_n$n.json 2> gpurun_out/r2bb/scale_n$n.err
done
python - <<'PY'
import json
base = None
for n in (1, 2, 4, 8):
    try:
        d = json.loads(open("gpurun_out/r2bb/scale_n%d.json" % n).read().strip().splitlines()[-1])
    except Exception as e:
        print(n, "failed", e); continue
    base = base or d["value"]
    print("N=%d value %.2f M agent-steps/s (%.1f %% of N x the 1-GPU value) e2e %.2f M ms/step %.3f" % (n, d["value"] / 1e6, 100 * d["value"] / (n * base), d["e2e"]["value"] / 1e6, d["ms_per_step"]))
PY
