"""BASELINE.json configs[3] / SURVEY 8(d) C4: GAE/returns sweep over T x N (stream = N segments of length T, last element
terminal) for three terminal mixes, device-timed through ggl_gae_compute, with the CPU side (the plain-C restatement of GAE.cpp,
oracle/gae_oracle.c, one host thread as the reference runs it) timed beside it on the streams it finishes in seconds
(L <= 2^22; the reference's truncation search is O(L * K), GAE.cpp:89-95). Streams smaller than L2 are timed with a 256 MiB
flush write between calls. GB/s = 21 B/element x L / time.   python tools/gae_sweep.py > profiles/<tag>_gae_sweep.txt"""
import importlib
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oracle_c   # noqa: E402

ggl = importlib.import_module("reinforcement-learning_b200")
dev = torch.device("cuda", 0)
PEAK = 6538.0
try:
    import json
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", PEAK))
except Exception:
    pass


def inputs(T, N, kind, seed=7):
    g = torch.Generator(device=dev); g.manual_seed(seed)
    L = T * N
    r = torch.randn(L, device=dev, generator=g)
    v = torch.randn(L, device=dev, generator=g)
    t = torch.zeros(L, dtype=torch.int8, device=dev)
    if kind == "normal":
        t[T - 1::T] = 1
    elif kind == "truncated":
        t[T - 1::T] = 2
    else:
        t[T - 1::T] = torch.randint(1, 3, (N,), device=dev, generator=g).to(torch.int8)
    K = int((t == 2).sum().item())
    tv = torch.randn(max(K, 1), device=dev, generator=g)[:K]
    return r, t, v, tv


def time_device(r, t, v, tv, reps):
    L = r.numel()
    out = tuple(torch.empty(L, device=dev) for _ in range(3))
    ws = None
    small = 21 * L < 200e6
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if small else None
    for _ in range(3):
        *_, ws = ggl.gae_compute_device(r, t, v, tv, 0.99, 0.95, 1.7, 10.0, out=out, workspace=ws)
    tot = 0.0
    for _ in range(reps):
        if small:
            flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ggl.gae_compute_device(r, t, v, tv, 0.99, 0.95, 1.7, 10.0, out=out, workspace=ws)
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / reps, out


def main():
    print("GAE sweep, gamma .99 lambda .95 returnStd 1.7 clip 10; GB/s at 21 B/element; peak %.0f GB/s; cpu = oracle/gae_oracle.c on 1 thread" % PEAK)
    print("%-10s %4s %8s %10s | %9s %8s %6s | %9s %8s | %8s  %s" % ("terminals", "T", "N", "L", "gpu ms", "GB/s", "frac", "cpu ms", "cpu GB/s", "speed-up", "max |adv diff| / max(1,|ref|)"))
    for kind in ("normal", "truncated", "mixed"):
        for T in (1, 4, 16, 64, 256):
            for N in (1 << 10, 1 << 14, 1 << 18, 1 << 20):
                L = T * N
                r, t, v, tv = inputs(T, N, kind)
                ms, out = time_device(r, t, v, tv, 10 if L < (1 << 26) else 4)
                gbs = 21.0 * L / (ms * 1e-3) / 1e9
                cpu_ms = cpu_gbs = err = float("nan")
                K = tv.numel()
                if L <= (1 << 22) and L * max(K, 1) <= (1 << 36):
                    hr, ht, hv, htv = r.cpu().numpy(), t.cpu().numpy(), v.cpu().numpy(), tv.cpu().numpy()
                    t0 = time.perf_counter()
                    adv, ret, tgt, clip, rc = oracle_c.gae(hr, ht, hv, htv, 0.99, 0.95, 1.7, 10.0)
                    cpu_ms = (time.perf_counter() - t0) * 1e3
                    cpu_gbs = 21.0 * L / (cpu_ms * 1e-3) / 1e9
                    err = float(np.max(np.abs(out[0].cpu().numpy() - adv) / np.maximum(1.0, np.abs(adv))))
                print("%-10s %4d %8d %10d | %9.4f %8.1f %6.3f | %9.2f %8.2f | %8.0f  %.1e" %
                      (kind, T, N, L, ms, gbs, gbs / PEAK, cpu_ms, cpu_gbs, cpu_ms / ms if cpu_ms == cpu_ms else float("nan"), err), flush=True)
                del r, t, v, tv, out
                torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
