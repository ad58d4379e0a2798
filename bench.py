#!/usr/bin/env python
"""Benchmark of the B200-native rollout hot path (BASELINE.json metric: agent-steps/s, whole box; GAE GB/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload env|gae]

One "step" = one env step (Reset + StepFirstHalf + StepSecondHalf, tick_skip 8) of every arena on the rank, on
BASELINE.json configs[1]: soccar 2v2, 4096 arenas per GPU, AdvancedObs + {VelocityPlayerToBall, TouchBall, Goal},
device-side random mask-valid actions.  Arenas shard across ranks with no data-path collective ("weak" scaling:
4096 arenas per GPU).  `value` = agent-steps of all ranks / max-over-ranks device time (CUDA events on the launch
stream, L2 flushed between steps).  `e2e` = the same metric through the host-buffer C-ABI calls in the Learner's order
(ggl_envset_begin_step_host: Reset + D2H of obs/masks + first half launched; host action pick; ggl_envset_finish_step_host:
pinned H2D of actions, second half, D2H of rewards/terminals -- all inside the timed wall-clock region).
`--workload gae` benches the GAE scan instead (config C4: T=256 x N=65536 stream, GB/s at 21 B/element).
`--impl reference` times the UNMODIFIED reference's CPU EnvSet (oracle/_ref, all host threads) on the same config.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
PKG = "reinforcement-learning_b200"

ARENAS_PER_GPU = 4096
# Both arms are pre-rolled this many env steps from the kickoff reset before warm-up, so that the timed steps see the
# steady-state mix of driving / aerial / wall / car-car contact states (episodes end on a goal or after the 10 s no-touch
# timeout = 150 env steps): the cost of a step grows ~5x from the kickoff state to steady state on both arms.
PREROLL_STEPS = 160
PLAYERS_PER_TEAM = 2
OBS_SIZE_ADV_2V2 = 167
# algorithmic bytes per agent-step of the fused env step, 2v2 AdvancedObs:
# obs 668 + mask 90 + reward 4 + state share ~260 + action 4
ENV_BYTES_PER_AGENT_STEP = 1026
GAE_BYTES_PER_ELEM = 21
# DRAM traffic per launch of the dominant kernels from one `ncu --set full` capture each (dram__bytes_read.sum +
# dram__bytes_write.sum; profiles/r2ba_env_step_ncu_summary.txt, profiles/r2ba_gae_scan_ncu_summary.txt). Env step at 4096 arenas
# 2v2: state blobs in, mesh and instruction fetch against 16.8 MB algorithmic; GAE scan: 151.3 + 144.6 MB at L = 2^24 against
# 352 MB algorithmic (part of the 201 MB of outputs is still in the 126 MB L2 when the kernel ends).
ENV_TRAFFIC_BYTES_NCU = 19.7e6      # profiles/r2ba_env_step_ncu_summary.txt: 19.33 MB read + 0.42 MB written per launch (the 12.5 MB of outputs
                                    # and the state written back are still in L2 when the kernel ends); round 1: 36.2 MB
GAE_TRAFFIC_BYTES_NCU = 296.0e6
# ExperienceBuffer gather (staged kernel, profiles/r1j_expbuf_ncu_summary.txt): 1.202 GB read + 0.773 GB written at 2^20 samples against
# 1.63 GB algorithmic: a random row drags in whole 64-byte DRAM bursts around its unaligned ends, and each of the four scalars
# costs a burst of its own.
EXPBUF_TRAFFIC_BYTES_NCU = 1.975e9


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def dist_setup(n_gpus):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def barrier_sync(world):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x, world):
    import torch
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(x, world):
    import torch
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


# ------------------------------------------------------------------------------------------------ env workload
def bench_env(args, rank, world, local):
    import numpy as np
    import torch
    ggl = importlib.import_module(PKG)
    dev = torch.device("cuda", local)
    par = importlib.import_module(PKG + ".parallel")
    # weak scaling: every rank owns ARENAS_PER_GPU arenas of the world*ARENAS_PER_GPU arena set; no data-path collective
    _, my_arenas = par.shard_arenas(ARENAS_PER_GPU * world, world, rank)
    env = ggl.EnvSet(my_arenas, PLAYERS_PER_TEAM, obs_kind=ggl.abi.OBS_ADVANCED, seed=par.rank_seed(123, rank) & 0x7FFFFFFFFFFFFFFF,
                     device=local)
    assert env.obs_size == OBS_SIZE_ADV_2V2
    acts = torch.zeros(env.num_players, dtype=torch.int32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)

    def one(i):
        env.random_actions(acts, seed=1000 + i)
        env.step(acts)

    for i in range(PREROLL_STEPS):
        one(100000 + i)
    for i in range(args.warmup):
        one(i)
    # ---- device-timed region: K steps, CUDA events on the launch stream, L2 flushed between steps
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ggl.launch_count()
    barrier_sync(world)
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        flush.fill_(i & 0xff)
        evs[i][0].record(stream)
        env.random_actions(acts, seed=5000 + i)
        kev[i][0].record(stream)
        env.step(acts)
        kev[i][1].record(stream)
        evs[i][1].record(stream)
    barrier_sync(world)
    t_wall = time.perf_counter() - t_wall0
    launches = ggl.launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    overflow = env.contact_overflow()          # contact rows exhausted anywhere since create (must be 0: same rows as the reference)
    dev_ms = sum(a.elapsed_time(b) for a, b in evs)
    kern_ms = sum(a.elapsed_time(b) for a, b in kev) / args.steps
    dev_ms = max_over_ranks(dev_ms, world)
    agent_steps = sum_over_ranks(float(env.num_players) * args.steps, world)
    value = agent_steps / (dev_ms * 1e-3)

    # ---- e2e: host-buffer C-ABI call (pinned), H2D + D2H inside the timed region
    h_act = torch.zeros(env.num_players, dtype=torch.int32).pin_memory()
    h_obs = torch.empty(env.num_players, env.obs_size, dtype=torch.float32).pin_memory()
    h_mask = torch.empty(env.num_players, env.num_actions, dtype=torch.uint8).pin_memory()
    h_rew = torch.empty(env.num_players, dtype=torch.float32).pin_memory()
    h_term = torch.empty(env.num_arenas, dtype=torch.uint8).pin_memory()
    e2e_i = [0]

    def one_host():
        # the policy stand-in: a mask-valid random action per agent from the HOST masks the previous step returned
        # Learner loop (Learner.cpp:669-793): Reset + StepFirstHalf(async), policy on the host copies of obs / masks while
        # the first half runs, then StepSecondHalf with the host action indices and the rewards / terminals read back.
        e2e_i[0] += 1
        env.begin_step_host(h_obs, h_mask)
        env.host_random_actions(h_mask, h_act, 7000 + 1000003 * rank + e2e_i[0])
        env.finish_step_host(h_act, h_rew, h_term)

    env.step_host(h_act, h_obs, h_mask, h_rew, h_term)
    e2e_steps = max(3, min(args.steps, 200))
    for _ in range(3):
        one_host()
    barrier_sync(world)
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        one_host()
    barrier_sync(world)
    e2e_s = max_over_ranks(time.perf_counter() - t0, world)
    e2e_val = sum_over_ranks(float(env.num_players) * e2e_steps, world) / e2e_s
    h2d = env.num_players * 4
    d2h = env.num_players * (env.obs_size * 4 + env.num_actions + 4) + env.num_arenas

    peak, peak_src = peaks()
    achieved = ENV_BYTES_PER_AGENT_STEP * env.num_players / (kern_ms * 1e-3) / 1e9
    line = {
        "metric": "agent-steps/sec", "value": value, "unit": "agent-steps/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "soccar 2v2, %d arenas per GPU, AdvancedObs(167) + DefaultAction(90) + "
                               "{VelocityPlayerToBall,TouchBall,Goal} + {GoalScore,NoTouch(10)}, tick_skip 8, "
                               "device random mask-valid actions (BASELINE.json configs[1])" % ARENAS_PER_GPU,
                   "arenas_per_gpu": ARENAS_PER_GPU, "agents_per_gpu": env.num_players, "l2": "flushed between steps (256 MiB write)",
                   "preroll_steps": PREROLL_STEPS,
                   "arena_mesh": "synthetic soccar (real .cmf meshes are not redistributable)",
                   "block_balancing": os.environ.get("GGL_BALANCE", "1") != "0",
                   "contact_points_dropped": int(overflow)},
        "e2e": {"value": e2e_val, "unit": "agent-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps, "api": "ggl_envset_begin_step_host + host action pick + ggl_envset_finish_step_host (pinned host buffers)"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ENV_TRAFFIC_BYTES_NCU, "kernel": "env_step_kernel", "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": ENV_BYTES_PER_AGENT_STEP * env.num_players,
                     "note": "physics ticks are instruction-fetch / barrier-latency bound, not HBM bound (ncu r2v: 12% issue utilisation, "
                             "4% of the FMA pipe, 50% of the stall samples at the per-phase block barriers, 14% instruction fetch; "
                             "DESIGN.md section 4); the HBM floor is the obs/mask/reward write (SURVEY 8d)"},
        "wall_s": t_wall,
    }
    if rank == 0:
        line["clocks"] = clocks
        if world == 1:
            line["cpu_baseline"] = cpu_baseline_env()
            line["gae"] = gae_numbers(ggl, dev, steps=10, warmup=3)
    env.close()
    return line


def cpu_baseline_env(arenas=None, steps=30, warmup=5, threads=0):
    """The UNMODIFIED reference EnvSet (oracle/_ref/libref_oracle.so) on the host cores; falls back to saying so
    when the compiled reference did not travel to this box."""
    try:
        import oracle_ref
        if not oracle_ref.available():
            return {"value": None, "unit": "agent-steps/s", "cores": 0, "kind": "reference",
                    "sample": "oracle/_ref/libref_oracle.so not present on this box"}
        L = oracle_ref.lib()
        L.ref_set_threads(threads)
        cores = L.ref_num_threads()
        arenas = arenas or ARENAS_PER_GPU
        env = oracle_ref.RefEnvSet(arenas, PLAYERS_PER_TEAM, obs_kind=oracle_ref.abi.OBS_ADVANCED)
        sec = env.bench(steps, warmup + PREROLL_STEPS, 123)
        v = env.num_players * steps / sec
        env.close()
        return {"value": v, "unit": "agent-steps/s", "cores": cores, "kind": "reference",
                "sample": "%d arenas 2v2 AdvancedObs, %d pre-roll + %d warm-up + %d timed env steps, reference thread pool"
                          % (arenas, PREROLL_STEPS, warmup, steps)}
    except Exception as e:  # noqa
        return {"value": None, "unit": "agent-steps/s", "cores": 0, "kind": "reference", "sample": "failed: %r" % (e,)}


# ------------------------------------------------------------------------------------------------ GAE workload
def gae_inputs(torch, dev, T=256, N=65536, seed=7):
    g = torch.Generator(device=dev); g.manual_seed(seed)
    L = T * N
    r = torch.randn(L, device=dev, generator=g)
    v = torch.randn(L, device=dev, generator=g)
    t = torch.zeros(L, dtype=torch.int8, device=dev)
    kinds = torch.randint(1, 3, (N,), device=dev, generator=g).to(torch.int8)
    t[T - 1::T] = kinds
    K = int((t == 2).sum().item())
    tv = torch.randn(max(K, 1), device=dev, generator=g)[:K]
    return r, t, v, tv


def gae_numbers(ggl, dev, steps, warmup, T=256, N=65536):
    import torch
    r, t, v, tv = gae_inputs(torch, dev, T, N)
    L = r.numel()
    out = tuple(torch.empty(L, device=dev) for _ in range(3))
    ws = None
    for _ in range(warmup):
        *_, ws = ggl.gae_compute_device(r, t, v, tv, 0.99, 0.95, 1.7, 10.0, out=out, workspace=ws)
    s = torch.cuda.current_stream(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(s)
    for _ in range(steps):
        ggl.gae_compute_device(r, t, v, tv, 0.99, 0.95, 1.7, 10.0, out=out, workspace=ws)
    e1.record(s)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    peak, src = peaks()
    gbs = GAE_BYTES_PER_ELEM * L / (ms * 1e-3) / 1e9
    return {"L": L, "T": T, "N": N, "ms": ms, "elements_per_s": L / (ms * 1e-3), "GB/s": gbs, "peak": peak,
            "frac": gbs / peak, "bytes_per_element": GAE_BYTES_PER_ELEM, "peak_source": src,
            "inputs": "%d MB stream > L2 (126 MB), no flush needed" % (GAE_BYTES_PER_ELEM * L >> 20)}


def bench_gae(args, rank, world, local):
    import numpy as np
    import torch
    ggl = importlib.import_module(PKG)
    dev = torch.device("cuda", local)
    T, N = 256, 65536
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ggl.launch_count()
    barrier_sync(world)
    g = gae_numbers(ggl, dev, args.steps, args.warmup, T, N)
    barrier_sync(world)
    launches = ggl.launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(g["ms"], world)
    L = g["L"]
    value = GAE_BYTES_PER_ELEM * L * world / (ms * 1e-3) / 1e9
    # e2e through host buffers (the reference call takes CPU tensors: Learner.cpp:949-953)
    r, t, v, tv = [x.cpu().numpy() for x in gae_inputs(torch, dev, T, N)]
    hr, ht, hv = [torch.from_numpy(x).pin_memory().numpy() for x in (r, t, v)]
    outs = tuple(torch.empty(L).pin_memory().numpy() for _ in range(3))
    ggl.gae_compute_host(hr, ht, hv, tv, 0.99, 0.95, 1.7, 10.0, out=outs)
    t0 = time.perf_counter()
    n_e2e = 3
    for _ in range(n_e2e):
        ggl.gae_compute_host(hr, ht, hv, tv, 0.99, 0.95, 1.7, 10.0, out=outs)
    e2e_s = max_over_ranks((time.perf_counter() - t0) / n_e2e, world)
    line = {
        "metric": "GAE GB/s", "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": "GAE/returns over a flat stream of N=65536 segments x T=256 (L=2^24), mixed NORMAL/TRUNCATED "
                               "terminals, gamma .99 lambda .95 returnStd 1.7 clip 10 (BASELINE.json configs[3])",
                   "l2": "inputs+outputs 352 MB > L2"},
        "e2e": {"value": GAE_BYTES_PER_ELEM * L * world / e2e_s / 1e9, "unit": "GB/s", "h2d_bytes_per_step": 9 * L + 4 * len(tv),
                "d2h_bytes_per_step": 12 * L + 8, "api": "ggl_gae_compute_host"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": g["GB/s"], "peak": g["peak"], "unit": "GB/s", "frac": g["frac"],
                     "traffic": GAE_TRAFFIC_BYTES_NCU, "kernel": "gae_scan_kernel", "peak_source": g["peak_source"],
                     "algorithmic_bytes_per_launch": GAE_BYTES_PER_ELEM * L},
    }
    if rank == 0:
        line["clocks"] = clocks
        if world == 1:
            import oracle_c
            Ls = 1 << 22
            t0 = time.perf_counter()
            oracle_c.gae(r[:Ls], t[:Ls], v[:Ls], None, 0.99, 0.95, 1.7, 10.0)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": GAE_BYTES_PER_ELEM * Ls / dt / 1e9, "unit": "GB/s", "cores": 1, "kind": "port",
                                    "sample": "first 2^22 elements of the same stream, C restatement of GAE.cpp, 1 thread"}
    return line


# ------------------------------------------------------------------------------------------------ ExperienceBuffer workload
def bench_expbuf(args, rank, world, local):
    """One epoch's worth of minibatches = one fused permuted gather of the experience stream:
    2^20 samples, obs 167, 90 actions; bytes = 2 x (4*obs + 90 + 16) + 8 per sample."""
    import numpy as np
    import torch
    ggl = importlib.import_module(PKG)
    dev = torch.device("cuda", local)
    n, obs, na = 1 << 20, OBS_SIZE_ADV_2V2, 90
    g = torch.Generator(device=dev); g.manual_seed(3 + rank)
    data = {"states": torch.randn(n, obs, device=dev, generator=g), "actions": torch.randint(0, na, (n,), device=dev, generator=g, dtype=torch.int32),
            "log_probs": torch.randn(n, device=dev, generator=g), "target_values": torch.randn(n, device=dev, generator=g),
            "action_masks": (torch.rand(n, na, device=dev, generator=g) < 0.5).to(torch.uint8), "advantages": torch.randn(n, device=dev, generator=g)}
    perms = [torch.randperm(n, device=dev, generator=g) for _ in range(4)]
    if os.environ.get("GGL_EXPBUF_IDENTITY"):          # experiment knob: sequential rows = the streaming-copy ceiling of this kernel
        perms = [torch.arange(n, device=dev) for _ in range(4)]
    out = ggl.expbuf_gather(data, perms[0])
    bytes_per_sample = 2 * (4 * obs + na + 16) + 8
    s = torch.cuda.current_stream(dev)
    for i in range(args.warmup):
        ggl.expbuf_gather(data, perms[i % 4], out=out)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ggl.launch_count()
    barrier_sync(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s)
    for i in range(args.steps):
        ggl.expbuf_gather(data, perms[i % 4], out=out)
    e1.record(s)
    barrier_sync(world)
    launches = ggl.launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(e0.elapsed_time(e1) / args.steps, world)
    gbs = bytes_per_sample * n / (ms * 1e-3) / 1e9
    peak, src = peaks()
    # e2e: the permutation comes from the host every epoch (std::shuffle stays on the host) and the result is read back
    h_perm = torch.randperm(n).pin_memory()
    h_adv = torch.empty(n).pin_memory()
    d_perm = torch.empty(n, dtype=torch.int64, device=dev)
    t0 = time.perf_counter()
    n_e2e = 5
    for _ in range(n_e2e):
        d_perm.copy_(h_perm, non_blocking=True)
        ggl.expbuf_gather(data, d_perm, out=out)
        h_adv.copy_(out["advantages"], non_blocking=True)
        torch.cuda.synchronize()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / n_e2e, world)
    line = {"metric": "ExperienceBuffer gather GB/s", "value": gbs * world, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "ExperienceBuffer epoch: fused permuted gather of 2^20 samples x (obs 167 f32, 90 mask bytes, 4 scalars)",
                       "l2": "sources + destinations 1.6 GB > L2"},
            "e2e": {"value": bytes_per_sample * n * world / e2e_s / 1e9, "unit": "GB/s", "h2d_bytes_per_step": 8 * n, "d2h_bytes_per_step": 4 * n,
                    "api": "H2D permutation + ggl_expbuf_gather + D2H advantages"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak, "traffic": EXPBUF_TRAFFIC_BYTES_NCU if n == (1 << 20) else None,
                         "kernel": "expbuf_gather_staged_kernel", "peak_source": src, "algorithmic_bytes_per_launch": bytes_per_sample * n}}
    if rank == 0:
        line["clocks"] = clocks
        if world == 1:
            import expbuf_oracle as eo
            ns = 1 << 17
            hd = {k: v[:ns].cpu().numpy() for k, v in data.items()}
            p = np.random.default_rng(0).permutation(ns)
            t0 = time.perf_counter()
            eo.get_samples(hd, p, -1)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": bytes_per_sample * ns / dt / 1e9, "unit": "GB/s", "cores": 1, "kind": "port",
                                    "sample": "2^17 samples of the same shape, numpy restatement of _GetSamples (six fancy-index gathers)"}
    return line



# ------------------------------------------------------------------------------------------------ trajectory-assembly workload
def bench_traj(args, rank, world, local):
    """Trajectory assembly at config C2's shape: 16 384 players (4 096 arenas 2v2), obs 167, 90 actions,
    maxEpisodeLength 1800 (120 s at tick_skip 8), arena terminals with probability 1/150 per step (mean episode = the no-touch
    timeout). A step = push_obs + push_result; bytes = every appended row in and out, plus every moved row in and out."""
    import numpy as np
    import torch
    ggl = importlib.import_module(PKG)
    dev = torch.device("cuda", local)
    arenas, ppa, obs, na, max_len = ARENAS_PER_GPU, 2 * PLAYERS_PER_TEAM, OBS_SIZE_ADV_2V2, 90, 1800
    n = arenas * ppa
    ts_per_itr = 1 << 20
    store = ggl.TrajectoryStore(n, ppa, obs, na, max_len, capacity_timesteps=ts_per_itr + (3 << 20), device=local)
    g = torch.Generator(device=dev); g.manual_seed(11 + rank)
    obs_d = torch.randn(n, obs, device=dev, generator=g); obs_f = torch.randn(n, obs, device=dev, generator=g)
    masks = (torch.rand(n, na, device=dev, generator=g) < 0.5).to(torch.uint8)
    acts = torch.randint(0, na, (n,), device=dev, generator=g, dtype=torch.int32)
    logp = torch.randn(n, device=dev, generator=g); rew = torch.randn(n, device=dev, generator=g)
    nterm = 64
    u = torch.rand(nterm, arenas, device=dev, generator=g)
    terms = torch.where(u < 0.5 / 150, 1, torch.where(u < 1.0 / 150, 2, 0)).to(torch.uint8)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    s = torch.cuda.current_stream(dev)
    row_bytes = 4 * obs + na + 13

    def one(i):
        store.push_obs(obs_d, masks)
        store.push_result(acts, logp, rew, terms[i % nterm], obs_f)

    store.begin_iteration()
    for i in range(300 + args.warmup):                  # pre-roll: episode lengths reach their steady-state mix
        one(i)
        if store.counts()[0] >= ts_per_itr:
            store.begin_iteration()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ggl.launch_count()
    barrier_sync(world)
    # the four launches of a step replayed as one CUDA graph per terminal pattern (a launch-bound inner loop: 32 MB per step)
    graphs = []
    side = torch.cuda.Stream(dev)
    side.wait_stream(s)
    with torch.cuda.stream(side):
        for j in range(nterm):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=side):
                one(j)
            graphs.append(gr)
    s.wait_stream(side)
    torch.cuda.synchronize()
    total_ms, moved, trunc = 0.0, 0, 0
    for i in range(args.steps):
        L0, K0 = store.counts()
        if L0 >= ts_per_itr:
            store.begin_iteration(); L0, K0 = 0, 0
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s); graphs[i % nterm].replay(); e1.record(s)
        L1, K1 = store.counts()
        total_ms += e0.elapsed_time(e1); moved += L1 - L0; trunc += K1 - K0
    barrier_sync(world)
    launches = ggl.launch_count() - l0
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(total_ms / args.steps, world)
    bytes_step = 2 * row_bytes * n + (2 * row_bytes * moved + 8 * obs * trunc) / args.steps
    gbs = bytes_step / (ms * 1e-3) / 1e9
    peak, src = peaks()
    # e2e: the env outputs arrive in pinned host buffers (a host-side env), the stream length goes back to the host every step
    hb = [x.cpu().pin_memory() for x in (obs_d, masks, acts, logp, rew, terms[0], obs_f)]
    db = [torch.empty_like(x, device=dev) for x in hb]
    torch.cuda.synchronize()
    n_e2e = max(5, args.steps)
    t0 = time.perf_counter()
    for i in range(n_e2e):
        for d, h in zip(db, hb):
            d.copy_(h, non_blocking=True)
        store.push_obs(db[0], db[1]); store.push_result(db[2], db[3], db[4], db[5], db[6])
        if store.counts()[0] >= ts_per_itr:
            store.begin_iteration()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / n_e2e, world)
    h2d = sum(h.numel() * h.element_size() for h in hb)
    line = {"metric": "trajectory assembly GB/s", "value": gbs * world, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "trajectory assembly, %d players (2v2 x %d arenas), obs 167, 90 actions, maxEpisodeLength 1800, arena terminal "
                                   "probability 1/150 per step, tsPerItr 2^20" % (n, arenas),
                       "l2": "256 MB flush before every timed step", "launch": "CUDA graph replay of the step's three launches", "agent_steps_per_s": n * world / (ms * 1e-3)},
            "e2e": {"value": bytes_step * world / e2e_s / 1e9, "unit": "GB/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 32,
                    "api": "H2D env outputs + ggl_traj_push_obs + ggl_traj_push_result + ggl_traj_counts"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak, "traffic": None,
                         "kernel": "traj_push_obs_kernel + traj_push_scalars_kernel + traj_scan_kernel + traj_move_kernel (one graph replay per step)", "peak_source": src,
                         "algorithmic_bytes_per_launch": bytes_step}}
    if rank == 0:
        line["clocks"] = clocks
        if world == 1:
            import traj_oracle as to
            pn, steps_cpu = 1024, 60
            o = to.TrajOracle(pn, ppa, 25)
            ho, hm = hb[0][:pn].numpy(), hb[1][:pn].numpy()
            ha, hl, hr, hf = hb[2][:pn].numpy(), hb[3][:pn].numpy(), hb[4][:pn].numpy(), hb[6][:pn].numpy()
            ht = np.zeros(pn // ppa, np.uint8)
            o.begin_iteration()
            t0 = time.perf_counter()
            for i in range(steps_cpu):
                o.push_obs(ho, hm); o.push_result(ha, hl, hr, ht, hf)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": 4 * row_bytes * pn * steps_cpu / dt / 1e9, "unit": "GB/s", "cores": 1, "kind": "port",
                                    "sample": "1024 players x 60 steps, maxEpisodeLength 25, Python restatement of Learner.cpp:701-712/823-863"}
    store.close()
    return line


# ------------------------------------------------------------------------------------------------ obs-statistics workload
def bench_obsstat(args, rank, world, local):
    """Obs statistics + in-place normalisation at config C2's shape: 16 384 rows x 167, 100 sampled rows per step
    (Learner defaults). Bytes = 8 per obs element (read + write) + the sampled rows."""
    import numpy as np
    import torch
    ggl = importlib.import_module(PKG)
    dev = torch.device("cuda", local)
    rows, width, ns = ARENAS_PER_GPU * 2 * PLAYERS_PER_TEAM, OBS_SIZE_ADV_2V2, 100
    g = torch.Generator(device=dev); g.manual_seed(5 + rank)
    src = torch.randn(rows, width, device=dev, generator=g) * 3 + 1
    obs = src.clone()
    idx = torch.randint(0, rows, (ns,), device=dev, generator=g, dtype=torch.int32)
    stat = ggl.ObsStat(width, device=local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    s = torch.cuda.current_stream(dev)

    def one():
        stat.increment_rows(obs, idx)
        stat.normalize_in_place(obs, 3.0, 0.1)

    for _ in range(args.warmup):
        obs.copy_(src); one()
    gr = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream(dev); side.wait_stream(s)
    with torch.cuda.stream(side):
        with torch.cuda.graph(gr, stream=side):
            one()
    s.wait_stream(side); torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ggl.launch_count()
    barrier_sync(world)
    total_ms = 0.0
    for i in range(args.steps):
        obs.copy_(src); flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s); gr.replay(); e1.record(s)
        torch.cuda.synchronize()
        total_ms += e0.elapsed_time(e1)
    barrier_sync(world)
    launches = 2 * args.steps + (ggl.launch_count() - l0)
    clocks = sampler.stop() if rank == 0 else None
    ms = max_over_ranks(total_ms / args.steps, world)
    bytes_step = 8 * rows * width + 4 * ns * width
    gbs = bytes_step / (ms * 1e-3) / 1e9
    peak, psrc = peaks()
    h_obs = src.cpu().pin_memory(); h_out = torch.empty_like(h_obs).pin_memory(); h_idx = idx.cpu().pin_memory()
    torch.cuda.synchronize()
    n_e2e = max(5, args.steps)
    t0 = time.perf_counter()
    for i in range(n_e2e):
        obs.copy_(h_obs, non_blocking=True); idx.copy_(h_idx, non_blocking=True)
        one()
        h_out.copy_(obs, non_blocking=True)
        torch.cuda.synchronize()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / n_e2e, world)
    line = {"metric": "obs normalisation GB/s", "value": gbs * world, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64 statistics / f32 obs", "data": "synthetic",
            "config": {"workload": "BatchedWelfordStat: 100 sampled rows + NormalizeInPlace over %d x %d obs" % (rows, width),
                       "l2": "256 MB flush before every timed step", "launch": "CUDA graph replay of the step's two launches"},
            "e2e": {"value": bytes_step * world / e2e_s / 1e9, "unit": "GB/s", "h2d_bytes_per_step": 4 * rows * width + 4 * ns,
                    "d2h_bytes_per_step": 4 * rows * width, "api": "H2D obs + ggl_obsstat_increment_rows + ggl_obsstat_normalize + D2H obs"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak, "traffic": None,
                         "kernel": "obsstat_increment_kernel + obsstat_apply_kernel", "peak_source": psrc,
                         "algorithmic_bytes_per_launch": bytes_step}}
    if rank == 0:
        line["clocks"] = clocks
        if world == 1:
            import welford_oracle as wo
            st = wo.BatchedWelfordStat(width)
            ho, hi = src.cpu().numpy(), idx.cpu().numpy()
            t0 = time.perf_counter()
            reps = 10
            for _ in range(reps):
                o = ho.copy(); st.increment_rows(o, hi); st.normalize_in_place(o, 3.0, 0.1)
            dt = (time.perf_counter() - t0) / reps
            line["cpu_baseline"] = {"value": bytes_step / dt / 1e9, "unit": "GB/s", "cores": 1, "kind": "port",
                                    "sample": "10 steps of the same shape, numpy restatement of WelfordStat.h:89-216 (bit-exact against the compiled reference)"}
    stat.close()
    return line

# ------------------------------------------------------------------------------------------------ reference arm
def bench_reference(args):
    """Times the reference's own CPU EnvSet on this box's host cores (oracle/_ref). Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return None
    n = max(1, args.gpus)
    cb = cpu_baseline_env(arenas=ARENAS_PER_GPU, steps=max(1, args.steps), warmup=max(1, args.warmup))
    v = cb["value"]
    line = {"impl": "reference", "metric": "agent-steps/sec", "value": v, "unit": "agent-steps/s", "n_gpus": n,
            "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": (ARENAS_PER_GPU * 2 * PLAYERS_PER_TEAM / v * 1e3) if v else None,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "soccar 2v2, %d arenas, AdvancedObs(167) + DefaultAction(90), tick_skip 8, random mask-valid "
                                   "actions: the reference RLGC::EnvSet on the host thread pool (CPU only; the reference has no "
                                   "multi-GPU path, so the same CPU number stands at every N)" % ARENAS_PER_GPU,
                       "arena_mesh": "synthetic soccar", "preroll_steps": PREROLL_STEPS},
            "cpu_baseline": cb,
            "e2e": {"value": v, "unit": "agent-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if v is None:
        line = {"impl": "reference", "unavailable": cb["sample"]}
    return line


_REAL_STDOUT = None


def _claim_stdout():
    """stdout carries exactly ONE line, the JSON: anything a library writes to fd 1 meanwhile (NCCL prints its version there
    when NCCL_DEBUG is set on the box) goes to stderr instead."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def _emit(line):
    sys.stdout.flush()
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=["env", "gae", "expbuf", "traj", "obsstat"])
    args = ap.parse_args()
    if args.impl == "reference":
        args.steps = args.steps if args.steps is not None else 30
        args.warmup = args.warmup if args.warmup is not None else 5
        line = bench_reference(args)
        if line is not None:
            _emit(line)
        return
    ggl = importlib.import_module(PKG)
    ggl.lib()   # fail loudly if the CUDA library is missing
    workload = args.workload or ("env" if hasattr(ggl, "EnvSet") else "gae")
    if args.steps is None:
        args.steps = 200 if workload == "env" else 20
    if args.warmup is None:
        args.warmup = 20 if workload == "env" else 5
    args.warmup = max(args.warmup, 3)
    rank, world, local = dist_setup(args.gpus)
    line = {"env": bench_env, "gae": bench_gae, "expbuf": bench_expbuf, "traj": bench_traj, "obsstat": bench_obsstat}[workload](args, rank, world, local)
    if rank == 0:
        _emit(line)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
