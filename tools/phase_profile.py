"""Per-phase cycle profile of the env step (profiling build: python reinforcement-learning_b200/build.py --prof).

    GGL_LIB_PATH=reinforcement-learning_b200/libgigaenv_cuda_prof.so python tools/phase_profile.py [preroll] [steps] [ppt] [arenas]
"""
import ctypes as C
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

NAMES = ["pre(snap,pads)", "car_pre_a", "wheel rays", "car_pre_c", "forces", "collide_body", "gather_contacts", "solve+free",
         "integrate", "car_post", "pad_check", "finish", "env layer", "stage_in", "-", "-"]


def main():
    preroll = int(sys.argv[1]) if len(sys.argv) > 1 else 160
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    ppt = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    arenas = int(sys.argv[4]) if len(sys.argv) > 4 else 4096
    ggl = importlib.import_module("reinforcement-learning_b200")
    L = ggl.lib()
    env = ggl.EnvSet(arenas, ppt, obs_kind=ggl.abi.OBS_ADVANCED, seed=123)
    acts = torch.zeros(env.num_players, dtype=torch.int32, device="cuda")
    buf = (C.c_uint64 * 16)()
    for i in range(preroll):
        env.random_actions(acts, seed=1000 + i)
        env.step(acts)
    L.ggl_debug_phase_cycles(buf, 1)
    import numpy as np
    na = min(arenas, 8192)
    per = np.zeros((na, 16), np.uint64)
    L.ggl_debug_arena_phase_cycles(C.c_void_p(per.ctypes.data), na, 1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        env.random_actions(acts, seed=5000 + i)
        env.step(acts)
    e1.record()
    torch.cuda.synchronize()
    have = L.ggl_debug_phase_cycles(buf, 1)
    print("ms/step %.3f (preroll %d, %d steps, %dv%d x %d arenas)%s" % (e0.elapsed_time(e1) / steps, preroll, steps, ppt, ppt, arenas,
                                                                     "" if have else "  [product build: no phase counters]"))
    tot = float(sum(buf)) or 1.0
    for n, v in zip(NAMES, buf):
        if v:
            print("%-18s %6.2f %%   %.0f cycles/arena/step" % (n, 100.0 * v / tot, v / arenas / steps))


    sub = (C.c_uint64 * 16)()
    if L.ggl_debug_phase_cycles(sub, 2) == 1:
        print("collide_body sub-phases (cycles summed over lanes, per arena-step): " +
              ", ".join("%s %.0f" % (n, v / arenas / steps) for n, v in zip(["car-ball", "car-mesh", "car-planes", "car-car", "ball-mesh"], sub)))
    if L.ggl_debug_arena_phase_cycles(C.c_void_p(per.ctypes.data), na, 1) == 1:
        per = per.astype(np.float64) / steps
        tot_a = per.sum(1)
        q = np.percentile(tot_a, [0, 10, 50, 90, 99, 100])
        print("per-arena cycles/step percentiles  min %.0f  p10 %.0f  p50 %.0f  p90 %.0f  p99 %.0f  max %.0f  mean %.0f" % (*q, tot_a.mean()))
        # warps = 4 consecutive arenas: the kernel ends with the slowest warp
        if na % 4 == 0:
            w = tot_a.reshape(-1, 4).max(1)
            print("per-warp (max of 4 arenas) p50 %.0f p99 %.0f max %.0f" % tuple(np.percentile(w, [50, 99, 100])))
        gpb = int(os.environ.get("GGL_GROUPS_PER_BLOCK", "28"))          # lane groups (arena positions) per block: 28 for 2v2
        nblk = (na + gpb - 1) // gpb
        blk = np.array([tot_a[b * gpb:(b + 1) * gpb].max() for b in range(nblk)])
        print("per-BLOCK cycles/step (positions of a block end together): min %.0f p10 %.0f p50 %.0f p90 %.0f max %.0f mean %.0f  -> max/mean %.2f" %
              (*np.percentile(blk, [0, 10, 50, 90, 100]), blk.mean(), blk.max() / blk.mean()))
        top = np.argsort(-tot_a)[: max(1, na // 100)]
        print("phase share in the slowest 1% of arenas vs all:")
        for i, n in enumerate(NAMES):
            if per[:, i].sum() > 0:
                print("  %-18s top1%% %9.0f   all %9.0f" % (n, per[top, i].mean(), per[:, i].mean()))


if __name__ == "__main__":
    main()
