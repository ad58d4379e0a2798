"""Workload census of steady-state play (debug): histograms over arenas of solver rows, island sizes, mesh candidates, ...
    python tools/tick_stats.py [preroll] [ticks] [ppt] [arenas]"""
import ctypes as C
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

NAMES = ["solver rows", "rows of largest island", "(car,tri) candidates", "islands with rows", "wheels in contact",
         "cars with world/car rows", "mesh hits", "largest per-car window"]


def main():
    preroll = int(sys.argv[1]) if len(sys.argv) > 1 else 160
    ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    ppt = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    arenas = int(sys.argv[4]) if len(sys.argv) > 4 else 4096
    ggl = importlib.import_module("reinforcement-learning_b200")
    L = ggl.lib()
    L.ggl_debug_tick_stats.argtypes = [C.c_void_p, C.c_void_p]
    env = ggl.EnvSet(arenas, ppt, obs_kind=ggl.abi.OBS_ADVANCED, seed=123)
    acts = torch.zeros(env.num_players, dtype=torch.int32, device="cuda")
    for i in range(preroll):
        env.random_actions(acts, seed=1000 + i)
        env.step(acts)
    tot = np.zeros((8, 40), np.uint64)
    buf = np.zeros((8, 40), np.uint32)
    for t in range(ticks):
        env.arena_step(1)
        assert L.ggl_debug_tick_stats(env.h, buf.ctypes.data) == 0
        tot += buf
    print("census over %d arenas x %d ticks after %d env steps (%dv%d)" % (arenas, ticks, preroll, ppt, ppt))
    for h, n in enumerate(NAMES):
        row = tot[h].astype(np.float64); s = row.sum() or 1.0
        mean = (row * np.arange(40)).sum() / s
        nz = [(i, 100 * v / s) for i, v in enumerate(row) if v]
        print("%-26s mean %5.2f  " % (n, mean) + " ".join("%d:%.1f%%" % (i, p) if p >= 0.05 else "%d:<" % i for i, p in nz))


if __name__ == "__main__":
    main()
