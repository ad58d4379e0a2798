"""Generates tests/golden/obsstat_golden.npz from the UNMODIFIED reference BatchedWelfordStat (oracle/_ref/libref_welford.so, built
by `make -C oracle`). Run in the build container: python tests/golden/make_obsstat_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import obsstat_cases as oc  # noqa: E402
import oracle_ref           # noqa: E402

out = {}
for ci, case in enumerate(oc.CASES):
    rows, width, ns, nsteps, mc, ms = case
    ref = oracle_ref.RefWelford(width)
    for si, (obs, idx) in enumerate(oc.steps(case, seed=ci)):
        ref.increment_rows(obs, idx)
        ref.normalize_in_place(obs, mc, ms)
        m, v, c = ref.state()
        out["c%d_s%d_means" % (ci, si)] = m
        out["c%d_s%d_vars" % (ci, si)] = v
        out["c%d_s%d_count" % (ci, si)] = np.int64(c)
        out["c%d_s%d_obs_head" % (ci, si)] = obs[:4].copy()                     # normalised rows, bit for bit
        out["c%d_s%d_obs_sum" % (ci, si)] = obs.astype(np.float64).sum(0)
    ref.close()
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "obsstat_golden.npz"), **out)
print("wrote", len(out), "arrays")
