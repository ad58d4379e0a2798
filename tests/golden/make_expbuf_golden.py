"""Generates tests/golden/expbuf_golden.npz from the UNMODIFIED reference ExperienceBuffer (oracle/_ref/libref_expbuf.so, built by
`make -C oracle`). Run in the build container: python tests/golden/make_expbuf_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import expbuf_cases as ec   # noqa: E402
import oracle_ref           # noqa: E402

out = {}
for ci, case in enumerate(ec.CASES):
    rows, rows_a, obs, na, bs, ob, maxa = case
    data = ec.make_data(rows, rows_a, obs, na, seed=ci)
    ref = oracle_ref.RefExpBuf(seed=1000 + ci, max_action_index=maxa)
    ref.set_data(data)
    for epoch in range(2):
        sizes, o = ref.get_batches(bs, ob)
        out["c%d_e%d_sizes" % (ci, epoch)] = sizes
        out["c%d_e%d_perm" % (ci, epoch)] = o["log_probs"].astype(np.int64)
        # the gathered tensors themselves, reduced where they are bulky (row sums of states in f64, row popcounts of masks)
        out["c%d_e%d_actions" % (ci, epoch)] = o["actions"]
        out["c%d_e%d_advantages" % (ci, epoch)] = o["advantages"]
        out["c%d_e%d_target_values" % (ci, epoch)] = o["target_values"]
        out["c%d_e%d_states_rowsum" % (ci, epoch)] = o["states"].astype(np.float64).sum(1)
        out["c%d_e%d_masks_rowcount" % (ci, epoch)] = o["action_masks"].sum(1).astype(np.int32)
    ref.close()
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "expbuf_golden.npz"), **out)
print("wrote", len(out), "arrays")
