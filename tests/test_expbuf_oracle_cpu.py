"""Pins the numpy restatement of ExperienceBuffer batching (oracle/expbuf_oracle.py) and the C-ABI partition rule against the
UNMODIFIED reference (oracle/_ref/libref_expbuf.so) when it is present, and against the committed golden vectors always."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import expbuf_oracle as eo      # noqa: E402
import expbuf_cases as ec       # noqa: E402
import oracle_ref               # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden", "expbuf_golden.npz")


def test_partition_rule_matches_c_abi(ggl):
    for n in list(range(0, 40)) + [255, 256, 257, 511, 512, 513, 1000, 65536, 100003]:
        for bs in (1, 7, 16, 256, 50000):
            for ob in (False, True):
                plan = eo.plan_batches(n, bs, ob)
                assert plan == ggl.expbuf_plan_batches(n, bs, ob), (n, bs, ob)
                assert all(s >= bs or (ob and st == 0) for st, s in plan)
                if plan and ob:
                    assert sum(s for _, s in plan) == n            # overbatching never drops a sample
    assert ggl.lib().ggl_expbuf_plan_batches(10, 0, 0, None, None, 0) == -1


def test_oracle_matches_golden_vectors():
    g = np.load(GOLDEN)
    for ci, case in enumerate(ec.CASES):
        rows, rows_a, obs, na, bs, ob, maxa = case
        data = ec.make_data(rows, rows_a, obs, na, seed=ci)
        for epoch in range(2):
            sizes = g["c%d_e%d_sizes" % (ci, epoch)]
            perm = g["c%d_e%d_perm" % (ci, epoch)]
            n = eo.exp_size(rows, rows_a)
            assert [s for _, s in eo.plan_batches(n, bs, ob)] == list(sizes)
            out = eo.get_samples(data, perm, maxa) if len(perm) else {}
            key = "c%d_e%d_" % (ci, epoch)
            if not out:
                assert len(g[key + "actions"]) == 0
                continue
            assert np.array_equal(out["actions"], g[key + "actions"]), (case, epoch)
            assert np.array_equal(out["advantages"], g[key + "advantages"]) and np.array_equal(out["target_values"], g[key + "target_values"])
            assert np.array_equal(out["states"].astype(np.float64).sum(1), g[key + "states_rowsum"])
            assert np.array_equal(out["action_masks"].sum(1).astype(np.int32), g[key + "masks_rowcount"])


@pytest.mark.skipif(not oracle_ref.expbuf_available(), reason="compiled reference not present")
def test_oracle_matches_compiled_reference():
    for ci, case in enumerate(ec.CASES):
        rows, rows_a, obs, na, bs, ob, maxa = case
        data = ec.make_data(rows, rows_a, obs, na, seed=100 + ci)
        ref = oracle_ref.RefExpBuf(seed=7 + ci, max_action_index=maxa)
        ref.set_data(data)
        prev = None
        for epoch in range(3):
            sizes, out = ref.get_batches(bs, ob)
            n = eo.exp_size(rows, rows_a)
            assert [s for _, s in eo.plan_batches(n, bs, ob)] == list(sizes), (case, epoch)
            perm = out["log_probs"].astype(np.int64)               # the id channel reveals the reference's permutation
            assert len(set(perm.tolist())) == len(perm) and (perm < n).all()
            if prev is not None and len(perm) == len(prev) and len(perm) > 8:
                assert not np.array_equal(perm, prev)              # re-shuffled every epoch
            prev = perm
            mine = eo.get_samples(data, perm, maxa) if len(perm) else {f: out[f][:0] for f in ec.FIELDS}
            for f in ec.FIELDS:
                assert np.array_equal(mine[f], out[f]), (case, epoch, f)
        ref.close()
