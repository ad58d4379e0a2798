"""GPU parity of the fused ExperienceBuffer gather (ggl_expbuf_gather) against the numpy oracle, the committed reference
vectors and, when it travelled to this box, the compiled reference itself. Bit-exact: this is index / byte / copy work."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import expbuf_oracle as eo      # noqa: E402
import expbuf_cases as ec       # noqa: E402
import oracle_ref               # noqa: E402


@pytest.fixture(params=["staged", "registers"], autouse=True)
def kernel_choice(request, monkeypatch):
    """Every test runs twice: through the staged kernel (aligned tensors, count >= 32) and through the general register kernel
    the library falls back to for unaligned tensors, very wide rows and tiny gathers (GGL_EXPBUF_REGISTER_KERNEL test hook)."""
    if request.param == "registers":
        monkeypatch.setenv("GGL_EXPBUF_REGISTER_KERNEL", "1")
    else:
        monkeypatch.delenv("GGL_EXPBUF_REGISTER_KERNEL", raising=False)
    return request.param


def _to_dev(data):
    import torch
    return {k: (None if v is None else torch.from_numpy(np.ascontiguousarray(v)).cuda()) for k, v in data.items()}


def _check(ggl, data, perm, maxa):
    import torch
    dev = _to_dev(data)
    idx = torch.from_numpy(np.asarray(perm, np.int64)).cuda()
    got = ggl.expbuf_gather(dev, idx, max_action_index=maxa)
    torch.cuda.synchronize()
    ref = eo.get_samples(data, perm, maxa)
    for f in ec.FIELDS:
        if data.get(f) is None:
            assert got[f] is None
        else:
            assert np.array_equal(got[f].cpu().numpy(), ref[f]), f


@pytest.mark.parametrize("ci", range(len(ec.CASES)))
def test_gather_matches_oracle_on_reference_permutations(ggl, ci):
    rows, rows_a, obs, na, bs, ob, maxa = ec.CASES[ci]
    g = np.load(os.path.join(ROOT, "tests", "golden", "expbuf_golden.npz"))
    data = ec.make_data(rows, rows_a, obs, na, seed=ci)
    for epoch in range(2):
        perm = g["c%d_e%d_perm" % (ci, epoch)]
        if len(perm):
            _check(ggl, data, perm, maxa)
            import torch
            got = ggl.expbuf_gather(_to_dev(data), torch.from_numpy(perm).cuda(), max_action_index=maxa)
            assert np.array_equal(got["actions"].cpu().numpy(), g["c%d_e%d_actions" % (ci, epoch)])
            assert np.array_equal(got["states"].cpu().numpy().astype(np.float64).sum(1), g["c%d_e%d_states_rowsum" % (ci, epoch)])


def test_edge_cases(ggl):
    rng = np.random.default_rng(5)
    data = ec.make_data(300, 300, 167, 90, seed=9)
    _check(ggl, data, rng.integers(0, 300, 1000), -1)                  # repeated rows
    _check(ggl, data, np.array([299, 300, 10**9, 0, 5]), -1)           # out of range: clamped to the last row
    _check(ggl, data, np.array([3]), 0)                                # single sample, every action clamped to 0
    d2 = dict(data); d2["advantages"] = None; d2["action_masks"] = None  # undefined tensors are skipped
    _check(ggl, d2, rng.permutation(300), 89)
    d3 = dict(data); d3["actions"] = data["actions"][:120]             # the shorter actions tensor sets the row limit
    _check(ggl, d3, rng.permutation(300), -1)
    import torch
    empty = ggl.expbuf_gather(_to_dev(data), torch.zeros(0, dtype=torch.int64, device="cuda"))
    assert empty["states"].shape == (0, 167)
    # tensors that do not start on a 16-byte boundary take the general kernel: same results
    dev = _to_dev(data)
    big = torch.empty(300 * 167 + 1, device="cuda"); big[1:].copy_(dev["states"].flatten())
    dev["states"] = big[1:].view(300, 167)
    perm = rng.permutation(300)
    got = ggl.expbuf_gather(dev, torch.from_numpy(perm).cuda())
    assert np.array_equal(got["states"].cpu().numpy(), data["states"][perm])


@pytest.mark.skipif(not oracle_ref.expbuf_available(), reason="compiled reference not present on this box")
def test_epoch_of_minibatches_matches_compiled_reference(ggl):
    import torch
    rows, obs, na, bs = 5000, 167, 90, 1024
    data = ec.make_data(rows, rows, obs, na, seed=77)
    ref = oracle_ref.RefExpBuf(seed=3, max_action_index=na - 1)
    ref.set_data(data)
    dev = _to_dev(data)
    for epoch in range(2):
        sizes, out = ref.get_batches(bs, True)
        plan = ggl.expbuf_plan_batches(rows, bs, True)
        assert [s for _, s in plan] == list(sizes)
        perm = torch.from_numpy(out["log_probs"].astype(np.int64)).cuda()
        got = ggl.expbuf_gather(dev, perm, max_action_index=na - 1)
        for (start, size) in plan:                                     # every minibatch is a view of the gathered stream
            for f in ec.FIELDS:
                assert np.array_equal(got[f][start:start + size].cpu().numpy(), out[f][start:start + size]), (epoch, start, f)
    ref.close()


def test_full_size_round_trip(ggl):
    """BASELINE-sized stream (2^20 samples x obs 167): gathering with a permutation and then with its inverse is the identity."""
    import torch
    n, obs, na = 1 << 20, 167, 90
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    data = {"states": torch.randn(n, obs, device="cuda", generator=g), "actions": torch.randint(0, na, (n,), device="cuda", generator=g, dtype=torch.int32),
            "log_probs": torch.randn(n, device="cuda", generator=g), "target_values": torch.randn(n, device="cuda", generator=g),
            "action_masks": (torch.rand(n, na, device="cuda", generator=g) < 0.5).to(torch.uint8), "advantages": torch.randn(n, device="cuda", generator=g)}
    perm = torch.randperm(n, device="cuda", generator=g)
    inv = torch.empty_like(perm); inv[perm] = torch.arange(n, device="cuda")
    a = ggl.expbuf_gather(data, perm)
    b = ggl.expbuf_gather(a, inv)
    for f in ec.FIELDS:
        assert torch.equal(b[f], data[f]), f
    assert torch.equal(a["states"][:1000], data["states"][perm[:1000]])
