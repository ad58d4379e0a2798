"""GPU parity of the device obs statistics (ggl_obsstat_*) against the numpy oracle, the committed reference vectors and, when it
travelled to this box, the compiled reference BatchedWelfordStat. Bit-exact: double-precision Welford in the reference's order,
float32 multiply and add without contraction."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import welford_oracle as wo     # noqa: E402
import obsstat_cases as oc      # noqa: E402
import oracle_ref               # noqa: E402


@pytest.mark.parametrize("ci", range(len(oc.CASES)))
def test_device_statistics_match_oracle_and_golden(ggl, ci):
    import torch
    case = oc.CASES[ci]
    rows, width, ns, nsteps, mc, ms = case
    g = np.load(os.path.join(ROOT, "tests", "golden", "obsstat_golden.npz"))
    dev, st = ggl.ObsStat(width), wo.BatchedWelfordStat(width)
    for si, (obs, idx) in enumerate(oc.steps(case, seed=ci)):
        d_obs = torch.from_numpy(obs.copy()).cuda()
        dev.increment_rows(d_obs, torch.from_numpy(idx).cuda())
        dev.normalize_in_place(d_obs, mc, ms)
        st.increment_rows(obs, idx); st.normalize_in_place(obs, mc, ms)
        m, v, c = dev.state()
        assert c == st.count and np.array_equal(m, st.means) and np.array_equal(v, st.vars)
        got = d_obs.cpu().numpy()
        assert got.tobytes() == obs.tobytes()
        assert np.array_equal(m, g["c%d_s%d_means" % (ci, si)]) and np.array_equal(got[:4], g["c%d_s%d_obs_head" % (ci, si)])
    dev.close()


@pytest.mark.skipif(not oracle_ref.welford_available(), reason="compiled reference not present on this box")
def test_device_statistics_match_compiled_reference(ggl):
    import torch
    case = (4096, 167, 100, 8, 3.0, 0.1)
    ref, dev = oracle_ref.RefWelford(167), ggl.ObsStat(167)
    for obs, idx in oc.steps(case, seed=7):
        d_obs = torch.from_numpy(obs.copy()).cuda()
        dev.increment_rows(d_obs, torch.from_numpy(idx).cuda()); dev.normalize_in_place(d_obs, 3.0, 0.1)
        ref.increment_rows(obs, idx); ref.normalize_in_place(obs, 3.0, 0.1)
        assert d_obs.cpu().numpy().tobytes() == obs.tobytes()
    m, v, c = dev.state(); rm, rv, rc = ref.state()
    assert c == rc and np.array_equal(m, rm) and np.array_equal(v, rv)
    ref.close(); dev.close()


def test_state_round_trip_and_envset_obs(ggl):
    """Checkpoint round trip (ToJSON / ReadFromJSON) and normalising the envset's own obs buffer in place."""
    import torch
    env = ggl.EnvSet(32, 2, obs_kind=ggl.abi.OBS_ADVANCED, seed=2)
    a, b = ggl.ObsStat(env.obs_size), ggl.ObsStat(env.obs_size)
    idx = torch.arange(0, env.num_players, 3, dtype=torch.int32, device="cuda")
    a.increment_rows(env.obs, idx)
    b.set_state(*a.state())
    raw = env.obs.clone()
    a.normalize_in_place(env.obs, 3.0, 0.1)
    b.normalize_in_place(raw, 3.0, 0.1)
    assert torch.equal(env.obs, raw) and torch.isfinite(raw).all()
    st = wo.BatchedWelfordStat(env.obs_size); m, v, c = a.state()
    assert c == len(idx)
    a.close(); b.close(); env.close()
