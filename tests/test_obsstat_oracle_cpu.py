"""Pins the numpy restatement of BatchedWelfordStat (oracle/welford_oracle.py) bit-exact against the UNMODIFIED reference
(oracle/_ref/libref_welford.so) when it is present, and against the committed golden vectors always."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import welford_oracle as wo     # noqa: E402
import obsstat_cases as oc      # noqa: E402
import oracle_ref               # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden", "obsstat_golden.npz")


def test_oracle_matches_golden_vectors():
    g = np.load(GOLDEN)
    for ci, case in enumerate(oc.CASES):
        rows, width, ns, nsteps, mc, ms = case
        st = wo.BatchedWelfordStat(width)
        for si, (obs, idx) in enumerate(oc.steps(case, seed=ci)):
            st.increment_rows(obs, idx)
            st.normalize_in_place(obs, mc, ms)
            assert st.count == int(g["c%d_s%d_count" % (ci, si)])
            assert np.array_equal(st.means, g["c%d_s%d_means" % (ci, si)]) and np.array_equal(st.vars, g["c%d_s%d_vars" % (ci, si)])
            assert np.array_equal(obs[:4], g["c%d_s%d_obs_head" % (ci, si)])
            assert np.array_equal(obs.astype(np.float64).sum(0), g["c%d_s%d_obs_sum" % (ci, si)])


@pytest.mark.skipif(not oracle_ref.welford_available(), reason="compiled reference not present")
def test_oracle_matches_compiled_reference():
    for ci, case in enumerate(oc.CASES):
        rows, width, ns, nsteps, mc, ms = case
        ref, st = oracle_ref.RefWelford(width), wo.BatchedWelfordStat(width)
        for obs, idx in oc.steps(case, seed=100 + ci):
            a, b = obs.copy(), obs.copy()
            ref.increment_rows(a, idx); ref.normalize_in_place(a, mc, ms)
            st.increment_rows(b, idx); st.normalize_in_place(b, mc, ms)
            m, v, c = ref.state()
            assert c == st.count and np.array_equal(m, st.means) and np.array_equal(v, st.vars)
            assert np.array_equal(a, b)
        ref.close()


def test_no_normalisation_before_two_samples():
    st = wo.BatchedWelfordStat(3)
    obs = np.array([[1.0, -0.0, 5.0]], np.float32)
    st.increment_rows(obs, [0])
    before = obs.copy()
    st.normalize_in_place(obs, 3.0, 0.1)                     # WelfordStat.h:133
    assert obs.tobytes() == before.tobytes()
