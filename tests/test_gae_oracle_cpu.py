"""CPU suite: pins the plain-C GAE restatement (oracle/gae_oracle.c) against (1) golden vectors produced by the
unmodified reference GAE.cpp and (2) the compiled reference itself when oracle/_ref/libref_gae.so is present."""
import os

import numpy as np
import pytest

import gae_cases
import oracle_c
import oracle_ref

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "gae_golden.npz"))


@pytest.mark.parametrize("case", gae_cases.GOLDEN_CASES, ids=[c[0] for c in gae_cases.GOLDEN_CASES])
def test_c_oracle_matches_golden_bit_exact(case):
    name, T, N, kind, ragged, g, lam, std, clip, pass_tv = case
    r, t, v, tv = gae_cases.make_stream(T, N, kind, ragged=ragged)
    adv, ret, tgt, cp, rc = oracle_c.gae(r, t, v, tv if pass_tv else None, g, lam, std, clip)
    assert rc == 0
    # same scalar float ops in the same order, both built without FMA: bit-exact
    assert np.array_equal(adv, GOLD[name + "/adv"])
    assert np.array_equal(ret, GOLD[name + "/ret"])
    assert np.array_equal(tgt, GOLD[name + "/tgt"])
    assert np.float32(cp) == GOLD[name + "/clip"]


@pytest.mark.skipif(not oracle_ref.gae_available(), reason="oracle/_ref/libref_gae.so not built")
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_c_oracle_matches_compiled_reference(seed):
    r, t, v, tv = gae_cases.make_stream(37, 211, "mixed", seed=seed, ragged=True)
    a0 = oracle_ref.gae(r, t, v, tv, 0.99, 0.95, 1.7, 10.0)
    a1 = oracle_c.gae(r, t, v, tv, 0.99, 0.95, 1.7, 10.0)
    for x, y in zip(a0[:3], a1[:3]):
        assert np.array_equal(x, y)
    assert np.float32(a0[3]) == np.float32(a1[3])


def test_c_oracle_known_answers():
    # hand-checkable: gamma=lambda=1, values 0 -> advantages == returns == suffix sums inside each segment
    r = np.array([1, 2, 3, 4, 5], np.float32)
    t = np.array([0, 0, 1, 0, 0], np.int8)
    v = np.zeros(5, np.float32)
    adv, ret, tgt, cp, rc = oracle_c.gae(r, t, v, None, 1.0, 1.0, 0.0, 0.0)
    assert rc == 0
    assert ret.tolist() == [6, 5, 3, 9, 5]
    assert adv.tolist() == [6, 5, 3, 9, 5]
    assert tgt.tolist() == [6, 5, 3, 9, 5]
    # truncation bootstraps the ADVANTAGE from truncValPreds but never the RETURN (GAE.cpp:180-182)
    t2 = np.array([0, 0, 2, 0, 0], np.int8)
    adv, ret, tgt, cp, rc = oracle_c.gae(r, t2, v, np.array([10.0], np.float32), 1.0, 1.0, 0.0, 0.0)
    assert rc == 0
    assert ret.tolist() == [6, 5, 3, 9, 5]
    assert adv.tolist() == [16, 15, 13, 9, 5]
    # truncation count mismatch is fatal in the reference (GAE.cpp:196-197)
    *_, rc = oracle_c.gae(r, t2, v, np.array([10.0, 1.0], np.float32), 1.0, 1.0, 0.0, 0.0)
    assert rc == -1


def test_empty_stream():
    e = np.zeros(0, np.float32)
    adv, ret, tgt, cp, rc = oracle_c.gae(e, np.zeros(0, np.int8), e, None)
    assert rc == 0 and len(adv) == 0 and cp == 0
