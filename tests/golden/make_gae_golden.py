"""Generates tests/golden/gae_golden.npz by running the UNMODIFIED reference GAE.cpp (oracle/_ref/libref_gae.so,
built from /root/reference by oracle/Makefile). Run in the build container: python tests/golden/make_gae_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import gae_cases  # noqa: E402
import oracle_ref  # noqa: E402

out = {}
for (name, T, N, kind, ragged, g, lam, std, clip, pass_tv) in gae_cases.GOLDEN_CASES:
    r, t, v, tv = gae_cases.make_stream(T, N, kind, ragged=ragged)
    adv, ret, tgt, cp = oracle_ref.gae(r, t, v, tv if pass_tv else None, g, lam, std, clip)
    out[name + "/adv"] = adv
    out[name + "/ret"] = ret
    out[name + "/tgt"] = tgt
    out[name + "/clip"] = np.float32(cp)
np.savez_compressed(os.path.join(HERE, "gae_golden.npz"), **out)
print("wrote", len(out), "arrays")
