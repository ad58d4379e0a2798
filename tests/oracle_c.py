"""TEST INFRASTRUCTURE: ctypes binding to the plain-C restatement oracle/_build/liboracle_c.so (built on demand
with gcc from oracle/*.c — no reference sources needed, so it also works on the GPU box)."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(ROOT, "oracle", "_build", "liboracle_c.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        srcs = [os.path.join(ROOT, "oracle", f) for f in os.listdir(os.path.join(ROOT, "oracle")) if f.endswith(".c")]
        if not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
            subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "restate"])
        L = C.CDLL(_SO)
        L.oracle_gae_compute.argtypes = [C.c_void_p] * 4 + [C.c_int64, C.c_int64] + [C.c_float] * 4 + [C.c_void_p] * 3 + [C.POINTER(C.c_float)]
        _lib = L
    return _lib


def gae(rews, terms, vals, trunc_vals, gamma=0.99, lam=0.95, return_std=0.0, clip_range=10.0):
    rews = np.ascontiguousarray(rews, np.float32)
    terms = np.ascontiguousarray(terms, np.int8)
    vals = np.ascontiguousarray(vals, np.float32)
    L = len(rews)
    adv, ret, tgt = (np.empty(L, np.float32) for _ in range(3))
    clip = C.c_float()
    tv, K = None, 0
    if trunc_vals is not None:
        tvv = np.ascontiguousarray(trunc_vals, np.float32)
        K = len(tvv)
        tv = tvv.ctypes.data if K > 0 else rews.ctypes.data
    rc = lib().oracle_gae_compute(rews.ctypes.data, terms.ctypes.data, vals.ctypes.data, tv, L, K,
                                  gamma, lam, return_std, clip_range, adv.ctypes.data, ret.ctypes.data, tgt.ctypes.data,
                                  C.byref(clip))
    return adv, ret, tgt, clip.value, rc
