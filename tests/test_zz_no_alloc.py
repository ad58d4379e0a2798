"""GPU test of ggl_gae_compute_static (include/gigaenv.h): the allocation-free variant of the GAE scan, whose scratch is a
file-scope static __device__ array instead of a caller workspace. No counterpart in the reference; checked against the existing
entry point ggl_gae_compute on the same inputs, on two calls in a row (the array keeps the last call's contents), with the
tolerance of tests/test_gae_gpu.py. Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

import gae_cases

pytestmark = pytest.mark.gpu
RTOL = 1e-4
MAX_L = 1 << 24


def close(x, ref):
    return np.all(np.abs(x - ref) <= RTOL * np.maximum(np.abs(ref), 1.0))


def _proto(L):
    args = [C.c_void_p] * 4 + [C.c_int64, C.c_int64] + [C.c_float] * 4 + [C.c_void_p] * 5
    L.ggl_gae_compute.argtypes = args + [C.c_void_p, C.c_void_p]
    L.ggl_gae_compute_static.argtypes = args + [C.c_void_p]
    L.ggl_gae_compute.restype = L.ggl_gae_compute_static.restype = C.c_int
    L.ggl_gae_workspace_bytes.restype = C.c_size_t
    L.ggl_gae_workspace_bytes.argtypes = [C.c_int64]
    L.ggl_last_error.restype = C.c_char_p


def _run(lib, torch, r, t, v, tv, std, clip, static):
    """One call through the C ABI on device copies of the inputs; every buffer is the caller's (torch is only the allocator)."""
    dev = "cuda:0"
    n = r.size
    d_r, d_t, d_v = (torch.from_numpy(x).to(dev) for x in (r, t, v))
    d_tv = torch.from_numpy(tv).to(dev) if tv is not None and tv.size else None
    K = 0 if tv is None else tv.size
    p_tv = None if tv is None else (d_tv.data_ptr() if d_tv is not None else d_r.data_ptr())
    outs = [torch.full((n,), float("nan"), dtype=torch.float32, device=dev) for _ in range(3)]
    scal = torch.zeros(2, dtype=torch.float32, device=dev)
    s = torch.cuda.current_stream(dev).cuda_stream
    head = (d_r.data_ptr(), d_t.data_ptr(), d_v.data_ptr(), p_tv, n, K, 0.99, 0.95, std, clip,
            outs[0].data_ptr(), outs[1].data_ptr(), outs[2].data_ptr(), scal.data_ptr(), scal.data_ptr() + 4)
    if static:
        rc = lib.ggl_gae_compute_static(*head, s)
    else:
        ws = torch.empty(lib.ggl_gae_workspace_bytes(n), dtype=torch.uint8, device=dev)
        rc = lib.ggl_gae_compute(*head, ws.data_ptr(), s)
    assert rc == 0, lib.ggl_last_error()
    torch.cuda.synchronize()
    return [o.cpu().numpy() for o in outs] + [scal[:1].cpu().numpy(), scal[1:].view(torch.int32).cpu().numpy()]


def test_static_scratch_gae_matches_workspace_gae_twice_in_a_row(ggl):
    import torch
    lib = ggl.lib()
    _proto(lib)
    # two calls in a row with different sizes and segment structure: the second one finds the first one's records in the
    # array (more tiles, then fewer), and must not be fooled by them
    cases = [(256, 1024, "mixed", 1.7, 10.0, 11), (100, 3001, "trunc", 0.25, 1.5, 12), (64, 1024, "normal", 0.0, 10.0, 13)]
    for T, N, kind, std, clip, seed in cases:
        r, t, v, tv = gae_cases.make_stream(T, N, kind, seed=seed, ragged=(N % 2 == 1))
        ref = _run(lib, torch, r, t, v, tv, std, clip, static=False)
        for _ in range(2):
            got = _run(lib, torch, r, t, v, tv, std, clip, static=True)
            for a, b in zip(got[:3], ref[:3]):
                assert close(a, b)
            assert abs(float(got[3][0]) - float(ref[3][0])) <= 1e-5
            assert int(got[4][0]) == int(ref[4][0])


def test_static_scratch_gae_rejects_sizes_beyond_its_maximum(ggl):
    import torch
    lib = ggl.lib()
    _proto(lib)
    d = torch.zeros(64, dtype=torch.float32, device="cuda:0")
    p = d.data_ptr()
    rc = lib.ggl_gae_compute_static(p, p, p, None, MAX_L + 1, 0, 0.99, 0.95, 1.0, 0.0, p, p, p, None, None, None)
    assert rc == -1
    assert b"GGL_GAE_STATIC_MAX_L" in lib.ggl_last_error()
    torch.cuda.synchronize()
