"""CPU suite: the C-ABI library loads without a GPU and exports every symbol include/gigaenv.h declares."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "gigaenv.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(ggl_[a-z_0-9]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(ggl):
    L = ggl.lib()
    syms = declared_symbols()
    assert len(syms) >= 20
    missing = [s for s in syms if not hasattr(L, s)]
    assert not missing, missing


def test_struct_sizes_match_header(ggl):
    # the ctypes mirror must agree with the compiled C structs
    L = ggl.lib()
    L.ggl_abi_sizeof.restype = ctypes.c_size_t
    L.ggl_abi_sizeof.argtypes = [ctypes.c_int]
    abi = ggl.abi
    for i, cls in enumerate([abi.EnvSetDesc, abi.ArenaState, abi.CarState, abi.BallState, abi.PadState,
                             abi.EnvSetViews, abi.RewardDesc, abi.TerminalDesc, abi.Controls, abi.WheelState]):
        assert L.ggl_abi_sizeof(i) == ctypes.sizeof(cls), cls.__name__


def test_version_and_error_strings(ggl):
    L = ggl.lib()
    assert b"sm_100a" in L.ggl_version()
    assert isinstance(L.ggl_last_error(), bytes)


def test_argument_validation_happens_before_any_device_call(ggl):
    """Bad arguments are refused with GGL_ERR_INVALID and a message, without touching CUDA (so this runs without a GPU)."""
    import numpy as np
    L = ggl.lib()
    buf = np.zeros(64, np.float32)
    p = buf.ctypes.data
    rc = L.ggl_gae_compute(p, p, p, None, 16, 0, -0.5, 0.95, 1.0, 0.0, p, p, p, None, None, p, None)
    assert rc == -1 and b"gamma" in L.ggl_last_error()
    rc = L.ggl_gae_compute(p, p, p, None, -1, 0, 0.99, 0.95, 1.0, 0.0, p, p, p, None, None, p, None)
    assert rc == -1
    assert L.ggl_envset_create(None, None) == -1
    assert L.ggl_envset_step(None, None, None) == -1


def test_host_action_picker_respects_masks(ggl):
    import numpy as np
    L = ggl.lib()
    rng = np.random.default_rng(3)
    n, na = 500, 90
    masks = (rng.random((n, na)) < 0.3).astype(np.uint8)
    masks[7] = 0                       # no valid action: index 0 like the device helper
    acts = np.full(n, -1, np.int32)
    L.ggl_host_random_actions.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_uint64]
    assert L.ggl_host_random_actions(masks.ctypes.data, acts.ctypes.data, n, na, 42) == 0
    ok = masks[np.arange(n), acts] != 0
    ok[7] = acts[7] == 0
    assert ok.all()
    acts2 = np.zeros(n, np.int32)
    L.ggl_host_random_actions(masks.ctypes.data, acts2.ctypes.data, n, na, 43)
    assert (acts != acts2).any()
    # uniform over the valid set: every valid action of a dense row is eventually chosen
    row = np.ones((1, na), np.uint8); seen = set(); a = np.zeros(1, np.int32)
    for s in range(2000):
        L.ggl_host_random_actions(row.ctypes.data, a.ctypes.data, 1, na, s)
        seen.add(int(a[0]))
    assert len(seen) == na
