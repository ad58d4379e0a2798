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
