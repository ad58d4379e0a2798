"""TEST INFRASTRUCTURE: ctypes binding to tests/hostsim/_build/libhostsim.so — the g++ compile of the device simulator
headers, used to lock-step the tick logic against the compiled reference in a container without a GPU."""
import ctypes as C
import importlib
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
abi = importlib.import_module("reinforcement-learning_b200.abi")
arena_mesh = importlib.import_module("reinforcement-learning_b200.arena_mesh")
_DIR = os.path.join(ROOT, "tests", "hostsim")
_SO = os.path.join(_DIR, "_build", "libhostsim.so")
_lib = None


def build():
    os.makedirs(os.path.join(_DIR, "_build"), exist_ok=True)
    csrc = os.path.join(ROOT, "reinforcement-learning_b200", "csrc")
    deps = [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith(".cuh")] + [os.path.join(_DIR, "hostsim.cpp")]
    if os.path.exists(_SO) and all(os.path.getmtime(d) <= os.path.getmtime(_SO) for d in deps):
        return
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-shared", "-fPIC", "-w",
                           os.path.join(_DIR, "hostsim.cpp"), "-o", _SO])


def order_mode(mode):
    """0 = as shipped, 1 = predict the reference's manifold array on every tick with a contact, 2 = never reorder"""
    lib().hostsim_order_mode(int(mode))


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.hostsim_create.restype = C.c_void_p
        L.hostsim_create.argtypes = [C.c_void_p, C.c_int]
        L.hostsim_destroy.argtypes = [C.c_void_p]
        L.hostsim_set_state.argtypes = [C.c_void_p, C.POINTER(abi.ArenaState)]
        L.hostsim_get_state.argtypes = [C.c_void_p, C.POINTER(abi.ArenaState)]
        L.hostsim_step.argtypes = [C.c_void_p, C.c_int]
        L.hostsim_set_controls.argtypes = [C.c_void_p, C.POINTER(abi.ArenaState)]
        L.hostsim_seed.argtypes = [C.c_void_p, C.c_uint32]
        L.hostsim_contacts.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.hostsim_sizeof_arena.restype = C.c_size_t
        _lib = L
    return _lib


class HostSim:
    def __init__(self, meshes=None):
        L = lib()
        meshes = arena_mesh.soccar_meshes() if meshes is None else meshes
        arr = (abi.MeshBlob * max(1, len(meshes)))()
        self._keep = []
        for i, b in enumerate(meshes):
            buf = C.create_string_buffer(b, len(b))
            self._keep.append(buf)
            arr[i].data = C.cast(buf, C.c_void_p)
            arr[i].size = len(b)
        self.h = L.hostsim_create(arr, len(meshes))
        assert self.h

    def set_state(self, st):
        lib().hostsim_set_state(self.h, C.byref(st))

    def get_state(self):
        st = abi.ArenaState()
        lib().hostsim_get_state(self.h, C.byref(st))
        return st

    def set_controls(self, st):
        lib().hostsim_set_controls(self.h, C.byref(st))

    def set_slot_order(self, ids):
        ids = np.ascontiguousarray(ids, np.int32)
        lib().hostsim_set_slot_order.argtypes = [C.c_void_p, C.c_void_p]
        lib().hostsim_set_slot_order(self.h, ids.ctypes.data)

    def step(self, ticks=1):
        lib().hostsim_step(self.h, ticks)

    def seed(self, s):
        lib().hostsim_seed(self.h, s)

    def set_game_mode(self, mode):
        lib().hostsim_set_game_mode.argtypes = [C.c_void_p, C.c_int]
        lib().hostsim_set_game_mode(self.h, mode)

    def predicted_manifolds(self, maxn=64):
        """(array [n][3] of body0, body1, island key; reordered?, unknown row?) of the last tick, or None if it predicted nothing"""
        out = np.zeros((maxn, 3), np.int32)
        flags = np.zeros(2, np.int32)
        n = lib().hostsim_predicted_manifolds(C.c_void_p(self.h), out.ctypes.data_as(C.c_void_p), maxn, flags.ctypes.data_as(C.c_void_p))
        return None if n < 0 else (out[:n], bool(flags[0]), bool(flags[1]))

    def contacts(self, maxn=64):
        out = np.zeros((maxn, 20), np.float32)
        n = lib().hostsim_contacts(self.h, out.ctypes.data, maxn)
        return out[:n]


class _T:
    """numpy array posing as the torch tensor the product EnvSet exposes (.cpu().numpy())."""

    def __init__(self, a):
        self.a = a

    def cpu(self):
        return self

    def numpy(self):
        return self.a


class HostEnvSet:
    """Same surface as the product's EnvSet (subset used by tests/env_parity.py), backed by the host compile."""

    def __init__(self, num_arenas, players_per_team, **kw):
        L = lib()
        L.hostenv_create.restype = C.c_void_p
        L.hostenv_create.argtypes = [C.POINTER(abi.EnvSetDesc)]
        for n in ("hostenv_set_state", "hostenv_get_state", "hostenv_set_reset_state"):
            getattr(L, n).argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.ArenaState)]
        L.hostenv_set_slot_order.argtypes = [C.c_void_p, C.c_void_p]
        L.hostenv_step.argtypes = [C.c_void_p, C.c_void_p]
        L.hostenv_copy_out.argtypes = [C.c_void_p] * 7
        L.hostenv_destroy.argtypes = [C.c_void_p]
        kw.pop("device", None)
        self.desc, self._keep = abi.make_desc(num_arenas, players_per_team, **kw)
        self.h = L.hostenv_create(C.byref(self.desc))
        self.num_arenas = num_arenas
        self.num_players = num_arenas * 2 * players_per_team
        if self.desc.obs_kind == abi.OBS_DEFAULT_PADDED:
            self.obs_size = 51 + 38 * self.desc.obs_max_players
        else:
            self.obs_size = 51 + (29 if self.desc.obs_kind == abi.OBS_ADVANCED else 19) * 2 * players_per_team
        self.num_actions = 90
        self.num_rewards = self.desc.num_rewards
        self.device = "cpu"
        self._refresh()

    def _refresh(self):
        o = np.empty((self.num_players, self.obs_size), np.float32)
        m = np.empty((self.num_players, 90), np.uint8)
        r = np.empty(self.num_players, np.float32)
        t = np.empty(self.num_arenas, np.uint8)
        e = np.empty((self.num_players, 9), np.uint8)
        lr = np.zeros((self.num_arenas, max(1, self.num_rewards)), np.float32)
        lib().hostenv_copy_out(self.h, o.ctypes.data, m.ctypes.data, r.ctypes.data, t.ctypes.data, e.ctypes.data, lr.ctypes.data)
        self.obs, self.action_masks, self.rewards, self.terminals, self.events, self.last_rewards = _T(o), _T(m), _T(r), _T(t), _T(e), _T(lr)

    def set_state(self, a, st):
        lib().hostenv_set_state(self.h, a, C.byref(st))

    def get_state(self, a):
        st = abi.ArenaState()
        lib().hostenv_get_state(self.h, a, C.byref(st))
        return st

    def set_reset_state(self, a, st):
        lib().hostenv_set_reset_state(self.h, a, C.byref(st))

    def set_slot_order(self, ids):
        ids = np.ascontiguousarray(ids, np.int32)
        lib().hostenv_set_slot_order(self.h, ids.ctypes.data)

    def set_rng(self, a, state):
        lib().hostenv_set_rng.argtypes = [C.c_void_p, C.c_int, C.c_uint32]
        lib().hostenv_set_rng(self.h, a, state)

    def get_rng(self, a):
        lib().hostenv_get_rng.restype = C.c_uint32
        lib().hostenv_get_rng.argtypes = [C.c_void_p, C.c_int]
        return int(lib().hostenv_get_rng(self.h, a))

    def contact_overflow(self):
        lib().hostenv_contact_overflow.restype = C.c_uint64
        lib().hostenv_contact_overflow.argtypes = [C.c_void_p]
        return int(lib().hostenv_contact_overflow(self.h))

    def reset_arena(self, a=-1):
        lib().hostenv_reset_arena.argtypes = [C.c_void_p, C.c_int]
        lib().hostenv_reset_arena(self.h, a)
        self._refresh()

    def step_np(self, acts):
        acts = np.ascontiguousarray(acts, np.int32)
        lib().hostenv_step(self.h, acts.ctypes.data)
        self._refresh()

    def sync(self):
        pass

    def close(self):
        if self.h:
            lib().hostenv_destroy(self.h)
            self.h = None
