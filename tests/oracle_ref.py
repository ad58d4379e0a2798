"""TEST INFRASTRUCTURE: ctypes binding to oracle/_ref/libref_oracle.so and libref_gae.so — the UNMODIFIED
reference compiled by oracle/Makefile. Never imported by the product package."""
import ctypes as C
import importlib
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
abi = importlib.import_module("reinforcement-learning_b200.abi")
arena_mesh = importlib.import_module("reinforcement-learning_b200.arena_mesh")

_REF_DIR = os.path.join(ROOT, "oracle", "_ref")
_lib = None
_gae = None


def available():
    return os.path.exists(os.path.join(_REF_DIR, "libref_oracle.so"))


def gae_available():
    return os.path.exists(os.path.join(_REF_DIR, "libref_gae.so"))


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(os.path.join(_REF_DIR, "libref_oracle.so"))
        L.ref_envset_create.restype = C.c_void_p
        L.ref_envset_create.argtypes = [C.POINTER(abi.EnvSetDesc)]
        L.ref_envset_bench.restype = C.c_double
        L.ref_envset_bench.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint32]
        L.ref_seed.argtypes = [C.c_uint64]
        for name in ("ref_envset_destroy", "ref_envset_sync", "ref_envset_reset"):
            getattr(L, name).argtypes = [C.c_void_p]
        L.ref_envset_info.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.ref_envset_step_first_half.argtypes = [C.c_void_p, C.c_int]
        L.ref_envset_step_second_half.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_envset_step.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_envset_reset_arena.argtypes = [C.c_void_p, C.c_int]
        L.ref_envset_copy_out.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.ref_envset_slot_info.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_envset_events.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_envset_last_rewards.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_envset_get_state.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.ArenaState)]
        L.ref_envset_set_state.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.ArenaState)]
        L.ref_envset_set_reset_state.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.ArenaState)]
        L.ref_arena_step.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.ref_envset_set_controls.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.ArenaState)]
        L.ref_action_table.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
        L.ref_rng_peek.argtypes = [C.c_int]
        L.ref_rng_peek.restype = C.c_uint32
        meshes = arena_mesh.soccar_meshes()
        arr = (abi.MeshBlob * len(meshes))()
        keep = []
        for i, b in enumerate(meshes):
            buf = C.create_string_buffer(b, len(b))
            keep.append(buf)
            arr[i].data = C.cast(buf, C.c_void_p)
            arr[i].size = len(b)
        assert L.ref_init(arr, len(meshes)) == 0
        L._keep = keep
        _lib = L
    return _lib


class RefEnvSet:
    """The reference's RLGC::EnvSet behind the harness ABI."""

    def __init__(self, num_arenas, players_per_team, **kw):
        L = lib()
        self.desc, self._keep = abi.make_desc(num_arenas, players_per_team, **kw)
        self.h = L.ref_envset_create(C.byref(self.desc))
        assert self.h, "ref_envset_create failed"
        a, b, c = C.c_int(), C.c_int(), C.c_int()
        L.ref_envset_info(self.h, C.byref(a), C.byref(b), C.byref(c))
        self.num_players, self.obs_size, self.num_actions = a.value, b.value, c.value
        self.num_arenas = num_arenas
        self.num_rewards = self.desc.num_rewards

    def close(self):
        if self.h:
            lib().ref_envset_destroy(self.h)
            self.h = None

    def step(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.int32)
        lib().ref_envset_step(self.h, a.ctypes.data)

    def step_first_half(self):
        lib().ref_envset_step_first_half(self.h, 0)

    def step_second_half(self, actions):
        a = np.ascontiguousarray(actions, dtype=np.int32)
        lib().ref_envset_step_second_half(self.h, a.ctypes.data)

    def reset(self):
        lib().ref_envset_reset(self.h)

    def reset_arena(self, i):
        lib().ref_envset_reset_arena(self.h, i)

    def outputs(self):
        obs = np.empty((self.num_players, self.obs_size), np.float32)
        masks = np.empty((self.num_players, self.num_actions), np.uint8)
        rew = np.empty(self.num_players, np.float32)
        term = np.empty(self.num_arenas, np.uint8)
        lib().ref_envset_copy_out(self.h, obs.ctypes.data, masks.ctypes.data, rew.ctypes.data, term.ctypes.data)
        return obs, masks, rew, term

    def slot_info(self):
        ids = np.empty(self.num_players, np.int32)
        teams = np.empty(self.num_players, np.uint8)
        lib().ref_envset_slot_info(self.h, ids.ctypes.data, teams.ctypes.data)
        return ids, teams

    def events(self):
        ev = np.empty((self.num_players, 9), np.uint8)
        lib().ref_envset_events(self.h, ev.ctypes.data)
        return ev

    def last_rewards(self):
        out = np.empty((self.num_arenas, self.num_rewards), np.float32)
        lib().ref_envset_last_rewards(self.h, out.ctypes.data)
        return out

    def get_state(self, arena):
        st = abi.ArenaState()
        lib().ref_envset_get_state(self.h, arena, C.byref(st))
        return st

    def set_state(self, arena, st):
        lib().ref_envset_set_state(self.h, arena, C.byref(st))

    def set_controls(self, arena, st):
        lib().ref_envset_set_controls(self.h, arena, C.byref(st))

    def set_reset_state(self, arena, st):
        lib().ref_envset_set_reset_state(self.h, arena, C.byref(st))

    def arena_step(self, ticks, arena=-1):
        lib().ref_arena_step(self.h, arena, ticks)

    def bench(self, steps, warmup, seed=123):
        return lib().ref_envset_bench(self.h, steps, warmup, seed)


def action_table():
    t = np.empty((90, 8), np.float32)
    n = C.c_int()
    lib().ref_action_table(t.ctypes.data, C.byref(n))
    return t[:n.value]


def gae(rews, terms, vals, trunc_vals, gamma=0.99, lam=0.95, return_std=0.0, clip_range=10.0):
    global _gae
    if _gae is None:
        import torch  # noqa: F401  (libtorch must be loaded first)
        _gae = C.CDLL(os.path.join(_REF_DIR, "libref_gae.so"))
        _gae.ref_gae_compute.argtypes = [C.c_void_p] * 4 + [C.c_int64, C.c_int64] + [C.c_float] * 4 + [C.c_void_p] * 3 + [C.POINTER(C.c_float)]
    rews = np.ascontiguousarray(rews, np.float32)
    terms = np.ascontiguousarray(terms, np.int8)
    vals = np.ascontiguousarray(vals, np.float32)
    L = len(rews)
    adv, ret, tgt = (np.empty(L, np.float32) for _ in range(3))
    clip = C.c_float()
    tv = None
    K = 0
    if trunc_vals is not None:
        tvv = np.ascontiguousarray(trunc_vals, np.float32)
        tv = tvv.ctypes.data
        K = len(tvv)
    rc = _gae.ref_gae_compute(rews.ctypes.data, terms.ctypes.data, vals.ctypes.data, tv, L, K,
                              gamma, lam, return_std, clip_range, adv.ctypes.data, ret.ctypes.data, tgt.ctypes.data,
                              C.byref(clip))
    if rc != 0:
        raise RuntimeError("ref_gae_compute failed")
    return adv, ret, tgt, clip.value


# ---- the reference ExperienceBuffer (oracle/_ref/libref_expbuf.so)
_expbuf = None


def expbuf_available():
    return os.path.exists(os.path.join(_REF_DIR, "libref_expbuf.so"))


class RefExpBuf:
    """GGL::ExperienceBuffer on CPU tensors; keeps its shuffled index vector between epochs like the reference."""

    def __init__(self, seed, max_action_index=-1):
        global _expbuf
        if _expbuf is None:
            _expbuf = C.CDLL(os.path.join(_REF_DIR, "libref_expbuf.so"))
            _expbuf.ref_expbuf_create.restype = C.c_void_p
            _expbuf.ref_expbuf_create.argtypes = [C.c_int, C.c_int64]
            _expbuf.ref_expbuf_destroy.argtypes = [C.c_void_p]
            _expbuf.ref_expbuf_set_data.argtypes = [C.c_void_p] + [C.c_int64] * 4 + [C.c_void_p] * 6
            _expbuf.ref_expbuf_get_batches.restype = C.c_int64
            _expbuf.ref_expbuf_get_batches.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int64] + [C.c_void_p] * 7
        self.h = _expbuf.ref_expbuf_create(seed, max_action_index)

    def close(self):
        if self.h:
            _expbuf.ref_expbuf_destroy(self.h)
            self.h = None

    def set_data(self, d):
        self.rows, self.obs = d["states"].shape
        self.na = d["action_masks"].shape[1]
        a = [np.ascontiguousarray(d[k]) for k in ("states", "actions", "log_probs", "target_values", "action_masks", "advantages")]
        _expbuf.ref_expbuf_set_data(self.h, self.rows, a[1].shape[0], self.obs, self.na, *[x.ctypes.data for x in a])

    def get_batches(self, batch_size, overbatching):
        n = self.rows
        sizes = np.zeros(n + 2, np.int64)
        out = {"states": np.zeros((n, self.obs), np.float32), "actions": np.zeros(n, np.int32), "log_probs": np.zeros(n, np.float32),
               "target_values": np.zeros(n, np.float32), "action_masks": np.zeros((n, self.na), np.uint8), "advantages": np.zeros(n, np.float32)}
        nb = _expbuf.ref_expbuf_get_batches(self.h, batch_size, int(overbatching), n + 2, sizes.ctypes.data,
                                            *[out[k].ctypes.data for k in ("states", "actions", "log_probs", "target_values", "action_masks", "advantages")])
        if nb < 0:
            raise RuntimeError("ref_expbuf_get_batches failed")
        sizes = sizes[:nb]
        used = int(sizes.sum())
        return sizes, {k: v[:used] for k, v in out.items()}


# ---- the reference BatchedWelfordStat (oracle/_ref/libref_welford.so)
_welford = None


def welford_available():
    return os.path.exists(os.path.join(_REF_DIR, "libref_welford.so"))


class RefWelford:
    """GGL::BatchedWelfordStat, unmodified (WelfordStat.h:69-243)."""

    def __init__(self, width):
        global _welford
        if _welford is None:
            _welford = C.CDLL(os.path.join(_REF_DIR, "libref_welford.so"))
            _welford.ref_welford_create.restype = C.c_void_p
            _welford.ref_welford_create.argtypes = [C.c_int]
            _welford.ref_welford_destroy.argtypes = [C.c_void_p]
            _welford.ref_welford_increment_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
            _welford.ref_welford_normalize.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double]
            _welford.ref_welford_state.restype = C.c_int64
            _welford.ref_welford_state.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self.width = width
        self.h = C.c_void_p(_welford.ref_welford_create(width))

    def increment_rows(self, obs, idx):
        idx = np.ascontiguousarray(idx, np.int32)
        _welford.ref_welford_increment_rows(self.h, obs.ctypes.data, self.width, idx.ctypes.data, len(idx))

    def normalize_in_place(self, obs, mean_clamp, min_std):
        _welford.ref_welford_normalize(self.h, obs.ctypes.data, obs.shape[0], self.width, mean_clamp, min_std)

    def state(self):
        m, v = np.zeros(self.width), np.zeros(self.width)
        c = _welford.ref_welford_state(self.h, m.ctypes.data, v.ctypes.data)
        return m, v, int(c)

    def close(self):
        if self.h:
            _welford.ref_welford_destroy(self.h)
            self.h = None
