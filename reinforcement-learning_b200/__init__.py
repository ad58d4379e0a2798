"""gigaenv-b200: B200-native rollout hot path (arena stepping, obs/reward/terminal, GAE) behind a C ABI.

This Python module is a thin ctypes binding over ``libgigaenv_cuda.so`` (include/gigaenv.h) used by the test
suite and bench.py; the product is the shared library and the C++ host headers under ``host/``.
There is NO CPU fallback: if the CUDA library is missing, loading fails loudly.
"""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GGL_LIB_PATH") or os.path.join(_HERE, "libgigaenv_cuda.so")   # GGL_LIB_PATH: profiling build
_lib = None


class GigaEnvError(RuntimeError):
    pass


def lib():
    """Loads libgigaenv_cuda.so (raises if it has not been built: run ``python __graft_entry__.py``)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GigaEnvError("libgigaenv_cuda.so not built (%s): no CPU fallback exists" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.ggl_last_error.restype = C.c_char_p
        L.ggl_version.restype = C.c_char_p
        L.ggl_launch_count.restype = C.c_uint64
        L.ggl_gae_workspace_bytes.restype = C.c_size_t
        L.ggl_gae_workspace_bytes.argtypes = [C.c_int64]
        L.ggl_gae_compute.argtypes = ([C.c_void_p] * 4 + [C.c_int64, C.c_int64] + [C.c_float] * 4 +
                                      [C.c_void_p] * 7)
        L.ggl_gae_compute_host.argtypes = ([C.c_void_p] * 4 + [C.c_int64, C.c_int64] + [C.c_float] * 4 +
                                           [C.c_void_p] * 3 + [C.POINTER(C.c_float)])
        if hasattr(L, "ggl_expbuf_gather"):
            L.ggl_expbuf_plan_batches.restype = C.c_int64
            L.ggl_expbuf_plan_batches.argtypes = [C.c_int64, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_int64]
            L.ggl_expbuf_gather.argtypes = [C.POINTER(abi.ExpBufTensors), C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_void_p,
                                            C.c_int64, C.c_int64, C.POINTER(abi.ExpBufTensors), C.c_void_p]
        if hasattr(L, "ggl_obsstat_create"):
            L.ggl_obsstat_create.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_void_p)]
            L.ggl_obsstat_destroy.argtypes = [C.c_void_p]
            L.ggl_obsstat_increment_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]
            L.ggl_obsstat_normalize.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_double, C.c_void_p]
            L.ggl_obsstat_get.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]
            L.ggl_obsstat_set.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        if hasattr(L, "ggl_traj_create"):
            L.ggl_traj_create.argtypes = [C.POINTER(abi.TrajDesc), C.POINTER(C.c_void_p)]
            L.ggl_traj_destroy.argtypes = [C.c_void_p]
            L.ggl_traj_views_get.argtypes = [C.c_void_p, C.POINTER(abi.TrajViews)]
            L.ggl_traj_begin_iteration.argtypes = [C.c_void_p, C.c_void_p]
            L.ggl_traj_push_obs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
            L.ggl_traj_push_result.argtypes = [C.c_void_p] * 7
            L.ggl_traj_counts.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
            L.ggl_traj_player_lengths.argtypes = [C.c_void_p, C.c_void_p]
        if hasattr(L, "ggl_envset_create"):
            L.ggl_envset_create.argtypes = [C.POINTER(abi.EnvSetDesc), C.POINTER(C.c_void_p)]
            L.ggl_envset_destroy.argtypes = [C.c_void_p]
            L.ggl_envset_step_first_half.argtypes = [C.c_void_p, C.c_void_p]
            L.ggl_envset_step_second_half.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
            L.ggl_envset_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
            L.ggl_envset_reset.argtypes = [C.c_void_p, C.c_void_p]
            L.ggl_envset_reset_arena.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
            L.ggl_envset_contact_overflow.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
            L.ggl_envset_sync.argtypes = [C.c_void_p]
            L.ggl_envset_views_get.argtypes = [C.c_void_p, C.POINTER(abi.EnvSetViews)]
            L.ggl_envset_set_state.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.ArenaState)]
            L.ggl_envset_get_state.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.ArenaState)]
            L.ggl_envset_set_reset_state.argtypes = [C.c_void_p, C.c_int, C.POINTER(abi.ArenaState)]
            L.ggl_envset_get_contacts.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
            L.ggl_envset_set_slot_order.argtypes = [C.c_void_p, C.c_void_p]
            L.ggl_envset_set_rng.argtypes = [C.c_void_p, C.c_int, C.c_uint32]
            L.ggl_envset_get_rng.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint32)]
            L.ggl_arena_step.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
            L.ggl_envset_random_actions.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
            L.ggl_envset_step_host.argtypes = [C.c_void_p] * 6
            L.ggl_envset_begin_step_host.argtypes = [C.c_void_p] * 3
            L.ggl_envset_finish_step_host.argtypes = [C.c_void_p] * 5
            L.ggl_action_table.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
            L.ggl_host_random_actions.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise GigaEnvError("gigaenv error %d: %s" % (rc, lib().ggl_last_error().decode()))


def launch_count():
    return int(lib().ggl_launch_count())


# ----------------------------------------------------------------------------------------------- GAE
def gae_compute_host(rews, terminals, val_preds, trunc_val_preds=None, gamma=0.99, lam=0.95, return_std=0.0,
                     clip_range=10.0, out=None):
    """GAE::Compute through HOST buffers (numpy arrays or anything exposing ``.ctypes``), mirroring the reference
    call at Learner.cpp:949-953 where all tensors live on the CPU. Returns (advantages, returns, target_values,
    clip_portion)."""
    rews = np.ascontiguousarray(rews, np.float32)
    terminals = np.ascontiguousarray(terminals, np.int8)
    val_preds = np.ascontiguousarray(val_preds, np.float32)
    L = rews.shape[0]
    if out is None:
        out = tuple(np.empty(L, np.float32) for _ in range(3))
    tv, K = None, 0
    if trunc_val_preds is not None:
        tvv = np.ascontiguousarray(trunc_val_preds, np.float32)
        K = tvv.shape[0]
        tv = tvv.ctypes.data if K > 0 else rews.ctypes.data  # non-NULL marks "defined tensor"
    clip = C.c_float()
    _check(lib().ggl_gae_compute_host(rews.ctypes.data, terminals.ctypes.data, val_preds.ctypes.data, tv, L, K,
                                      gamma, lam, return_std, clip_range,
                                      out[0].ctypes.data, out[1].ctypes.data, out[2].ctypes.data, C.byref(clip)))
    return out[0], out[1], out[2], clip.value


def gae_compute_device(rews, terminals, val_preds, trunc_val_preds=None, gamma=0.99, lam=0.95, return_std=0.0,
                       clip_range=10.0, out=None, workspace=None, stream=None):
    """Device-resident GAE on torch CUDA tensors (zero-copy: raw pointers go straight to the C ABI).
    Returns (advantages, returns, target_values, clip_portion[1], trunc_count[1]) as CUDA tensors."""
    import torch
    L = rews.numel()
    dev = rews.device
    assert rews.dtype == torch.float32 and val_preds.dtype == torch.float32 and terminals.dtype == torch.int8
    if out is None:
        out = tuple(torch.empty(L, dtype=torch.float32, device=dev) for _ in range(3))
    scal = torch.zeros(2, dtype=torch.float32, device=dev)
    need = lib().ggl_gae_workspace_bytes(L)
    if workspace is None or workspace.numel() < need:
        workspace = torch.empty(need, dtype=torch.uint8, device=dev)
    K = 0 if trunc_val_preds is None else trunc_val_preds.numel()
    tv = None if trunc_val_preds is None else (trunc_val_preds.data_ptr() if K > 0 else rews.data_ptr())
    s = torch.cuda.current_stream(dev).cuda_stream if stream is None else stream
    _check(lib().ggl_gae_compute(rews.data_ptr(), terminals.data_ptr(), val_preds.data_ptr(), tv, L, K,
                                 gamma, lam, return_std, clip_range,
                                 out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(),
                                 scal.data_ptr(), scal.data_ptr() + 4, workspace.data_ptr(), s))
    return out[0], out[1], out[2], scal[:1], scal[1:].view(torch.int32), workspace


def action_table():
    t = np.empty((abi.GGL_NUM_ACTIONS, 8), np.float32)
    n = C.c_int()
    _check(lib().ggl_action_table(t.ctypes.data, C.byref(n)))
    return t[:n.value]


# ------------------------------------------------------------------------------------ ExperienceBuffer
EXPBUF_FIELDS = ("states", "actions", "log_probs", "target_values", "action_masks", "advantages")


def expbuf_plan_batches(exp_size, batch_size, overbatching):
    """[(start, size)] of GetAllBatchesShuffled's minibatches over the permuted stream (ExperienceBuffer.cpp:143-160)."""
    L = lib()
    n = L.ggl_expbuf_plan_batches(exp_size, batch_size, int(bool(overbatching)), None, None, 0)
    if n < 0:
        _check(int(n))
    starts = np.zeros(max(n, 1), np.int64); sizes = np.zeros(max(n, 1), np.int64)
    L.ggl_expbuf_plan_batches(exp_size, batch_size, int(bool(overbatching)), starts.ctypes.data, sizes.ctypes.data, n)
    return [(int(starts[i]), int(sizes[i])) for i in range(n)]


def expbuf_gather(data, indices, max_action_index=-1, out=None, stream=None):
    """One fused permuted gather of the experience tensors on the device. data: dict over EXPBUF_FIELDS of CUDA tensors (or
    None); indices: CUDA int64 tensor. Returns the dict of gathered tensors (minibatches are slices of them)."""
    import torch
    L = lib()
    states = data.get("states")
    dev = indices.device
    count = int(indices.numel())
    src, dst = abi.ExpBufTensors(), abi.ExpBufTensors()
    res = {}
    for f in EXPBUF_FIELDS:
        t = data.get(f)
        if t is None:
            res[f] = None
            continue
        assert t.is_cuda and t.is_contiguous()
        o = None if out is None else out.get(f)
        if o is None:
            o = torch.empty((count,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev)
        res[f] = o
        setattr(src, f, t.data_ptr()); setattr(dst, f, o.data_ptr())
    rows = 0 if states is None else int(states.shape[0])
    rows_a = rows if data.get("actions") is None else int(data["actions"].shape[0])
    obs = 1 if states is None else int(states.shape[1])
    na = 1 if data.get("action_masks") is None else int(data["action_masks"].shape[1])
    s = torch.cuda.current_stream(dev).cuda_stream if stream is None else stream
    _check(L.ggl_expbuf_gather(C.byref(src), rows, rows_a, obs, na, indices.data_ptr(), count, max_action_index, C.byref(dst), s))
    return res


# -------------------------------------------------------------------------------------------- EnvSet
class _DevView:
    """Zero-copy torch view of a device buffer owned by the envset (via __cuda_array_interface__)."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class EnvSet:
    """Device-resident RLGC::EnvSet (EnvSet.h:62-110) behind the C ABI. Buffers (`obs`, `action_masks`, `rewards`,
    `terminals`) are torch CUDA tensors aliasing the library's memory: what libtorch would wrap with from_blob."""

    def __init__(self, num_arenas, players_per_team, device=0, **kw):
        import torch
        L = lib()
        self.desc, self._keep = abi.make_desc(num_arenas, players_per_team, device=device, **kw)
        h = C.c_void_p()
        _check(L.ggl_envset_create(C.byref(self.desc), C.byref(h)))
        self.h = h
        v = abi.EnvSetViews()
        _check(L.ggl_envset_views_get(self.h, C.byref(v)))
        self.num_arenas, self.num_players, self.obs_size = v.num_arenas, v.num_players, v.obs_size
        self.num_actions, self.players_per_arena = v.num_actions, v.players_per_arena
        self.num_rewards = self.desc.num_rewards
        self.device = torch.device("cuda", device)
        with torch.cuda.device(self.device):
            t = lambda p, shape, ts: torch.as_tensor(_DevView(p, shape, ts), device=self.device)
            self.obs = t(v.d_obs, (self.num_players, self.obs_size), "<f4")
            self.action_masks = t(v.d_action_masks, (self.num_players, self.num_actions), "|u1")
            self.rewards = t(v.d_rewards, (self.num_players,), "<f4")
            self.terminals = t(v.d_terminals, (self.num_arenas,), "|u1")
            self.slot_car_id = t(v.d_slot_car_id, (self.num_players,), "<i4")
            self.slot_team = t(v.d_slot_team, (self.num_players,), "|u1")
            self.events = t(v.d_events, (self.num_players, 9), "|u1")
            self.last_rewards = t(v.d_last_rewards, (self.num_arenas, max(1, self.num_rewards)), "<f4")

    def _stream(self, stream):
        import torch
        return torch.cuda.current_stream(self.device).cuda_stream if stream is None else stream

    def close(self):
        if self.h:
            lib().ggl_envset_destroy(self.h)
            self.h = None

    def reset(self, stream=None):
        _check(lib().ggl_envset_reset(self.h, self._stream(stream)))

    def contact_overflow(self):
        """contact points dropped since create because an arena ran out of contact rows (0 = none)"""
        n = C.c_uint64()
        _check(lib().ggl_envset_contact_overflow(self.h, C.byref(n)))
        return int(n.value)

    def reset_arena(self, arena=-1, stream=None):
        """EnvSet::ResetArena(arena) whatever the terminal flag; -1 = every arena."""
        _check(lib().ggl_envset_reset_arena(self.h, int(arena), self._stream(stream)))

    def step_first_half(self, stream=None):
        _check(lib().ggl_envset_step_first_half(self.h, self._stream(stream)))

    def step_second_half(self, actions, stream=None):
        _check(lib().ggl_envset_step_second_half(self.h, actions.data_ptr(), self._stream(stream)))

    def step(self, actions, stream=None, reset=True):
        """Learner loop order (Learner.cpp:669-793): Reset(), then both halves fused in one launch."""
        s = self._stream(stream)
        if reset:
            _check(lib().ggl_envset_reset(self.h, s))
        _check(lib().ggl_envset_step(self.h, actions.data_ptr(), s))

    def arena_step(self, ticks, stream=None):
        _check(lib().ggl_arena_step(self.h, ticks, self._stream(stream)))

    def random_actions(self, actions, seed, stream=None):
        _check(lib().ggl_envset_random_actions(self.h, actions.data_ptr(), seed, self._stream(stream)))

    def host_random_actions(self, h_masks, h_actions, seed):
        p = lambda x: x.data_ptr() if hasattr(x, "data_ptr") else x.ctypes.data
        _check(lib().ggl_host_random_actions(p(h_masks), p(h_actions), self.num_players, self.num_actions, seed))

    def sync(self):
        _check(lib().ggl_envset_sync(self.h))

    def step_host(self, h_actions, h_obs=None, h_masks=None, h_rewards=None, h_terminals=None):
        p = lambda x: None if x is None else (x.data_ptr() if hasattr(x, "data_ptr") else x.ctypes.data)
        _check(lib().ggl_envset_step_host(self.h, p(h_actions), p(h_obs), p(h_masks), p(h_rewards), p(h_terminals)))

    def begin_step_host(self, h_obs=None, h_masks=None):
        """Reset + post-reset obs/masks to the host + first half launched (still running on return)."""
        p = lambda x: None if x is None else (x.data_ptr() if hasattr(x, "data_ptr") else x.ctypes.data)
        _check(lib().ggl_envset_begin_step_host(self.h, p(h_obs), p(h_masks)))

    def finish_step_host(self, h_actions, h_rewards=None, h_terminals=None, h_final_obs=None):
        p = lambda x: None if x is None else (x.data_ptr() if hasattr(x, "data_ptr") else x.ctypes.data)
        _check(lib().ggl_envset_finish_step_host(self.h, p(h_actions), p(h_rewards), p(h_terminals), p(h_final_obs)))

    def set_state(self, arena, st):
        _check(lib().ggl_envset_set_state(self.h, arena, C.byref(st)))

    def get_state(self, arena):
        st = abi.ArenaState()
        _check(lib().ggl_envset_get_state(self.h, arena, C.byref(st)))
        return st

    def get_contacts(self, arena, maxn=64):
        out = np.zeros((maxn, 20), np.float32)
        n = lib().ggl_envset_get_contacts(self.h, arena, out.ctypes.data, maxn)
        if n < 0:
            _check(n)
        return out[:n]

    def set_reset_state(self, arena, st):
        _check(lib().ggl_envset_set_reset_state(self.h, arena, C.byref(st)))

    def set_slot_order(self, car_ids):
        a = np.ascontiguousarray(car_ids, np.int32)
        _check(lib().ggl_envset_set_slot_order(self.h, a.ctypes.data))

    def set_rng(self, arena, state):
        """state of the arena's std::minstd_rand0 stream (1..2147483646); arena -1 = every arena"""
        _check(lib().ggl_envset_set_rng(self.h, int(arena), int(state)))

    def get_rng(self, arena):
        s = C.c_uint32()
        _check(lib().ggl_envset_get_rng(self.h, int(arena), C.byref(s)))
        return int(s.value)


class TrajectoryStore:
    """Device-side `trajectories[]` + `combinedTraj` of Learner::Start (Learner.cpp:504-547, 701-712, 823-863) over ggl_traj_*.
    `tensors()` returns torch views of the first `length` rows of the combined stream (zero-copy)."""

    def __init__(self, num_players, players_per_arena, obs_size, num_actions, max_episode_length, capacity_timesteps,
                 capacity_truncated=None, device=0, capacity_player_rows=0):
        import torch
        d = abi.TrajDesc()
        d.device, d.num_players, d.players_per_arena, d.obs_size, d.num_actions = device, num_players, players_per_arena, obs_size, num_actions
        d.max_episode_length, d.capacity_timesteps = max_episode_length, capacity_timesteps
        d.capacity_truncated = capacity_truncated if capacity_truncated is not None else capacity_timesteps
        d.capacity_player_rows = capacity_player_rows      # 0 = worst case (players x max_episode_length); else the steps in flight
        self.desc = d
        self.device = torch.device("cuda", device)
        h = C.c_void_p()
        _check(lib().ggl_traj_create(C.byref(d), C.byref(h)))
        self.h = h
        v = abi.TrajViews()
        _check(lib().ggl_traj_views_get(self.h, C.byref(v)))
        self.views = v

    def _stream(self, stream):
        import torch
        return C.c_void_p(stream if stream is not None else torch.cuda.current_stream(self.device).cuda_stream)

    def begin_iteration(self, stream=None):
        _check(lib().ggl_traj_begin_iteration(self.h, self._stream(stream)))

    def push_obs(self, obs, masks, stream=None):
        _check(lib().ggl_traj_push_obs(self.h, obs.data_ptr(), masks.data_ptr(), self._stream(stream)))

    def push_result(self, actions, log_probs, rewards, terminals, obs_final, stream=None):
        _check(lib().ggl_traj_push_result(self.h, actions.data_ptr(), log_probs.data_ptr(), rewards.data_ptr(), terminals.data_ptr(),
                                          obs_final.data_ptr(), self._stream(stream)))

    def counts(self, stream=None):
        a, b = C.c_int64(), C.c_int64()
        _check(lib().ggl_traj_counts(self.h, self._stream(stream), C.byref(a), C.byref(b)))
        return a.value, b.value

    def player_lengths(self):
        out = np.zeros(self.desc.num_players, np.int32)
        _check(lib().ggl_traj_player_lengths(self.h, out.ctypes.data))
        return out

    def tensors(self, stream=None):
        import torch
        L, K = self.counts(stream)
        d, v = self.desc, self.views
        t = lambda p, shape, ts: (torch.as_tensor(_DevView(p, shape, ts), device=self.device) if shape[0] > 0 else
                                  torch.empty(shape, dtype={"<f4": torch.float32, "|u1": torch.uint8, "<i4": torch.int32, "|i1": torch.int8}[ts], device=self.device))
        return {"states": t(v.states, (L, d.obs_size), "<f4"), "action_masks": t(v.action_masks, (L, d.num_actions), "|u1"),
                "actions": t(v.actions, (L,), "<i4"), "log_probs": t(v.log_probs, (L,), "<f4"), "rewards": t(v.rewards, (L,), "<f4"),
                "terminals": t(v.terminals, (L,), "|i1"), "next_states": t(v.next_states, (K, d.obs_size), "<f4")}

    def close(self):
        if self.h:
            lib().ggl_traj_destroy(self.h)
            self.h = None


class ObsStat:
    """Device-side GGL::BatchedWelfordStat (WelfordStat.h:69-243) as the Learner uses it (Learner.cpp:680-694)."""

    def __init__(self, width, device=0):
        import torch
        self.width, self.device = width, torch.device("cuda", device)
        h = C.c_void_p()
        _check(lib().ggl_obsstat_create(device, width, C.byref(h)))
        self.h = h

    def _stream(self, stream):
        import torch
        return C.c_void_p(stream if stream is not None else torch.cuda.current_stream(self.device).cuda_stream)

    def increment_rows(self, obs, row_indices, stream=None):
        _check(lib().ggl_obsstat_increment_rows(self.h, obs.data_ptr(), row_indices.data_ptr(), int(row_indices.numel()), self._stream(stream)))

    def normalize_in_place(self, obs, mean_clamp, min_std, stream=None):
        _check(lib().ggl_obsstat_normalize(self.h, obs.data_ptr(), int(obs.shape[0]), float(mean_clamp), float(min_std), self._stream(stream)))

    def state(self):
        m, v, c = np.zeros(self.width), np.zeros(self.width), C.c_int64()
        _check(lib().ggl_obsstat_get(self.h, m.ctypes.data, v.ctypes.data, C.byref(c)))
        return m, v, c.value

    def set_state(self, means, variances, count):
        m, v = np.ascontiguousarray(means, np.float64), np.ascontiguousarray(variances, np.float64)
        _check(lib().ggl_obsstat_set(self.h, m.ctypes.data, v.ctypes.data, int(count)))

    def close(self):
        if self.h:
            lib().ggl_obsstat_destroy(self.h)
            self.h = None
