"""TEST INFRASTRUCTURE: lock-steps the random plugins of SURVEY row f4 (FuzzedKickoffState, RandomState, CombinedState, the
kickoff shuffle itself, DefaultObsPadded's per-row shuffles) against the compiled reference (oracle/_ref).

The reference draws from one std::default_random_engine per thread (RocketSim Math.cpp:59-64); the device keeps one
std::minstd_rand0 state per arena. With one arena, a one-worker pool and the same engine state installed on both sides before an
operation, both must draw the same numbers in the same order: states, obs rows and the engine state afterwards are compared."""
import numpy as np

import oracle_ref
from oracle_ref import abi

M = 2147483647


def _phys(st, n):
    out = [np.array(st.ball.pos[:]), np.array(st.ball.vel[:]), np.array(st.ball.ang_vel[:])]
    for i in range(n):
        c = st.cars[i]
        out += [np.array(c.pos[:]), np.array(c.rot[:]), np.array(c.vel[:]), np.array(c.ang_vel[:]),
                np.array([c.boost, float(c.is_on_ground), float(c.has_jumped), float(c.is_demoed)])]
    out.append(np.array([float(st.pads[i].is_active) for i in range(abi.GGL_NUM_PADS)]))
    return np.concatenate(out).astype(np.float64)


def make_pair(env_cls, ppt, **kw):
    L = oracle_ref.lib()
    L.ref_set_threads(1)
    ref = oracle_ref.RefEnvSet(1, ppt, **kw)
    dev = env_cls(1, ppt, **kw)
    ids, _ = ref.slot_info()
    dev.set_slot_order(ids)
    return ref, dev


def _sync_engines(dev, state):
    oracle_ref.lib().ref_seed(int(state))
    dev.set_rng(0, int(state))


def _next(state):
    return (16807 * int(state)) % M


def _engine_after(start):
    """next draw of whichever reference engine moved (calling thread = 0, pool worker = 1); that of `start` when neither did"""
    L = oracle_ref.lib()
    moved = [s for s in (L.ref_rng_peek(0), L.ref_rng_peek(1)) if s != _next(start)]
    assert len(moved) <= 1, "both reference engines drew: the comparison needs one arena on one thread"
    return moved[0] if moved else _next(start)


def _dev_obs(dev):
    o = dev.obs
    return o.cpu().numpy() if hasattr(o, "cpu") else np.asarray(o)


def run_resets(env_cls, ppt, trials=40, seed=0, atol=0.0, **kw):
    """trials x (install one engine state on both sides, ResetArena(0), compare). Returns the worst absolute differences."""
    ref, dev = make_pair(env_cls, ppt, **kw)
    rng = np.random.default_rng(seed)
    n = 2 * ppt
    worst_state = worst_obs = 0.0
    kinds = set()
    for _ in range(trials):
        s0 = int(rng.integers(1, M - 1))
        _sync_engines(dev, s0)
        ref.reset_arena(0)
        dev.reset_arena(0)
        if hasattr(dev, "sync"):
            dev.sync()
        a, b = _phys(ref.get_state(0), n), _phys(dev.get_state(0), n)
        worst_state = max(worst_state, float(np.max(np.abs(a - b))))
        ro = ref.outputs()[0]
        do = _dev_obs(dev)
        worst_obs = max(worst_obs, float(np.max(np.abs(ro - do))))
        assert _engine_after(s0) == _next(dev.get_rng(0)), "engine state after the reset differs: a draw is missing or in excess"
        kinds.add(round(float(a[2]), 2))           # ball z: 93.15 for the kickoffs, random for RandomState
    ref.close()
    return {"state": worst_state, "obs": worst_obs, "ball_z_values": len(kinds)}


def run_padded_steps(env_cls, ppt, max_players, steps=30, seed=0, **kw):
    """DefaultObsPadded through env steps: same engine state before every step, obs rows (which teammate / opponent sits in which
    shuffled slot, zero rows where nobody is) compared after it. Physics parity is the other tests' business: the reference's
    post-step state is handed to the device before each step."""
    kw = dict(kw, obs_kind=abi.OBS_DEFAULT_PADDED, obs_max_players=max_players, state_setter=abi.SETTER_INJECTED)
    ref, dev = make_pair(env_cls, ppt, **kw)
    rng = np.random.default_rng(seed)
    worst = 0.0
    perms = set()
    for _ in range(steps):
        st = ref.get_state(0)
        dev.set_state(0, st)
        masks = ref.outputs()[1]
        acts = np.array([rng.choice(np.flatnonzero(m)) for m in masks], np.int32)
        s0 = int(rng.integers(1, M - 1))
        _sync_engines(dev, s0)
        ref.step(acts)
        if hasattr(dev, "step_np"):
            dev.step_np(acts)
        else:
            import torch
            dev.step(torch.from_numpy(acts).to("cuda:%d" % dev.desc.device))
            dev.sync()
        ro, do = ref.outputs()[0], _dev_obs(dev)
        # which slots are occupied must agree exactly (a row is all zeros or a player)
        occ_r = np.abs(ro[:, 70:].reshape(ro.shape[0], -1, 19)).sum(-1) > 0
        occ_d = np.abs(do[:, 70:].reshape(do.shape[0], -1, 19)).sum(-1) > 0
        assert np.array_equal(occ_r, occ_d), "DefaultObsPadded: a player sits in another shuffled slot than in the reference"
        perms.add(occ_r.tobytes())
        worst = max(worst, float(np.max(np.abs(ro - do))))
        assert _engine_after(s0) == _next(dev.get_rng(0))
    ref.close()
    return {"obs": worst, "distinct_layouts": len(perms)}
