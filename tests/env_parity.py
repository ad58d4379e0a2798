"""TEST INFRASTRUCTURE: parity drivers comparing the CUDA EnvSet (through the C ABI) with the compiled reference
(oracle/_ref/libref_oracle.so). Used by tests/test_env_gpu.py and __graft_entry__.smoke()."""
import os

import numpy as np

import oracle_ref

if not oracle_ref.available():
    raise ImportError("oracle/_ref/libref_oracle.so not built")

OBS_RTOL = 1e-4      # BASELINE.json north_star: 1e-4 relative on obs / reward


def close_rel(x, ref, rtol=OBS_RTOL, atol=1e-4):
    return np.abs(x - ref) <= rtol * np.abs(ref) + atol


def random_valid_actions(rng, masks):
    cnt = masks.sum(1)
    k = (rng.random(len(masks)) * cnt).astype(np.int64)
    k = np.minimum(k, np.maximum(cnt - 1, 0))
    cs = np.cumsum(masks, 1)
    return (cs > k[:, None]).argmax(1).astype(np.int32)


def _compare_steps(ref, dev, steps, rng, res, ppa, tick_skip=8, verbose=False, track_touches=True, resync=False):
    """Steps both sides with the same mask-valid random actions and accumulates the worst-case deviations in `res`.
    resync: before every step the device gets the reference's physics state of every arena (ref.get_state -> dev.set_state),
    so that each step's deviation is that step's own error and not the chaotic growth of an earlier one; the env-layer
    bookkeeping (previous state, event tracker, timers, terminal flags) is never copied and must stay equal on its own."""
    host_mode = hasattr(dev, "step_np")
    acts_d = None
    if not host_mode:
        import torch
        acts_d = torch.zeros(dev.num_players, dtype=torch.int32, device=dev.device)
    bad_arenas = res.setdefault("_bad", set())
    obs_r, masks_r, _, _ = ref.outputs()
    for s in range(steps):
        acts = random_valid_actions(rng, masks_r)
        snaps = None
        if resync:
            snaps = [ref.get_state(a) for a in range(ref.num_arenas)]
            for a in range(ref.num_arenas):
                dev.set_state(a, snaps[a])
        ref.step(acts)
        if host_mode:
            dev.step_np(acts)
        else:
            import torch
            acts_d.copy_(torch.from_numpy(acts))
            dev.step(acts_d)
        dev.sync()
        obs_r, masks_r, rew_r, term_r = ref.outputs()
        obs_d = dev.obs.cpu().numpy(); masks_d = dev.action_masks.cpu().numpy()
        rew_d = dev.rewards.cpu().numpy(); term_d = dev.terminals.cpu().numpy()
        ev_r = ref.events(); ev_d = dev.events.cpu().numpy()
        ok = close_rel(obs_d, obs_r)
        # an arena that left the tolerance stays out of the later comparisons (chaotic growth after the first miss)
        live = np.ones(ref.num_arenas, bool)
        if bad_arenas and not resync:
            live[list(bad_arenas)] = False
        if snaps is not None:
            # a demolished car that respawns during this step gets its spawn slot from the reference's global random engine
            # (Car::Respawn -> Math::RandInt, Car.cpp:43-55), the device draws from the arena's own stream: position, obs and
            # masks of that arena are not comparable in this one step (the next step starts from the reference's state again)
            for a in range(ref.num_arenas):
                st = snaps[a]
                if any(st.cars[c].is_demoed and st.cars[c].demo_respawn_timer <= (tick_skip + 1) / 120.0 for c in range(st.num_cars)):
                    live[a] = False
                    res["respawn_arena_steps"] += 1
        live_p = np.repeat(live, ppa)
        # arenas whose obs left the tolerance in THIS step: the physics diverged at a knife edge (or through a documented
        # ordering difference), so flags derived from it (turtle bit of the mask, touch events ...) may differ too. Integer
        # outputs are required to be exact on every other arena-step; on these they are only counted.
        ok_arena = ok.reshape(ref.num_arenas, -1).all(1)
        clean = live & (ok_arena | (not resync))
        clean_p = np.repeat(clean, ppa)
        res["obs"] = max(res["obs"], float(np.max(np.abs(obs_d - obs_r)[live_p]))) if live_p.any() else res["obs"]
        res["obs_bad"] += int((~ok)[live_p].sum())
        res["rew"] = max(res["rew"], float(np.max(np.abs(rew_d - rew_r)[clean_p]))) if clean_p.any() else res["rew"]
        res["rew_abs_max"] = max(res["rew_abs_max"], float(np.abs(rew_r).max()))
        res["rew_nonzero"] += int((rew_r != 0).sum())
        res["mask_mismatch"] += int((masks_d != masks_r)[clean_p].sum())
        res["term_mismatch"] += int((term_d != term_r)[clean].sum())
        res["event_mismatch"] += int((ev_d != ev_r)[clean_p].sum())
        res["int_mismatch_on_bad_steps"] += int((masks_d != masks_r)[live_p & ~clean_p].sum() + (ev_d != ev_r)[live_p & ~clean_p].sum()
                                                + (term_d != term_r)[live & ~clean].sum())
        res["terminals_seen"] += int((term_r != 0).sum())
        res["goals_seen"] += int((term_r == 1).sum())
        res["events"] += ev_r[live_p].sum(0).astype(np.int64)
        res["masks_with_jump_reset"] += int(masks_r.sum())
        if track_touches:
            for a in range(ref.num_arenas):
                st = ref.get_state(a)
                for c in range(st.num_cars):
                    car = st.cars[c]
                    if car.hit_valid and car.hit_tick + tick_skip >= st.tick_count:
                        res["touches"] += 1
        res["steps"] += 1
        dump = os.environ.get("GGL_PARITY_DUMP")
        if dump and snaps is not None and ((masks_d != masks_r).any() or (ev_d != ev_r).any() or (term_d != term_r).any()):
            # debug aid: the pre-step reference state of the first arena with an integer mismatch, for tools/replay_arena_step.py
            import pickle
            pl = np.argwhere((masks_d != masks_r).any(1) | (ev_d != ev_r).any(1))[:, 0]
            a = int(pl[0] // ppa) if len(pl) else int(np.argwhere(term_d != term_r)[0, 0])
            ids, _ = ref.slot_info()
            with open(dump, "wb") as f:
                pickle.dump({"state": bytes(snaps[a]), "acts": acts[a * ppa:(a + 1) * ppa], "ids": ids[a * ppa:(a + 1) * ppa], "ppa": ppa,
                             "step": s, "arena": a}, f)
        res["arena_steps"] += ref.num_arenas
        if not ok[live_p].all():
            newbad = set((np.argwhere(~ok & live_p[:, None])[:, 0] // ppa).tolist())
            res["bad_arena_steps"] += len(newbad)
            if verbose:
                print("step", s, "obs bad arenas", sorted(newbad - bad_arenas), "first", np.argwhere(~ok)[:3].tolist())
            bad_arenas.update(newbad)
        if verbose and ((masks_d != masks_r).any() or (term_d != term_r).any() or (ev_d != ev_r).any()):
            print("step", s, "masks", np.argwhere(masks_d != masks_r)[:5].tolist(), "term", np.argwhere(term_d != term_r)[:5].tolist(),
                  "events", np.argwhere(ev_d != ev_r)[:5].tolist())
        res["bad_arenas"] = len(bad_arenas)
    return res


def _new_result():
    return {"obs": 0.0, "rew": 0.0, "mask_mismatch": 0, "term_mismatch": 0, "event_mismatch": 0, "obs_bad": 0, "steps": 0,
            "terminals_seen": 0, "goals_seen": 0, "touches": 0, "bad_arenas": 0, "events": np.zeros(9, np.int64),
            "masks_with_jump_reset": 0, "arena_steps": 0, "bad_arena_steps": 0, "respawn_arena_steps": 0, "int_mismatch_on_bad_steps": 0, "rew_abs_max": 0.0, "rew_nonzero": 0}


def run_env_steps(ggl, num_arenas=8, ppt=2, steps=40, seed=0, obs_kind=1, verbose=False, env_cls=None, **kw):
    """Steps both sides from the kickoff with the same mask-valid random actions. Returns a dict of worst-case deviations."""
    L = oracle_ref.lib()
    L.ref_set_threads(1)
    L.ref_seed(99 + seed)
    ref = oracle_ref.RefEnvSet(num_arenas, ppt, obs_kind=obs_kind, state_setter=oracle_ref.abi.SETTER_INJECTED, **kw)
    dev = (env_cls or ggl.EnvSet)(num_arenas, ppt, obs_kind=obs_kind, state_setter=oracle_ref.abi.SETTER_INJECTED, seed=seed, **kw)
    ids, teams = ref.slot_info()
    dev.set_slot_order(ids)
    for a in range(num_arenas):
        st = ref.get_state(a)
        dev.set_state(a, st)
        ref.set_reset_state(a, st)
        dev.set_reset_state(a, st)
    rng = np.random.default_rng(seed)
    res = _compare_steps(ref, dev, steps, rng, _new_result(), 2 * ppt, verbose=verbose, track_touches=num_arenas <= 64)
    res.pop("_bad", None)
    ref.close(); dev.close()
    return res


def run_steady_state(ggl, num_arenas=16, ppt=2, preroll=160, steps=40, seed=0, obs_kind=1, verbose=False, env_cls=None,
                     scripts=None, threads=1, resync=True, **kw):
    """Parity in steady-state play (the kickoff runs end before any car reaches the ball or another car).

    The REFERENCE is pre-rolled `preroll` env steps from the kickoff with random mask-valid actions (kickoff resets as they
    come), then every arena's full physics state is read out (`ref.get_state`), optionally overwritten by a scripted state
    (`scripts[a](st)`: goal shots, bumps, demolitions ...), installed on BOTH sides as that arena's injected reset state and
    both sides run EnvSet::ResetArena on it, so that the env-layer bookkeeping (prev state, event tracker, no-touch timers)
    starts equal. Then `steps` env steps are compared like run_env_steps; an arena that hits a terminal falls back to the
    same injected state on both sides."""
    L = oracle_ref.lib()
    L.ref_set_threads(threads)
    L.ref_seed(199 + seed)
    ref = oracle_ref.RefEnvSet(num_arenas, ppt, obs_kind=obs_kind, state_setter=oracle_ref.abi.SETTER_INJECTED, **kw)
    rng = np.random.default_rng(1000 + seed)
    masks_r = ref.outputs()[1]
    for s in range(preroll):
        ref.step(random_valid_actions(rng, masks_r))
        masks_r = ref.outputs()[1]
    L.ref_set_threads(1)
    dev = (env_cls or ggl.EnvSet)(num_arenas, ppt, obs_kind=obs_kind, state_setter=oracle_ref.abi.SETTER_INJECTED, seed=seed, **kw)
    ids, teams = ref.slot_info()
    dev.set_slot_order(ids)
    # The game event tracker keeps the ball update count of its last Update() across resets (GameEventTracker.cpp:48-158;
    # ResetPersistentInfo does not clear it) and is not part of any state struct. Both sides have played before the hand-over
    # (the device two idle steps) and the injected ball counter is 0, as after Ball::SetState in a real StateSetter: both
    # trackers then see the count go DOWN in the first compared step, reset themselves and continue in step from there.
    idle = np.full(num_arenas * 2 * ppt, 8, np.int32)
    for _ in range(2):
        if hasattr(dev, "step_np"):
            dev.step_np(idle)
        else:
            import torch
            dev.step(torch.from_numpy(idle).to(dev.device))
    for a in range(num_arenas):
        st = ref.get_state(a)
        if scripts and a in scripts:
            scripts[a](st)
        st.ball.update_counter = 0
        ref.set_reset_state(a, st)
        dev.set_reset_state(a, st)
        ref.reset_arena(a)
    dev.reset_arena(-1)
    dev.sync()
    res = _new_result()
    obs_r, masks_r, _, _ = ref.outputs()
    assert close_rel(dev.obs.cpu().numpy(), obs_r).all(), "obs differ right after the hand-over reset"
    assert np.array_equal(dev.action_masks.cpu().numpy(), masks_r), "masks differ right after the hand-over reset"
    _compare_steps(ref, dev, steps, rng, res, 2 * ppt, verbose=verbose, track_touches=num_arenas <= 64, resync=resync)
    res.pop("_bad", None)
    res["contact_overflow"] = dev.contact_overflow() if hasattr(dev, "contact_overflow") else -1
    ref.close(); dev.close()
    return res


def smoke_check(ggl):
    r = run_env_steps(ggl, num_arenas=4, ppt=2, steps=6, seed=3)
    assert r["mask_mismatch"] == 0 and r["term_mismatch"] == 0, r
    assert r["obs_bad"] == 0, r
    return r
