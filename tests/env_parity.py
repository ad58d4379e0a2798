"""TEST INFRASTRUCTURE: parity drivers comparing the CUDA EnvSet (through the C ABI) with the compiled reference
(oracle/_ref/libref_oracle.so). Used by tests/test_env_gpu.py and __graft_entry__.smoke()."""
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


def run_env_steps(ggl, num_arenas=8, ppt=2, steps=40, seed=0, obs_kind=1, verbose=False, env_cls=None, **kw):
    """Steps both sides with the same mask-valid random actions. Returns a dict of worst-case deviations."""
    import torch
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
    res = {"obs": 0.0, "rew": 0.0, "mask_mismatch": 0, "term_mismatch": 0, "event_mismatch": 0, "obs_bad": 0, "steps": 0,
           "terminals_seen": 0, "touches": 0, "bad_arenas": 0}
    bad_arenas = set()
    ppa = 2 * ppt
    host_mode = hasattr(dev, "step_np")
    acts_d = None if host_mode else torch.zeros(dev.num_players, dtype=torch.int32, device=dev.device)
    # first step: masks from the reference reset
    obs_r, masks_r, _, _ = ref.outputs()
    for s in range(steps):
        acts = random_valid_actions(rng, masks_r)
        ref.step(acts)
        if host_mode:
            dev.step_np(acts)
        else:
            acts_d.copy_(torch.from_numpy(acts))
            dev.step(acts_d)
        dev.sync()
        obs_r, masks_r, rew_r, term_r = ref.outputs()
        obs_d = dev.obs.cpu().numpy(); masks_d = dev.action_masks.cpu().numpy()
        rew_d = dev.rewards.cpu().numpy(); term_d = dev.terminals.cpu().numpy()
        ev_r = ref.events(); ev_d = dev.events.cpu().numpy()
        ok = close_rel(obs_d, obs_r)
        res["obs"] = max(res["obs"], float(np.max(np.abs(obs_d - obs_r))))
        res["obs_bad"] += int((~ok).sum())
        if not ok.all():
            bad_arenas.update((np.argwhere(~ok)[:, 0] // ppa).tolist())
            res["bad_arenas"] = len(bad_arenas)
        res["rew"] = max(res["rew"], float(np.max(np.abs(rew_d - rew_r))))
        res["mask_mismatch"] += int((masks_d != masks_r).sum())
        res["term_mismatch"] += int((term_d != term_r).sum())
        res["event_mismatch"] += int((ev_d != ev_r).sum())
        res["terminals_seen"] += int((term_r != 0).sum())
        res["steps"] += 1
        if verbose and ((~ok).any() or (masks_d != masks_r).any() or (term_d != term_r).any()):
            bad = np.argwhere(~ok)
            print("step", s, "obs bad", bad[:5], "masks", np.argwhere(masks_d != masks_r)[:5], "term", term_d, term_r)
    ref.close(); dev.close()
    return res


def smoke_check(ggl):
    r = run_env_steps(ggl, num_arenas=4, ppt=2, steps=6, seed=3)
    assert r["mask_mismatch"] == 0 and r["term_mismatch"] == 0, r
    assert r["obs_bad"] == 0, r
    return r
