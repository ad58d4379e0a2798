"""GPU parity of the device trajectory store (ggl_traj_*) against the oracle restatement of Learner.cpp:504-547, 701-712,
823-863: bit-exact (index, byte and copy work). The oracle itself is PARITY UNPINNED (oracle/traj_oracle.py header)."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import traj_oracle as to        # noqa: E402
import traj_cases as tc         # noqa: E402


def _dev(x):
    import torch
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


def _compare(store, oracle, obs, na):
    got = store.tensors()
    ref = oracle.tensors(obs, na)
    for k, v in ref.items():
        assert np.array_equal(got[k].cpu().numpy(), v), k
    assert np.array_equal(store.player_lengths(), oracle.player_lengths())


@pytest.mark.parametrize("pool", ["worst_case", "in_flight"])
@pytest.mark.parametrize("ci", range(len(tc.CASES)))
def test_store_matches_oracle(ggl, ci, pool):
    """pool = in_flight: the per-player pages come from a pool sized for the steps in flight (players x the longest episode the
    run can reach, page-rounded) instead of players x max_episode_length; pages are recycled as buffers flush."""
    case = tc.CASES[ci]
    arenas, ppa, obs, na, max_len, steps, _, _ = case
    n = arenas * ppa
    rows = 0 if pool == "worst_case" else n * ((min(max_len, steps) + 15) // 16 * 16)
    store = ggl.TrajectoryStore(n, ppa, obs, na, max_len, capacity_timesteps=n * steps + 16, capacity_player_rows=rows)
    oracle = to.TrajOracle(n, ppa, max_len)
    store.begin_iteration(); oracle.begin_iteration()
    for t, step in enumerate(tc.rollout(case, seed=ci)):
        if t == steps // 2:                         # a new collection iteration in the middle: unfinished buffers carry over
            _compare(store, oracle, obs, na)
            store.begin_iteration(); oracle.begin_iteration()
        oracle.push_obs(step[0], step[1]); oracle.push_result(*step[2:])
        store.push_obs(_dev(step[0]), _dev(step[1]))
        store.push_result(*[_dev(x) for x in step[2:]])
        assert store.counts()[0] == oracle.length()        # the stop rule of :665 reads this after every step
    _compare(store, oracle, obs, na)
    store.close()


def test_capacity_is_enforced_loudly(ggl):
    store = ggl.TrajectoryStore(8, 2, 5, 3, 2, capacity_timesteps=10)
    store.begin_iteration()
    step = next(tc.rollout((4, 2, 5, 3, 2, 1, 0.0, 0.0)))
    for _ in range(2):                              # 8 players flush 2 steps each = 16 > 10 rows
        store.push_obs(_dev(step[0]), _dev(step[1])); store.push_result(*[_dev(x) for x in step[2:]])
    with pytest.raises(ggl.GigaEnvError):
        store.counts()
    store.close()


def test_player_pool_exhaustion_is_reported(ggl):
    """A pool of one 16-step page per player and episodes that run longer: the 17th step finds no page, ggl_traj_counts fails
    with the pool named (nothing is written out of bounds, nothing is dropped silently)."""
    store = ggl.TrajectoryStore(8, 2, 5, 3, 1000, capacity_timesteps=4096, capacity_player_rows=8 * 16)
    store.begin_iteration()
    case = (4, 2, 5, 3, 1000, 20, 0.0, 0.0)
    for t, step in enumerate(tc.rollout(case)):
        store.push_obs(_dev(step[0]), _dev(step[1])); store.push_result(*[_dev(x) for x in step[2:]])
        if t < 16:
            assert store.counts()[0] == 0
    with pytest.raises(ggl.GigaEnvError, match="pages"):
        store.counts()
    store.close()


def test_pages_are_recycled_over_many_episodes(ggl):
    """2000 steps of 64 players with episodes of at most 24 steps through a pool of 2 pages per player: every page is handed out
    and returned dozens of times; the stream still matches the oracle bit for bit."""
    case = (16, 4, 12, 6, 24, 2000, 0.04, 0.04)
    n = 64
    store = ggl.TrajectoryStore(n, 4, 12, 6, 24, capacity_timesteps=n * 250, capacity_player_rows=n * 32)
    oracle = to.TrajOracle(n, 4, 24)
    store.begin_iteration(); oracle.begin_iteration()
    for t, step in enumerate(tc.rollout(case, seed=3)):
        if t % 200 == 199:
            _compare(store, oracle, 12, 6)
            store.begin_iteration(); oracle.begin_iteration()
        oracle.push_obs(step[0], step[1]); oracle.push_result(*step[2:])
        store.push_obs(_dev(step[0]), _dev(step[1])); store.push_result(*[_dev(x) for x in step[2:]])
    _compare(store, oracle, 12, 6)
    store.close()


def test_with_the_envset(ggl):
    """The store fed by the real envset buffers, the way the Learner loop would: Reset, push obs, step, push result."""
    import torch
    env = ggl.EnvSet(64, 2, obs_kind=ggl.abi.OBS_ADVANCED, seed=5)
    n, obs, na = env.num_players, env.obs_size, env.num_actions
    store = ggl.TrajectoryStore(n, 4, obs, na, 20, capacity_timesteps=n * 64)
    oracle = to.TrajOracle(n, 4, 20)
    store.begin_iteration(); oracle.begin_iteration()
    acts = torch.zeros(n, dtype=torch.int32, device="cuda")
    logp = torch.zeros(n, device="cuda")
    for i in range(45):
        env.reset(); env.sync()
        store.push_obs(env.obs, env.action_masks)
        oracle.push_obs(env.obs.cpu().numpy(), env.action_masks.cpu().numpy())
        env.random_actions(acts, seed=900 + i)
        env.step_first_half(); env.step_second_half(acts); env.sync()
        store.push_result(acts, logp, env.rewards, env.terminals, env.obs)
        oracle.push_result(acts.cpu().numpy(), logp.cpu().numpy(), env.rewards.cpu().numpy(), env.terminals.cpu().numpy(), env.obs.cpu().numpy())
    _compare(store, oracle, obs, na)
    assert store.counts()[0] == oracle.length() >= n * 20      # forced truncations at steps 20 and 40 (plus any real terminal)
    store.close(); env.close()
