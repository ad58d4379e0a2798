"""CPU checks of the trajectory-assembly oracle (oracle/traj_oracle.py, PARITY UNPINNED: see its header) on cases small enough
to follow by hand from Learner.cpp:823-863, plus the invariants any correct assembly has."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import traj_oracle as to        # noqa: E402
import traj_cases as tc         # noqa: E402


def test_hand_checked_order():
    # 2 arenas x 1 player, max_episode_length 3. Arena 1 ends (NORMAL) at step 1; player 0 is force-truncated at step 2.
    o = to.TrajOracle(2, 1, 3)
    o.begin_iteration()
    terms = [[0, 0], [0, 1], [0, 0], [0, 0]]
    for t in range(4):
        obs = np.array([[10 * t + 0], [10 * t + 1]], np.float32)
        o.push_obs(obs, np.ones((2, 1), np.uint8))
        o.push_result([t, t], [0.5, 0.5], [1.0, 2.0], terms[t], obs + 0.5)
    d = o.tensors(1, 1)
    # stream: player 1 steps 0-1 (appended at step 1), then player 0 steps 0-2 (forced truncation at step 2: Length() == 3)
    assert d["states"][:, 0].tolist() == [1, 11, 0, 10, 20]
    assert d["terminals"].tolist() == [0, 1, 0, 0, 2]
    assert d["next_states"][:, 0].tolist() == [20.5]                 # final obs of the truncated segment only
    assert o.player_lengths().tolist() == [1, 2]                     # unfinished buffers stay with their player
    o.begin_iteration()                                              # :631 clears the combined stream only
    assert o.length() == 0 and o.player_lengths().tolist() == [1, 2]


def test_invariants():
    for ci, case in enumerate(tc.CASES):
        arenas, ppa, obs, na, max_len, steps, _, _ = case
        o = to.TrajOracle(arenas * ppa, ppa, max_len)
        o.begin_iteration()
        for step in tc.rollout(case, seed=ci):
            o.push_obs(step[0], step[1]); o.push_result(*step[2:])
        d = o.tensors(obs, na)
        L = o.length()
        assert L + int(o.player_lengths().sum()) == arenas * ppa * steps          # every timestep is in exactly one place
        assert (d["terminals"] == 2).sum() == len(d["next_states"])
        ends = np.flatnonzero(d["terminals"])
        assert L == 0 or ends[-1] == L - 1                                       # the stream is whole segments
        start = 0
        for e in ends:                                                            # a segment = one player's consecutive steps
            seg = d["states"][start:e + 1]
            assert len(set(seg[:, 1].tolist())) == 1 and np.all(np.diff(seg[:, 0]) == 1) and len(seg) <= max(max_len, 1)
            start = e + 1
