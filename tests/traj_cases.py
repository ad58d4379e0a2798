"""Shared trajectory-assembly test material: a seeded synthetic rollout (obs rows carry (step, player) so order errors show)."""
import numpy as np

# (arenas, players_per_arena, obs, num_actions, max_episode_length, steps, p_normal, p_truncated)
CASES = [
    (3, 2, 5, 3, 4, 12, 0.10, 0.10),       # tiny: forced truncation every 4 steps, odd mask rows
    (16, 4, 167, 90, 25, 60, 0.03, 0.03),  # 2v2 AdvancedObs shape
    (8, 6, 225, 90, 1000, 40, 0.05, 0.05), # 3v3, no forced truncation inside the run
    (64, 2, 89, 90, 7, 30, 0.0, 0.0),      # only forced truncations: every buffer flushes at 7
    (40, 4, 33, 8, 9, 50, 0.2, 0.0),       # frequent NORMAL terminals
    (750, 4, 9, 6, 5, 14, 0.1, 0.1),       # 3000 players: the ordered scan runs over several 1024-player chunks
]


def rollout(case, seed=0):
    arenas, ppa, obs, na, max_len, steps, pn, pt = case
    n = arenas * ppa
    rng = np.random.default_rng(seed)
    for t in range(steps):
        obs_pre = rng.standard_normal((n, obs)).astype(np.float32)
        obs_pre[:, 0] = t; obs_pre[:, 1] = np.arange(n)
        masks = (rng.random((n, na)) < 0.5).astype(np.uint8)
        actions = rng.integers(0, na, n).astype(np.int32)
        logp = rng.standard_normal(n).astype(np.float32)
        rew = rng.standard_normal(n).astype(np.float32)
        u = rng.random(arenas)
        terms = np.where(u < pn, 1, np.where(u < pn + pt, 2, 0)).astype(np.uint8)
        obs_final = rng.standard_normal((n, obs)).astype(np.float32)
        obs_final[:, 0] = -t - 1
        yield obs_pre, masks, actions, logp, rew, terms, obs_final
