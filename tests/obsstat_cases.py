"""Shared obs-statistics test material: a seeded sequence of collection steps (obs rows, sampled row indices)."""
import numpy as np

# (rows, width, samples per step, steps, mean_clamp, min_std)   Learner defaults: maxObsSamples 100, maxObsMeanRange 3, minObsSTD 0.1
CASES = [
    (64, 89, 100, 6, 3.0, 0.1),        # more samples than rows: rows repeat
    (512, 167, 100, 5, 3.0, 0.1),      # 2v2 AdvancedObs
    (300, 225, 37, 4, 0.5, 0.5),       # clamps that bite
    (7, 5, 1, 3, 3.0, 0.1),            # one sample per step: the first normalise is a no-op (count < 2); rows*width not a multiple of 4
]


def steps(case, seed=0):
    rows, width, ns, nsteps, _, _ = case
    rng = np.random.default_rng(seed)
    scale = rng.uniform(0.01, 8.0, width).astype(np.float32)
    shift = rng.uniform(-6.0, 6.0, width).astype(np.float32)
    scale[::11] = 0.0                                      # constant columns: variance 0 -> std 1 (:121)
    for _ in range(nsteps):
        obs = (rng.standard_normal((rows, width)).astype(np.float32) * scale + shift).astype(np.float32)
        idx = rng.integers(0, rows, ns).astype(np.int32)
        yield obs, idx
