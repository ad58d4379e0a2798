"""Seeded GAE input generator shared by the golden-vector script and the tests:
a stream of N segments of length T, last element of each segment terminal."""
import numpy as np


def make_stream(T, N, kind, seed=7, ragged=False):
    """kind: 'normal' | 'trunc' | 'mixed' | 'open' (last segment left unterminated)."""
    rng = np.random.default_rng(seed)
    if ragged:
        lens = rng.integers(1, 2 * T + 1, size=N)
    else:
        lens = np.full(N, T)
    L = int(lens.sum())
    rews = rng.standard_normal(L).astype(np.float32)
    vals = rng.standard_normal(L).astype(np.float32)
    terms = np.zeros(L, np.int8)
    ends = np.cumsum(lens) - 1
    if kind == "normal":
        terms[ends] = 1
    elif kind == "trunc":
        terms[ends] = 2
    else:
        terms[ends] = rng.integers(1, 3, size=N).astype(np.int8)
    if kind == "open":
        terms[ends[-1]] = 0
    K = int((terms == 2).sum())
    tv = rng.standard_normal(K).astype(np.float32)
    return rews, terms, vals, tv


GOLDEN_CASES = [
    # (name, T, N, kind, ragged, gamma, lambda, return_std, clip_range, pass_trunc_vals)
    ("t1_n64_normal", 1, 64, "normal", False, 0.99, 0.95, 0.0, 10.0, True),
    ("t8_n32_mixed_std", 8, 32, "mixed", False, 0.99, 0.95, 1.7, 10.0, True),
    ("t8_n32_mixed_clip", 8, 32, "mixed", False, 0.99, 0.95, 0.3, 1.0, True),
    ("t16_n16_trunc", 16, 16, "trunc", False, 0.99, 0.95, 1.7, 10.0, True),
    ("t16_n16_trunc_undef", 16, 16, "trunc", False, 0.99, 0.95, 1.0, 10.0, False),
    ("ragged_open", 12, 40, "open", True, 0.995, 0.9, 1.7, 0.0, True),
    ("t256_n20_mixed", 256, 20, "mixed", False, 0.99, 0.95, 1.7, 10.0, True),
]
