"""Shared ExperienceBuffer test material (seeded inputs; the sample id rides in log_probs so the permutation can be read back)."""
import numpy as np

FIELDS = ("states", "actions", "log_probs", "target_values", "action_masks", "advantages")
# (rows, rows_actions, obs, num_actions, batch_size, overbatching, max_action_index)
CASES = [
    (1000, 1000, 89, 90, 256, True, -1),
    (1000, 1000, 89, 90, 256, False, -1),
    (1024, 1024, 167, 90, 256, True, -1),       # exact multiple
    (100, 100, 51, 90, 256, True, -1),          # fewer samples than one batch: one overbatch
    (100, 100, 51, 90, 256, False, -1),         # ... or nothing at all
    (777, 700, 33, 7, 100, True, 3),            # actions shorter than states (row limit), odd mask rows, action clamp
    (513, 513, 225, 90, 512, True, 50),
    (1, 1, 5, 3, 1, False, -1),
]


def make_data(rows, rows_actions, obs, na, seed=0):
    rng = np.random.default_rng(seed)
    return {"states": rng.standard_normal((rows, obs)).astype(np.float32),
            "actions": rng.integers(-2, na + 3, rows_actions).astype(np.int32),
            "log_probs": np.arange(rows, dtype=np.float32),                 # sample id (exact in f32 below 2^24)
            "target_values": rng.standard_normal(rows).astype(np.float32),
            "action_masks": (rng.random((rows, na)) < 0.4).astype(np.uint8),
            "advantages": rng.standard_normal(rows).astype(np.float32)}
