"""TEST INFRASTRUCTURE — NOT PART OF THE PRODUCT (only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline may import it).

numpy restatement of GGL::ExperienceBuffer batching (reference: GigaLearnCPP/src/private/GigaLearnCPP/PPO/ExperienceBuffer.cpp):
  plan_batches   GetAllBatchesShuffled's partition of the permuted stream, :132-160
  get_samples    _GetSamples, :15-98 (row limit, index clamp, index_select of the six tensors, action clamp)
The permutation itself is std::shuffle over std::default_random_engine (libstdc++-specific, :129): it is DATA here, taken from
the compiled reference in the tests (oracle/_ref/libref_expbuf.so, tests/test_expbuf_oracle_cpu.py pins this file against it).
"""
import numpy as np


def exp_size(rows_states, rows_actions=None):
    """:108-115 — rows of `states`, limited by the rows of `actions` when that tensor is defined."""
    n = int(rows_states)
    if rows_actions is not None:
        n = min(n, int(rows_actions))
    return max(n, 0)


def plan_batches(n, batch_size, overbatching):
    """[(start, size)] over the permuted stream, :143-160. With overbatching the last batch takes everything that is left once
    fewer than 2*batch_size samples remain; without it the remainder is dropped."""
    out = []
    start = 0
    while start < n:
        end = start + batch_size
        if end + batch_size > n and overbatching:
            end = n
        if end > n:
            break
        if end - start <= 0:
            break
        out.append((start, end - start))
        if end == n:
            break
        start += batch_size
    return out


def get_samples(data, indices, max_action_index=-1):
    """data: dict of the six arrays (any may be None). :15-98."""
    states = data.get("states")
    if states is None or len(indices) == 0:
        return {}
    row_limit = exp_size(states.shape[0], None if data.get("actions") is None else data["actions"].shape[0])
    if row_limit <= 0:
        return {}
    idx = np.asarray(indices, np.int64)
    if idx.max(initial=0) >= row_limit:                 # :42-69 clamp to the last valid row
        idx = np.minimum(idx, row_limit - 1)
    out = {k: (None if v is None else v[idx]) for k, v in data.items()}
    if max_action_index >= 0 and out.get("actions") is not None:   # :88-90
        out["actions"] = np.clip(out["actions"], 0, max_action_index)
    return out
