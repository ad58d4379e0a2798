"""Multi-GPU plumbing of the rollout path: one process per GPU, arenas sharded with NO data-path collective.

Arenas are independent units (no cross-arena state in RLGC::EnvSet), so every rank owns a contiguous range of
arenas, its own obs/reward/terminal buffers and its own slice of the experience stream; GAE segments never span ranks.
The only exchanges are three tiny statistics reductions per PPO iteration, which the reference computes on one process:
  * advantage normalisation  (GigaLearnCPP/src/private/GigaLearnCPP/PPO/PPOLearner.cpp:364-370): mean / unbiased std
  * return-std Welford state (GigaLearnCPP/src/private/GigaLearnCPP/Util/WelfordStat.h:16-25, Learner.cpp:959-967)
  * benchmark timing         (max over ranks of the device time, sum over ranks of the work)
They work on whatever backend the process group has (nccl on the GPUs, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist


def shard_arenas(total_arenas, world, rank):
    """Contiguous arena range [start, start+count) of `rank`; the first total % world ranks get one extra arena."""
    if world < 1 or not (0 <= rank < world) or total_arenas < 0:
        raise ValueError("bad shard request")
    base, extra = divmod(total_arenas, world)
    count = base + (1 if rank < extra else 0)
    start = rank * base + min(rank, extra)
    return start, count


def rank_seed(seed, rank):
    """Decorrelated per-rank seed for the device RNG (kickoff permutations, respawn slots)."""
    return (int(seed) * 0x9E3779B97F4A7C15 + (rank + 1) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF


def _world():
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def _reduce(t, op):
    if _world() > 1:
        dist.all_reduce(t, op=op)
    return t


def max_over_ranks(x, device="cpu"):
    return float(_reduce(torch.tensor([x], dtype=torch.float64, device=device), dist.ReduceOp.MAX).item())


def sum_over_ranks(x, device="cpu"):
    return float(_reduce(torch.tensor([x], dtype=torch.float64, device=device), dist.ReduceOp.SUM).item())


def global_advantage_stats(adv):
    """(mean, std) of the advantages of ALL ranks: one all-reduce of (sum, sum of squares, n) in float64.
    std is the unbiased estimator torch.std uses in PPOLearner.cpp:364-370."""
    a = adv.double()
    m = torch.stack([a.sum(), (a * a).sum(), torch.tensor(float(a.numel()), dtype=torch.float64, device=a.device)])
    _reduce(m, dist.ReduceOp.SUM)
    n = m[2]
    mean = m[0] / n
    var = (m[1] - n * mean * mean) / torch.clamp(n - 1, min=1.0)
    return float(mean), float(torch.sqrt(torch.clamp(var, min=0.0)))


def welford_merge_across_ranks(count, mean, m2, device="cpu"):
    """Chan's parallel merge of per-rank Welford states (count, mean, M2) -> the state of the union, so every rank feeds
    the same returnStd to its GAE kernel. Exact (no dependence on reduction order beyond float64 addition)."""
    t = torch.tensor([count, count * mean, 0.0], dtype=torch.float64, device=device)
    _reduce(t[:2], dist.ReduceOp.SUM)
    n = float(t[0])
    gmean = float(t[1]) / n if n > 0 else 0.0
    # M2_total = sum_r (M2_r + n_r * (mean_r - gmean)^2)
    u = torch.tensor([m2 + count * (mean - gmean) ** 2], dtype=torch.float64, device=device)
    _reduce(u, dist.ReduceOp.SUM)
    return n, gmean, float(u[0])


def welford_std(count, m2):
    """WelfordStat::GetSTD (WelfordStat.h:52-60): sqrt(M2 / (count - 1)), 1 until two samples are in."""
    return (m2 / (count - 1)) ** 0.5 if count > 1 else 1.0
