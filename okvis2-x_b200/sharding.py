"""Host-side sharding rules of the multi-GPU path; one process per GPU, torch.distributed plumbing.

* Frames / cameras: independent units, sharded round-robin, no data-path collective.
* Matcher sweep: map rows block-partitioned; every rank computes best[q] = (dist << 32) | GLOBAL row index (or
  NO_MATCH) over its rows; one all-gather + element-wise min reproduces the reference's single-machine result,
  first-wins ties included, because the key orders by distance and then by global index.
"""
import numpy as np

NO_MATCH_I64 = -1  # 0xFFFFFFFFFFFFFFFF viewed as int64


def frame_of(step, rank, world, n_frames):
    """Frame index processed by `rank` at its local step `step` (round-robin over ranks, cyclic over the pool)."""
    return (rank + step * world) % n_frames


def map_row_shard(rank, world, nm):
    """[lo, hi) of the map rows owned by `rank` (block partition, last shard may be short or empty)."""
    per = (nm + world - 1) // world
    lo = min(nm, rank * per)
    return lo, min(nm, lo + per)


def merge_best(best_local, world, dist=None):
    """All-gather + min of per-shard best keys. best_local: torch int64 tensor [nq] holding the uint64 keys' bits
    (NO_MATCH = -1). Works on CUDA tensors with NCCL and on CPU tensors with gloo. Returns int64 keys, NO_MATCH = -1."""
    import torch
    if world == 1:
        return best_local
    gathered = torch.empty(world * best_local.numel(), dtype=torch.int64, device=best_local.device)
    dist.all_gather_into_tensor(gathered, best_local.contiguous())
    big = torch.iinfo(torch.int64).max
    # keys are < 2^63 (dist <= 384), so int64 order == uint64 order except for NO_MATCH (-1): lift it above all keys
    m = torch.where(gathered < 0, torch.full_like(gathered, big), gathered).view(world, -1).min(0).values
    return torch.where(m == big, torch.full_like(m, NO_MATCH_I64), m)


def merge_best_numpy(parts):
    """Same rule on host arrays of uint64 keys (tests, single-process emulation of the ranks)."""
    return np.minimum.reduce([np.asarray(p, np.uint64) for p in parts])
