"""Clip sharding across ranks (one process per GPU) - pure data parallelism, no per-step collective.

The reference parallelises over files with ``multiprocessing.Pool`` (scripts/smplx_to_robot_dataset.py:241-242);
here the unit is the clip and the worker is a GPU rank.  Clips are independent (each starts cold from qpos0,
scripts/bvh_to_robot_dataset.py:84-88), so a rank needs nothing from the others; the only optional exchange is
a final gather of the results.
"""
from __future__ import annotations

import numpy as np


def shard_clips(lengths, world_size: int):
    """Partition clips over ranks balancing total frames (longest-processing-time first).
    -> list (one entry per rank) of int64 index arrays, each sorted by decreasing length
    (the order the device work queue wants for ragged batches)."""
    lengths = np.asarray(lengths, np.int64)
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    order = np.argsort(-lengths, kind="stable")
    load = np.zeros(world_size, np.int64)
    count = np.zeros(world_size, np.int64)
    buckets = [[] for _ in range(world_size)]
    for i in order:
        # least-loaded rank; ties -> fewest clips, then lowest rank (deterministic on every rank)
        r = int(np.lexsort((np.arange(world_size), count, load))[0])
        buckets[r].append(int(i))
        load[r] += lengths[i]
        count[r] += 1
    return [np.asarray(b, np.int64) for b in buckets]


def work_order(lengths) -> np.ndarray:
    """Device queue order for a ragged batch: longest clips first (int32 permutation)."""
    return np.argsort(-np.asarray(lengths, np.int64), kind="stable").astype(np.int32)


def gather_results(local_qpos, local_index, n_total: int, group=None):
    """Optional final gather: every rank contributes ``local_qpos [n_r, T, nq]`` (torch tensor, any device the
    process group supports) for the clips ``local_index [n_r]``; returns the full ``[n_total, T, nq]`` tensor
    on every rank.  One padded all_gather - the only collective of the whole job."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n_r = torch.tensor([local_qpos.shape[0]], dtype=torch.int64, device=local_qpos.device)
    counts = [torch.zeros_like(n_r) for _ in range(world)]
    dist.all_gather(counts, n_r, group=group)
    n_max = int(max(c.item() for c in counts))
    T, nq = local_qpos.shape[1:]
    pad_q = torch.zeros((n_max, T, nq), dtype=local_qpos.dtype, device=local_qpos.device)
    pad_q[:local_qpos.shape[0]] = local_qpos
    pad_i = torch.full((n_max,), -1, dtype=torch.int64, device=local_qpos.device)
    pad_i[:local_qpos.shape[0]] = torch.as_tensor(local_index, dtype=torch.int64, device=local_qpos.device)
    all_q = [torch.zeros_like(pad_q) for _ in range(world)]
    all_i = [torch.zeros_like(pad_i) for _ in range(world)]
    dist.all_gather(all_q, pad_q, group=group)
    dist.all_gather(all_i, pad_i, group=group)
    out = torch.zeros((n_total, T, nq), dtype=local_qpos.dtype, device=local_qpos.device)
    for q, i in zip(all_q, all_i):
        keep = i >= 0
        out[i[keep]] = q[keep]
    return out
