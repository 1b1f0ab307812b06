"""Multi-rank host logic on CPU (gloo, world_size 2): clip sharding and the optional result gather.
The data path itself has no collective - each rank retargets its own clips."""
import os
import socket

import numpy as np
import pytest


def test_shard_balances_frames_and_is_a_partition():
    from gmr_b200.shard import shard_clips, work_order
    rng = np.random.default_rng(0)
    lengths = np.clip(rng.lognormal(np.log(240), 0.6, size=1000), 60, 3000).astype(np.int64)
    for world in (1, 2, 4, 8):
        parts = shard_clips(lengths, world)
        allidx = np.concatenate(parts)
        assert sorted(allidx.tolist()) == list(range(1000))
        loads = np.array([lengths[p].sum() for p in parts])
        assert loads.max() - loads.min() <= lengths.max()
        for p in parts:
            assert (np.diff(lengths[p]) <= 0).all()          # longest first inside a shard
    o = work_order(lengths)
    assert o.dtype == np.int32 and (np.diff(lengths[o]) <= 0).all()
    # empty and tiny inputs
    assert [len(p) for p in shard_clips([], 4)] == [0, 0, 0, 0]
    assert sorted(np.concatenate(shard_clips([5, 7], 4)).tolist()) == [0, 1]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, lengths, q):
    import torch
    import torch.distributed as dist
    from gmr_b200.shard import gather_results, shard_clips
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        parts = shard_clips(lengths, world)
        mine = parts[rank]
        T, nq = 5, 7
        # stand-in for the rank's retargeted clips: value encodes the global clip id
        local = torch.stack([torch.full((T, nq), float(i)) for i in mine]) if len(mine) else torch.zeros((0, T, nq))
        full = gather_results(local.double(), mine, len(lengths))
        ok = all(bool((full[i] == float(i)).all()) for i in range(len(lengths)))
        q.put((rank, ok, [int(x) for x in mine]))
    finally:
        dist.destroy_process_group()


def test_two_rank_gather_gloo():
    import torch.multiprocessing as mp
    lengths = [300, 120, 900, 60, 450, 300, 75]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lengths, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res)
    got = sorted(i for _, _, idx in res for i in idx)
    assert got == list(range(len(lengths)))
