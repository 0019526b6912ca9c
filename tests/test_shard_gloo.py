"""CPU, world_size 2, gloo: the host-side logic of the multi-GPU path (frame ownership, descriptor all-gather in global
frame order, pair partition).  No compute kernel runs here."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from os1_b200 import shard
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        F, cap = 3, 5
        mine = shard.frames_for_rank(F * world, rank, world)
        desc = torch.zeros((F, cap, 32), dtype=torch.uint8)
        for li, gf in enumerate(mine):
            desc[li] = int(gf) + 1                      # tag every slab with its global frame number
        counts = torch.tensor([int(gf) % cap for gf in mine], dtype=torch.int32)
        d_all, c_all = shard.allgather_descriptors(desc, counts)
        ok = d_all.shape == (F * world, cap, 32)
        for gf in range(F * world):
            ok = ok and bool((d_all[gf] == gf + 1).all()) and int(c_all[gf]) == gf % cap
            ok = ok and shard.global_frame_index(gf // world, gf % world, world) == gf
        pairs = shard.pair_partition(F * world, rank, world)
        gathered = [None] * world
        dist.all_gather_object(gathered, pairs.tolist())
        q.put((rank, ok, gathered))
    finally:
        dist.destroy_process_group()


def test_world2_gloo_sharding():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in ps]
    res = [q.get(timeout=120) for _ in ps]
    [p.join(timeout=60) for p in ps]
    assert all(r[1] for r in res)
    allpairs = sorted(tuple(p) for part in res[0][2] for p in part)
    n = 6
    assert allpairs == sorted((i, j) for i in range(n) for j in range(i + 1, n))     # every pair exactly once


def test_frame_ownership_partitions_the_batch():
    from os1_b200 import shard
    for world in (1, 2, 4, 8):
        got = np.sort(np.concatenate([shard.frames_for_rank(37, r, world) for r in range(world)]))
        assert np.array_equal(got, np.arange(37))
