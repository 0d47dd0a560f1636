"""CPU suite: the N>1 host logic (batch sharding, max/sum over ranks) under gloo with
world_size 2; the sharded oracle results must reassemble to the unsharded ones."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mfl_b200 import sharding, synthetic


def test_shard_range_partitions():
    for n in (0, 1, 5, 16, 17):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import msda_oracle
        full = synthetic.make_inputs(5, 7, 2, 8, 2, [(1, 9), (1, 4)], torch.float32, "cpu", seed=11)
        mine = sharding.shard_batch(full, rank, world)
        a, b = sharding.shard_range(5, rank, world)
        out = msda_oracle.forward(mine["value"], mine["shapes"], mine["lsi"], mine["loc"], mine["aw"])
        ref = msda_oracle.forward(full["value"], full["shapes"], full["lsi"], full["loc"], full["aw"])
        ok = np.array_equal(out, ref[a:b])
        mx = sharding.max_over_ranks(10.0 + rank)
        sm = sharding.sum_over_ranks(b - a)
        q.put((rank, ok, mx, sm))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharding():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [True, True]
    assert all(r[2] == 11.0 for r in res)     # max over ranks
    assert all(r[3] == 5.0 for r in res)      # shards cover the batch exactly
