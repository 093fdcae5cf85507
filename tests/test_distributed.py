"""CPU, world_size 2 over gloo: the multi-process plumbing of the sharded env path — contiguous env
shards that tile [0, N), the statistics all-reduce, max-over-ranks timing (safelife_k2_b200/distributed.py).
The data path itself has no collective."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from safelife_k2_b200 import distributed as D


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_total, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    r, w, _ = D.init_from_env(backend="gloo")
    assert (r, w) == (rank, world)
    lo, hi = D.shard_range(n_total, rank, world)
    # every rank owns the envs [lo, hi); its statistics are a function of the GLOBAL env index only
    idx = np.arange(lo, hi)
    stats = torch.tensor([idx.sum(), len(idx), (idx % 7).sum(), 0, 0, 0, len(idx), 0], dtype=torch.int64)
    D.all_reduce_stats(stats)
    t = D.max_over_ranks(1.0 + rank)
    out.put((rank, lo, hi, stats.tolist(), t))
    dist.barrier()
    dist.destroy_process_group()


def test_sharding_and_stats_allreduce_world2():
    n_total, world = 4097, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, world, port, n_total, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, lo0, hi0, s0, t0), (r1, lo1, hi1, s1, t1) = res
    assert lo0 == 0 and hi0 == lo1 and hi1 == n_total  # shards tile [0, N) without overlap
    allidx = np.arange(n_total)
    want = [allidx.sum(), n_total, (allidx % 7).sum(), 0, 0, 0, n_total, 0]
    assert s0 == want and s1 == want                    # both ranks hold the global sums
    assert t0 == t1 == 2.0                              # slowest rank's time


def test_shard_range_properties():
    for n in [0, 1, 7, 4096, 65536, 1000003]:
        for w in [1, 2, 3, 4, 8]:
            edges = [D.shard_range(n, r, w) for r in range(w)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(edges[:-1], edges[1:]))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1
