#!/usr/bin/env python
"""BASELINE config 5, the scaling sweep: a FIXED total of envs sharded by index over the GPUs of one box
(strong scaling), one process per GPU, no data-path collective, NCCL all-reduce of the statistics vector.

    python bench_scaling.py --workload mixed_26 [--total-envs 65536] [--steps 300]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench_scaling.py --workload big_256 [--total-envs 4096] [--steps 20]

workloads: mixed_26 = 26x26 levels cycled over all 8 v1.0 archives (config 5a);
           big_256  = the synthetic 256x256 levels of SURVEY 8(d) (config 5b).
Rank 0 prints one JSON line: whole-job env-steps/s and cell-updates/s, time = max over ranks (CUDA events).
The headline bench (weak scaling, 4096 envs per GPU) is bench.py; this file only covers the sweep.
"""
import argparse
import json

import numpy as np
import torch

from bench_configs import synth_256
from safelife_k2_b200 import _lib, distributed
from safelife_k2_b200.safelife_env import SafeLifeEnv
from safelife_k2_b200.safelife_game import LevelPool
from tests import golden_util as gu

L2_BYTES = 126e6


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="mixed_26", choices=["mixed_26", "big_256"])
    ap.add_argument("--total-envs", type=int, default=None)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    # keep stdout for the one JSON line: NCCL prints its version banner there when the communicator is made
    import os
    import sys
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    try:
        rank, world, local = distributed.init_from_env()
        torch.cuda.set_device(local)
        dev = torch.device("cuda", local)
        if world > 1:
            torch.distributed.barrier(device_ids=[local])
            torch.cuda.synchronize()
    finally:
        sys.stdout.flush()
        os.dup2(saved, 1)
        os.close(saved)
    if args.workload == "mixed_26":
        parts = [gu.levels(a) for a in gu.ARCHIVES]
        lv = {k: np.concatenate([p[k] for p in parts]) for k in parts[0]}
        total, h, w = args.total_envs or 65536, 26, 26
        steps = args.steps or 300
    else:
        lv = synth_256(np.random.RandomState(5), np.random.RandomState(6), 8)
        total, h, w = args.total_envs or 4096, 256, 256
        steps = args.steps or 20
    lo, hi = distributed.shard_range(total, rank, world)
    n = hi - lo
    pool = LevelPool(lv, device=dev)
    # independent batches of this rank's share until their state exceeds L2 (one is enough for big boards)
    per = n * (2 * pool.resident_bytes + 64)
    ring = max(1, int(np.ceil(2.2 * L2_BYTES / per)))
    envs = []
    for r in range(ring):
        e = SafeLifeEnv(pool, num_envs=n, env_id0=lo + r * total, global_counter=None, compute_obs=False,
                        auto_reset=True)
        e.seed(0)
        e.reset()
        envs.append(e)
    acts = torch.from_numpy(np.random.RandomState(3 + rank).randint(9, size=(16, n)).astype(np.uint8)).to(dev)
    stats = torch.zeros(8, dtype=torch.int64, device=dev)
    lib, st = _lib.load(), torch.cuda.current_stream().cuda_stream

    def step(i):
        e = envs[i % ring]
        g = e.game
        _lib.check(lib.slb_env_step(pool.c, n, g.board_res.data_ptr(), g.goals_res.data_ptr(), g.state.data_ptr(),
                                    acts[i % 16].data_ptr(), 1000, 0, g.env_id0, e._reward.data_ptr(),
                                    e._done.data_ptr(), e._stride(), st))

    for i in range(max(3, args.warmup)):
        step(i)
    torch.cuda.synchronize()
    # one pass over the ring as a CUDA graph (as in bench.py): at 8 192 envs per GPU a step is ~9 us, below
    # the host's ctypes launch rate
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        st = side.cuda_stream
        with torch.cuda.graph(graph, stream=side):
            for i in range(ring):
                step(i)
    st = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()
    graph.replay()
    torch.cuda.synchronize()
    reps = max(1, steps // ring)
    steps = reps * ring
    if world > 1:
        torch.distributed.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        graph.replay()
    last = envs[(steps - 1) % ring]
    _lib.check(lib.slb_accumulate_stats(n, last.game.state.data_ptr(), last._reward.data_ptr(),
                                        last._done.data_ptr(), stats.data_ptr(), st))
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    distributed.all_reduce_stats(stats)
    if world > 1:
        torch.distributed.barrier()
    ms = distributed.max_over_ranks(ms, dev)
    if rank == 0:
        value = total * steps / (ms * 1e-3)
        print(json.dumps({"workload": args.workload, "n_gpus": world, "scaling": "strong", "total_envs": total,
                          "envs_per_gpu": n, "board": [h, w], "steps": steps, "ms_per_step": ms / steps,
                          "env_steps_per_s": value, "cell_updates_per_s": value * 2 * h * w,
                          "ring_batches_per_gpu": ring, "stats_allreduce": distributed.stats_dict(stats)}))
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
