"""A few launches of each kernel that profiles/r02 documents, for `ncu -k regex:<kernel>`:

    python scripts/profile_kernels.py step_prune_spawn big_256 advance_pairs side_effects emd

Small and fixed: the same command is run once plainly and once under ncu (B200_PROFILING.md)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import bench_configs as bc
from safelife_k2_b200 import _lib
from safelife_k2_b200 import side_effects as se
from safelife_k2_b200.safelife_env import SafeLifeEnv
from safelife_k2_b200.safelife_game import LevelPool
from tests import golden_util as gu

names = sys.argv[1:] or ["step_prune_spawn", "big_256", "advance_pairs", "side_effects", "emd"]
dev = torch.device("cuda", 0)
lib = _lib.load()
st = torch.cuda.current_stream().cuda_stream


def steps(pool, n, count, ring=3):
    envs, _ = bc.ring_envs(pool, n, min_bytes=0, min_ring=ring)
    acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(16, n)).astype(np.uint8)).to(dev)
    launch = bc.step_launcher(pool, envs, acts)
    for j in range(count):
        launch(j, st)
    torch.cuda.synchronize()


if "step_prune_spawn" in names:   # BASELINE config 3: res_env_rollout_kernel<15,3,26,1>, 16 384 envs per launch
    steps(LevelPool(gu.levels("prune-spawn")), 16384, 12)
if "step_mixed" in names:         # config 5a share of one of 8 GPUs
    steps(LevelPool(bc.mixed_levels()), 8192, 12)
if "big_256" in names:            # config 5b: big_rows_env_kernel<8> (SLB_BIG_ROWS=0: big_env_kernel), 512 envs
    steps(LevelPool(bc.synth_256(np.random.RandomState(5), np.random.RandomState(6), 8)), 512, 6, ring=1)
if "advance_pairs" in names:      # the speedups.advance_board contract, k = 1: advance_pairs_kernel
    n = 131072
    p = torch.full((n,), 0.3, dtype=torch.float32, device=dev)
    for arch in ("append-still", "prune-spawn"):
        lv = gu.levels(arch)
        idx = torch.arange(n, device=dev) % 100
        src = torch.from_numpy(lv["board"]).to(dev).view(torch.int16)[idx].contiguous().view(torch.uint16)
        dst = torch.empty_like(src)
        for _ in range(4):
            _lib.check(lib.slb_advance_board(src.data_ptr(), dst.data_ptr(), n, 26, 26, p.data_ptr(), 0, 0, None, 1,
                                             None, st))
    torch.cuda.synchronize()
if "side_effects" in names or "emd" in names:
    pool = LevelPool(gu.levels("prune-spawn"))
    env = SafeLifeEnv(pool, num_envs=2048, global_counter=None, compute_obs=False)
    env.seed(0)
    env.reset()
    acts = torch.from_numpy(np.random.RandomState(3).randint(9, size=(300, 2048)).astype(np.uint8)).to(dev)
    env.rollout(acts)
    g = env.game
    b0 = pool.board.view(torch.int16)[g.level_id.long()].view(torch.uint16)
    keys, counts, nkeys = se.side_effect_counts(b0, g.board, g.num_steps.contiguous(), g.spawn_prob, 1000, 0, 0, None, 16)
    if "emd" in names:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        emd = se.emd_batch(keys, counts, 1000)
        e1.record()
        torch.cuda.synchronize()
        print("emd_batch: %d problems in %.2f ms, max %.3f" % (int((keys != 0).sum()), e0.elapsed_time(e1), float(emd.max())))
torch.cuda.synchronize()
print("PROFILE KERNELS DONE", names)
