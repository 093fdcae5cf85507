"""One small pass over the kernels with hand-rolled protocols, for compute-sanitizer (VERDICT r01 item 6 iii):

    compute-sanitizer --tool memcheck  python scripts/sanitize_paths.py
    compute-sanitizer --tool racecheck python scripts/sanitize_paths.py

smoke() (fused step + observation vs the oracle), a host-stepper loop (the last-CTA ticket / system fence /
host-spin hand-off of res_env_step_host_kernel), one side-effect call (the shared nibble counters of
side_effect_fast_kernel), one EMD batch, one 256x256 step (the sliding-window big-board kernel) and one
uint16 advance_board call.  Small sizes: the tools run the kernels 10-100x slower."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

import __graft_entry__ as ge
from bench_configs import synth_256
from safelife_k2_b200 import side_effects as se
from safelife_k2_b200 import speedups as sp
from safelife_k2_b200.safelife_env import SafeLifeEnv
from safelife_k2_b200.safelife_game import LevelPool
from tests import golden_util as gu

ge.smoke()
lv = gu.level_slice(gu.levels("prune-spawn"), 0, 16)
pool = LevelPool(lv)
env = SafeLifeEnv(pool, num_envs=96, global_counter=None, compute_obs=False, auto_reset=True, time_limit=20)
env.seed(1)
env.reset()
rng = np.random.RandomState(0)
streams = [torch.cuda.Stream() for _ in range(2)]
hs = [env.host_stepper(s) for s in streams]
acts = torch.from_numpy(rng.randint(9, size=(8, 96)).astype(np.uint8)).pin_memory()
outs = [torch.zeros(5 * 96, dtype=torch.uint8).pin_memory() for _ in range(2)]
for t in range(24):  # strictly alternating streams on one batch: step, wait
    k = t % 2
    hs[k].step(acts[t % 8].data_ptr(), outs[k].data_ptr(), outs[k].data_ptr() + 4 * 96)
    hs[k].wait()
print("host stepper ok")
res = se.side_effect_score(env.game, num_samples=50)
print("side effects + emd ok:", len(res), "episodes")
big = LevelPool(synth_256(np.random.RandomState(5), np.random.RandomState(6), 2))
benv = SafeLifeEnv(big, num_envs=3, global_counter=None, compute_obs=False, auto_reset=True)
benv.seed(2)
benv.reset()
for t in range(3):
    benv.step(rng.randint(9, size=3))
print("256x256 ok")
b = sp.advance_board(lv["board"][:5], 0.3)
b = sp.advance_board(torch.from_numpy(lv["board"][:40]).cuda(), 0.3)
torch.cuda.synchronize()
print("SANITIZE PATHS DONE")
