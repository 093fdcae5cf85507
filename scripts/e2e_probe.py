"""Where does a host-buffer step spend its time?  (GPU box)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tests import golden_util as gu
from safelife_k2_b200 import _lib
from safelife_k2_b200.safelife_env import SafeLifeEnv
from safelife_k2_b200.safelife_game import LevelPool

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
pool = LevelPool(gu.levels("append-still"))
env = SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False, auto_reset=True)
env.seed(0); env.reset()
g = env.game
lib = _lib.load()
st = torch.cuda.current_stream().cuda_stream
acts = torch.randint(0, 9, (n,), dtype=torch.uint8, device="cuda")
h_act = torch.randint(0, 9, (n,), dtype=torch.uint8).pin_memory()
h_out = torch.zeros(5 * n, dtype=torch.uint8).pin_memory()
hr, hd = h_out.data_ptr(), h_out.data_ptr() + 4 * n
args = (pool.c, n, g.board_res.data_ptr(), g.goals_res.data_ptr(), g.state.data_ptr())
tail = (1000, 0, 0)

def bench(name, fn, iters=2000):
    for _ in range(20): fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(iters): fn()
    torch.cuda.synchronize()
    print("%-40s %7.2f us/step" % (name, (time.perf_counter() - t0) / iters * 1e6))

bench("device step, no sync (launch rate)", lambda: lib.slb_env_step(*args, acts.data_ptr(), *tail, env._reward.data_ptr(), env._done.data_ptr(), 100, st))
def dsync():
    lib.slb_env_step(*args, acts.data_ptr(), *tail, env._reward.data_ptr(), env._done.data_ptr(), 100, st)
    torch.cuda.current_stream().synchronize()
bench("device step + stream sync", dsync)
for z in ("1", "0"):
    os.environ["SLB_E2E_ZEROCOPY"] = z
bench("step_host (env default)", lambda: lib.slb_env_step_host(*args, h_act.data_ptr(), *tail, hr, hd, 100, st))
bench("env.step_host wrapper", lambda: env.step_host(h_act, h_out[:4 * n].view(torch.int32), h_out[4 * n:]))
# zero-copy actions only / outputs only
def zc_in():
    lib.slb_env_step(*args, h_act.data_ptr(), *tail, env._reward.data_ptr(), env._done.data_ptr(), 100, st)
    torch.cuda.current_stream().synchronize()
bench("kernel reads pinned actions, dev outputs", zc_in)
def zc_out():
    lib.slb_env_step(*args, acts.data_ptr(), *tail, hr, hd, 100, st)
    torch.cuda.current_stream().synchronize()
bench("dev actions, kernel writes pinned outputs", zc_out)
