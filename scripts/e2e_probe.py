"""Where does a host-buffer step spend its time?  (GPU box)

    python scripts/e2e_probe.py [n_envs]        # SLB_HOST_GRAPH=0/1 selects the path of step_host
"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tests import golden_util as gu
from safelife_k2_b200 import _lib
from safelife_k2_b200.safelife_env import SafeLifeEnv
from safelife_k2_b200.safelife_game import LevelPool

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nb = int(sys.argv[2]) if len(sys.argv) > 2 else 14
pool = LevelPool(gu.levels("append-still"))
envs = []
for b in range(nb):
    e = SafeLifeEnv(pool, num_envs=n, env_id0=b * n, global_counter=None, compute_obs=False, auto_reset=True)
    e.seed(0); e.reset()
    envs.append(e)
lib = _lib.load()
h_act = torch.randint(0, 9, (n,), dtype=torch.uint8).pin_memory()
outs = [torch.zeros(5 * n, dtype=torch.uint8).pin_memory() for _ in range(4)]
ap = h_act.data_ptr()
op = [(o.data_ptr(), o.data_ptr() + 4 * n) for o in outs]
pc = time.perf_counter
ALL_STREAMS = [torch.cuda.Stream() for _ in range(4)]


def pipelined(depth, steps, same_batch=False):
    streams = ALL_STREAMS[:depth]
    t_step = t_wait = 0.0

    def loop(count, timed):
        nonlocal t_step, t_wait
        for i in range(count):
            s = i % depth
            e = envs[0 if same_batch else i % nb]
            a = pc()
            if i >= depth:
                e.wait(streams[s])
            b = pc()
            e.step_host(ap, op[s][0], op[s][1], sync=False, stream=streams[s])
            c = pc()
            if timed:
                t_wait += b - a
                t_step += c - b
        for s in range(depth):
            envs[0].wait(streams[s])

    loop(2 * nb + depth, False)
    t0 = pc()
    loop(steps, True)
    dt = pc() - t0
    print("depth %d%s: %6.2f us/step  (in step call %5.2f, in wait %5.2f)  -> %.0f M env-steps/s" % (
        depth, " same batch" if same_batch else "", dt / steps * 1e6, t_step / steps * 1e6, t_wait / steps * 1e6,
        n * steps / dt / 1e6))


print("SLB_HOST_GRAPH=%s n=%d batches=%d" % (os.environ.get("SLB_HOST_GRAPH", "1"), n, nb))
for depth in (1, 2, 3, 4):
    pipelined(depth, 2000)
pipelined(2, 2000, same_batch=True)
