"""Longer randomized parity runs than the test suite affords (GPU box): every v1.0 archive, several
seeds, with and without auto-reset, plus dense random levels on awkward shapes.  Exits non-zero on the
first mismatch.    python scripts/soak_parity.py [steps] [envs]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import slo
from tests import golden_util as gu
from tests.test_gpu_parity import _random_levels
from safelife_k2_b200.safelife_env import SafeLifeEnv
from safelife_k2_b200.safelife_game import LevelPool

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 400
n = int(sys.argv[2]) if len(sys.argv) > 2 else 384


def run(name, lv, seed, auto_reset, time_limit, n=n, steps=steps, env0=11):
    pool = LevelPool(lv)
    P = pool.n_levels
    env = SafeLifeEnv(pool, num_envs=n, env_id0=env0, global_counter=None, compute_obs=False,
                      auto_reset=auto_reset, time_limit=time_limit)
    env.seed(seed); env.reset()
    lvl = (np.arange(n) + env0) % P
    o = slo.OracleBatchEnv(slo.OraclePool(**lv), lvl, seed=seed, env_id0=env0, threads=8, time_limit=time_limit)
    rng = np.random.RandomState(seed + 1000)
    stride = n % P or P
    t0 = time.time()
    for t in range(steps):
        acts = rng.randint(9, size=n)
        _, r, d, _ = env.step(acts)
        orr, od = o.step(acts)
        if not ((r.cpu().numpy() == orr).all() and (d.cpu().numpy() == (od != 0)).all()):
            print("MISMATCH reward/done", name, seed, auto_reset, "step", t); sys.exit(1)
        if auto_reset and od.any():
            idx = np.nonzero(od)[0]
            lvl[idx] = (lvl[idx] + stride) % P
            o.reset(idx, lvl[idx])
        if t % 40 == 0 or t == steps - 1:
            st, ost = env.game.state_numpy(), o.states
            ok = (env.game.board.cpu().numpy() == o.boards).all() and (env.game.goals.cpu().numpy() == o.goals).all()
            ok = ok and all((st[f] == ost[f]).all() for f in st.dtype.names)
            if not ok:
                print("MISMATCH state", name, seed, auto_reset, "step", t); sys.exit(1)
    print("ok %-24s seed %d auto_reset %d: %d envs x %d steps (%.1f s)" % (name, seed, auto_reset, n, steps, time.time() - t0))


for a in gu.ARCHIVES:
    lv = gu.levels(a)
    for seed in (1, 2):
        run(a, lv, seed, False, 1000)
        run(a, lv, seed, True, 37)
for (h, w) in [(26, 26), (32, 32), (7, 31), (31, 7), (4, 4), (12, 64), (33, 96), (5, 5), (3, 7), (50, 45)]:
    rng = np.random.RandomState(h * 131 + w)
    run("random %dx%d" % (h, w), _random_levels(rng, 16, h, w), h + w, True, 29, n=96, steps=min(steps, 200))
# the sliding-window shapes (rows of 2 / 4 / 8 words): fused step kernel of slb_big_rows.cuh
for (h, w) in [(256, 256), (128, 128), (256, 64), (192, 128), (384, 64)]:
    rng = np.random.RandomState(h * 131 + w)
    lv = _random_levels(rng, 6, h, w, exits=16)
    if h == 256:  # spawners in the goals too: the separate score pass
        m = rng.rand(*lv["goals"].shape)
        lv["goals"][m < 0.01] = 1688
    run("random %dx%d" % (h, w), lv, h + w, True, 23, n=24, steps=min(steps, 150))
print("SOAK OK")
