"""Side-effect count maps of finished episodes of EVERY v1.0 archive against the oracle (GPU box): the
kernel's run-length / fixed-point shortcuts see still, oscillating and spawner boards here.  Exits
non-zero on the first mismatch.    python scripts/soak_side_effects.py [envs] [checked per archive]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import slo
from tests import golden_util as gu
from safelife_k2_b200.safelife_env import SafeLifeEnv
from safelife_k2_b200.safelife_game import LevelPool
from safelife_k2_b200 import side_effects as se

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
checked = int(sys.argv[2]) if len(sys.argv) > 2 else 32
for a in gu.ARCHIVES:
    lv = gu.levels(a)
    pool = LevelPool(lv)
    for T, seed in ((60, 1), (400, 2)):
        env = SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False)
        env.seed(seed); env.reset()
        acts = torch.from_numpy(np.random.RandomState(seed).randint(9, size=(T, n)).astype(np.uint8)).cuda()
        env.rollout(acts)
        g = env.game
        st = g.state_numpy()
        b0 = lv["board"][st["level_id"]]
        b1 = g.board.cpu().numpy()
        ctr0 = st["rng_counter"].astype(np.int64)
        ctr = torch.from_numpy(ctr0.copy()).cuda()
        spawn = lv["spawn_prob"][st["level_id"]].astype(np.float32)
        t0 = time.time()
        keys, counts, nkeys = se.side_effect_counts(torch.from_numpy(b0).cuda(), g.board, st["num_steps"].astype(np.int32),
                                                    spawn, num_samples=1000, seed=seed, env_id0=0, counters=ctr, kmax=32)
        torch.cuda.synchronize()
        dt = time.time() - t0
        keys, counts, nk, ctr = keys.cpu().numpy(), counts.cpu().numpy(), nkeys.cpu().numpy(), ctr.cpu().numpy()
        for i in np.linspace(0, n - 1, checked).astype(int):
            ok, oc, octr = slo.side_effect_counts(b0[i], b1[i], int(st["num_steps"][i]), num_samples=1000,
                                                  spawn_prob=float(spawn[i]), seed=seed, env=int(i), counter=int(ctr0[i]), kmax=32)
            mine = {int(k): counts[i, :, s] for s, k in enumerate(keys[i]) if k}
            good = nk[i] == len(ok) and ctr[i] == octr and sorted(mine) == sorted(int(k) for k in ok)
            good = good and all((mine[int(k)] == oc[:, s]).all() for s, k in enumerate(ok))
            if not good:
                print("MISMATCH", a, "T", T, "episode", i); sys.exit(1)
        print("ok %-20s after %3d actions: %d episodes in %.1f ms, %d checked, up to %d cell types" % (a, T, n, dt * 1e3, checked, nk.max()))
print("SOAK OK")
