"""Find the first (step, env, cell) where the CUDA env diverges from the oracle (GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import slo
from tests import golden_util as gu
from safelife_k2_b200.safelife_env import SafeLifeEnv
from safelife_k2_b200.safelife_game import LevelPool

archive, n, steps, seed, aseed = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
lv = gu.levels(archive)
pool = LevelPool(lv)
env = SafeLifeEnv(pool, num_envs=n, env_id0=77, global_counter=None, compute_obs=False)
env.seed(seed); env.reset()
lvl = (np.arange(n) + 77) % pool.n_levels
oenv = slo.OracleBatchEnv(slo.OraclePool(**lv), lvl, seed=seed, env_id0=77, threads=8)
rng = np.random.RandomState(aseed)
prev_b = oenv.boards.copy()
for t in range(steps):
    acts = rng.randint(9, size=n)
    pst = oenv.states.copy()
    env.step(acts); oenv.step(acts)
    b = env.game.board.cpu().numpy(); g = env.game.goals.cpu().numpy()
    bad = np.nonzero((b != oenv.boards).reshape(n, -1).any(1) | (g != oenv.goals).reshape(n, -1).any(1))[0]
    if len(bad):
        i = bad[0]
        print("step", t, "env", i, "level", lvl[i], "action", acts[i], "n bad envs", len(bad))
        print("agent before", pst[i]["agent_x"], pst[i]["agent_y"], "orient", pst[i]["orientation"])
        ys, xs = np.nonzero(b[i] != oenv.boards[i])
        print("board diffs (y,x,dev,oracle,prev):", [(int(y), int(x), int(b[i][y, x]), int(oenv.boards[i][y, x]), int(prev_b[i][y, x])) for y, x in zip(ys, xs)][:20])
        ys, xs = np.nonzero(g[i] != oenv.goals[i])
        print("goal diffs:", [(int(y), int(x), int(g[i][y, x]), int(oenv.goals[i][y, x])) for y, x in zip(ys, xs)][:20])
        y0 = max(0, ys.min() - 2) if len(ys) else 0
        np.set_printoptions(linewidth=250)
        print("prev board rows:\n", prev_b[i][max(0, pst[i]["agent_y"] - 3):pst[i]["agent_y"] + 4])
        print("st dev", env.game.state_numpy()[i], "\nst ora", oenv.states[i])
        break
    prev_b = oenv.boards.copy()
else:
    print("no divergence in", steps, "steps")
