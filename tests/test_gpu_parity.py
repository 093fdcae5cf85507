"""GPU parity tests: the CUDA path (through the C ABI / the host mirror of the reference interface)
against the golden vectors generated from the reference and against the CPU oracle on the same
seeded inputs.  Bit-exact: this is integer / byte work."""
import numpy as np
import pytest
import torch

from oracle import slo
from tests import golden_util as gu

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sl():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from safelife_k2_b200 import _lib, safelife_env, safelife_game, side_effects, speedups

    _lib.load()  # fails loudly if the CUDA library is missing
    return type("M", (), dict(lib=_lib, env=safelife_env, game=safelife_game, se=side_effects,
                              speedups=speedups))


def dev_advance(sl, boards, p, seed, env0, counters=None, k=1):
    """boards: numpy [N,H,W] -> (numpy out, counters) through slb_advance_board on the device."""
    b = torch.from_numpy(np.ascontiguousarray(boards, dtype=np.uint16)).cuda()
    n, h, w = b.shape
    out = torch.empty_like(b)
    pp = torch.full((n,), float(np.float32(p)), dtype=torch.float32, device="cuda")
    ctr = torch.zeros(n, dtype=torch.int64, device="cuda")
    if counters is not None:
        ctr.copy_(torch.from_numpy(np.asarray(counters, dtype=np.int64)))
    scratch = torch.empty_like(b) if k > 1 else None
    sl.lib.check(sl.lib.load().slb_advance_board(b.data_ptr(), out.data_ptr(), n, h, w, pp.data_ptr(),
                                                 seed, env0, ctr.data_ptr(), k,
                                                 sl.lib.ptr(scratch), sl.lib.current_stream()))
    torch.cuda.synchronize()
    return out.cpu().numpy(), ctr.cpu().numpy()


def test_advance_board_known_answers(sl):
    n = 0
    for name, env, seed, board, outs, p in gu.kat_cases():
        b, ctr = board[None], None
        for t in range(outs.shape[0]):
            b, ctr = dev_advance(sl, b, p, seed, env, ctr)
            assert (b[0] == outs[t]).all(), "%s step %d" % (name, t)
        # k fused steps give the same final board and counter
        bk, ctrk = dev_advance(sl, board[None], p, seed, env, None, k=outs.shape[0])
        assert (bk[0] == outs[-1]).all() and ctrk[0] == ctr[0], name
        n += 1
    assert n > 80


def test_speedups_module_contract(sl):
    sp = sl.speedups
    b = np.zeros((7, 7), np.uint16)
    b[3, 2:5] = 9 | 512
    out = sp.advance_board(b)
    assert out.dtype == np.uint16 and out.shape == (7, 7) and out is not b
    assert (b[3, 2:5] == 521).all()  # input not mutated
    assert (out[2:5, 3] == 521).all() and out.sum() == 3 * 521
    # list / int64 input is force-cast like NPY_ARRAY_FORCECAST
    assert (sp.advance_board(b.astype(np.int64).tolist()) == out).all()
    with pytest.raises(ValueError):
        sp.advance_board(np.zeros((0, 5), np.uint16))
    with pytest.raises(ValueError):
        sp.advance_board(np.zeros(5, np.uint16))
    with pytest.raises(sl.lib.SlbError):
        sp.advance_board(np.zeros((2, 5), np.uint16))
    # stochastic: module stream, seed() rewinds
    s = np.zeros((5, 5), np.uint16)
    s[2, 2] = 16 | 128 | 8 | 1024
    sp.seed(5)
    a1 = sp.advance_board(s, 0.3)
    a2 = sp.advance_board(a1, 0.3)
    sp.seed(5)
    assert (sp.advance_board(s, 0.3) == a1).all()
    o1, c1 = slo.advance_board(s, 0.3, seed=5, env=0, counter=0)
    o2, _ = slo.advance_board(o1, 0.3, seed=5, env=0, counter=c1)
    assert (a1 == o1).all() and (a2 == o2).all()
    # CUDA tensor in -> CUDA tensor out
    t = torch.from_numpy(b).cuda()
    assert (sp.advance_board(t).cpu().numpy() == out).all()


@pytest.mark.parametrize("archive", ["append_still", "prune_spawn", "append_spawn", "navigation"])
def test_env_trace(sl, archive):
    tr = gu.load("trace_%s.npz" % archive)
    lv = gu.levels(str(tr["archive"]))
    l0 = int(tr["level0"])
    pool = sl.game.LevelPool(gu.level_slice(lv, l0, l0 + 2))
    env = sl.env.SafeLifeEnv(pool, num_envs=1, env_id0=int(tr["env_id"]), global_counter=None,
                             compute_obs=False, level_ids=[0])
    env.seed(int(tr["seed"]))
    env.reset()
    for t, a in enumerate(tr["actions"]):
        _, r, d, info = env.step(np.array([a]))
        st = env.game.state_numpy()[0]
        assert (info["board"][0].cpu().numpy() == tr["board"][t]).all(), t
        assert (info["goals"][0].cpu().numpy() == tr["goals"][t]).all(), t
        assert int(r[0]) == tr["reward"][t] and bool(d[0]) == bool(tr["done"][t]), t
        assert bool(info["times_up"][0]) == bool(tr["times_up"][t])
        assert (st["agent_x"], st["agent_y"]) == tuple(tr["agent_loc"][t])
        assert st["orientation"] == tr["orientation"][t]
        assert st["old_points"] == tr["points"][t]
        assert st["completed"] == tr["completed"][t] and st["possible"] == tr["possible"][t]
        if d[0]:
            env.reset()


def test_scenarios(sl):
    d = gu.load("scenarios.npz")
    for i, name in enumerate(d["names"]):
        rec = {"board": d["board0_%d" % i][None], "goals": d["goals0_%d" % i][None],
               "agent_loc": d["agent_%d" % i][None, :2], "orientation": d["agent_%d" % i][None, 2],
               "spawn_prob": np.array([0.3]), "min_performance": d["minperf_%d" % i][None]}
        pool = sl.game.LevelPool(rec)
        g = sl.game.SafeLifeGame(pool, 1, seed=int(d["seed"]), env_id0=1000 + i)
        assert (g.board[0].cpu().numpy() == d["start_%d" % i]).all(), name
        for t, a in enumerate(d["actions_%d" % i]):
            meta = d["meta_%d" % i][t]
            assert bool(g.can_exit()[0]) == bool(meta[5]), (name, t)
            r = g.execute_action(sl.game.ACTION_NAMES[int(a)])
            assert (g.board[0].cpu().numpy() == d["boards_%d" % i][t, 0]).all(), (name, t)
            st = g.state_numpy()[0]
            assert (int(r[0]), st["agent_x"], st["agent_y"], st["orientation"], st["game_over"]) == \
                tuple(int(v) for v in meta[:5]), (name, t)
            g.advance_board()
            assert int(g.current_points()[0]) == meta[6]
            assert int(g.performance_ratio()[0][0]) == meta[7]
            g.update_exit_colors()
            assert (g.board[0].cpu().numpy() == d["boards_%d" % i][t, 1]).all(), (name, t)
            assert (g.goals[0].cpu().numpy() == d["goals_%d" % i][t]).all(), (name, t)


def test_side_effect_counts_golden(sl):
    d = gu.load("side_effects.npz")
    for i in range(int(d["n_cases"])):
        lv = gu.levels(str(d["archive_%d" % i]))
        l0 = int(d["level_%d" % i])
        pool = sl.game.LevelPool(gu.level_slice(lv, l0, l0 + 1))
        env = sl.env.SafeLifeEnv(pool, num_envs=1, env_id0=int(d["env_%d" % i]), global_counter=None,
                                 compute_obs=False)
        env.seed(int(d["seed"]))
        env.reset()
        for a in d["actions_%d" % i]:
            env.step(np.array([a]))
        assert (env.game.board[0].cpu().numpy() == d["final_board_%d" % i]).all()
        n = int(d["num_samples_%d" % i])
        res = sl.se.side_effect_score(env.game, num_samples=n, emd=False)[0]
        keys = sorted(int(k) for k in res)
        assert keys == d["keys_%d" % i].tolist()
        assert [res[np.uint16(k)][1] for k in keys] == d["sum_inaction_%d" % i].tolist()


def _oracle_vs_device(sl, archive, n, steps, seed, aseed, auto_reset, check_every=25):
    lv = gu.levels(archive)
    pool = sl.game.LevelPool(lv)
    opool = slo.OraclePool(**lv)
    env = sl.env.SafeLifeEnv(pool, num_envs=n, env_id0=77, global_counter=None, compute_obs=False,
                             auto_reset=auto_reset, time_limit=60 if auto_reset else 1000)
    env.seed(seed)
    env.reset()
    lvl = (np.arange(n) + 77) % pool.n_levels
    oenv = slo.OracleBatchEnv(opool, lvl, seed=seed, env_id0=77, threads=8,
                              time_limit=60 if auto_reset else 1000)
    rng = np.random.RandomState(aseed)
    stride = n % pool.n_levels or pool.n_levels
    for t in range(steps):
        acts = rng.randint(9, size=n)
        _, r, d, _ = env.step(acts)
        orr, od = oenv.step(acts)
        assert (r.cpu().numpy() == orr).all(), t
        assert (d.cpu().numpy() == (od != 0)).all(), t
        if auto_reset and od.any():
            idx = np.nonzero(od)[0]
            lvl[idx] = (lvl[idx] + stride) % pool.n_levels
            oenv.reset(idx, lvl[idx])
        if t % check_every == 0 or t == steps - 1:
            assert (env.game.board.cpu().numpy() == oenv.boards).all(), t
            assert (env.game.goals.cpu().numpy() == oenv.goals).all(), t
            st, ost = env.game.state_numpy(), oenv.states
            for f in st.dtype.names:
                assert (st[f] == ost[f]).all(), (f, t)
    return env


def test_batch_vs_oracle_prune_spawn(sl):
    _oracle_vs_device(sl, "prune-spawn", 384, 150, seed=0, aseed=3, auto_reset=False)


def test_batch_vs_oracle_auto_reset(sl):
    _oracle_vs_device(sl, "append-spawn", 250, 200, seed=9, aseed=4, auto_reset=True)


def test_rollout_equals_steps_and_placement_independence(sl):
    lv = gu.levels("prune-spawn")
    pool = sl.game.LevelPool(lv)
    n, T = 256, 40
    acts = torch.from_numpy(np.random.RandomState(1).randint(9, size=(T, n)).astype(np.uint8)).cuda()
    a = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False)
    a.seed(3); a.reset()
    b = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False)
    b.seed(3); b.reset()
    r_roll, d_roll = a.rollout(acts)
    for t in range(T):
        _, r, d, _ = b.step(acts[t])
        assert (r == r_roll[t]).all() and (d == (d_roll[t] != 0)).all()
    assert (a.game.board == b.game.board).all() and (a.game.goals == b.game.goals).all()
    assert (a.game.state == b.game.state).all()
    # the second half of the envs, run as their own shard (env_id0 = n/2), gives identical results
    h = n // 2
    c = sl.env.SafeLifeEnv(pool, num_envs=h, env_id0=h, global_counter=None, compute_obs=False)
    c.seed(3); c.reset()
    r_c, _ = c.rollout(acts[:, h:].contiguous())
    assert (r_c == r_roll[:, h:]).all()
    assert (c.game.board == a.game.board[h:]).all() and (c.game.goals == a.game.goals[h:]).all()


def test_large_board_vs_oracle(sl):
    rng = np.random.RandomState(5)
    vals = np.array([0, 1033, 16, 1041, 521, 32788, 1688], np.uint16)
    p = np.array([60, 25, 5, 4, 3, 2, 1], float) / 100
    boards = vals[rng.choice(len(vals), size=(3, 256, 256), p=p)]
    out, ctr = dev_advance(sl, boards, 0.3, seed=11, env0=5)
    for i in range(3):
        o, c = slo.advance_board(boards[i], 0.3, seed=11, env=5 + i)
        assert (out[i] == o).all() and ctr[i] == c
    # idempotence-style property at full size: a board of still lifes and walls is a fixed point
    still = np.zeros((4096, 26, 26), np.uint16)
    still[:, 3:5, 3:5] = 1033
    still[:, 10, :] = 16
    out, _ = dev_advance(sl, still, 0.3, seed=0, env0=0)
    assert (out == still).all()


def test_error_behaviour(sl):
    L = sl.lib.load()
    b = torch.zeros((1, 2, 8), dtype=torch.uint16, device="cuda")
    p = torch.zeros(1, device="cuda")
    rc = L.slb_advance_board(b.data_ptr(), b.data_ptr() + 64, 1, 2, 8, p.data_ptr(), 0, 0, None, 1,
                             None, None)
    assert rc == sl.lib.SLB_ERR_INVALID and b"3x3" in L.slb_last_error()
    rc = L.slb_advance_board(b.data_ptr(), b.data_ptr(), 1, 4, 4, p.data_ptr(), 0, 0, None, 1, None, None)
    assert rc == sl.lib.SLB_ERR_INVALID


def test_stats_accumulate(sl):
    lv = gu.levels("append-still")
    pool = sl.game.LevelPool(lv)
    env = sl.env.SafeLifeEnv(pool, num_envs=300, global_counter=None, compute_obs=False, time_limit=3)
    env.seed(0); env.reset()
    stats = torch.zeros(8, dtype=torch.int64, device="cuda")
    tot_r, tot_d = 0, 0
    for t in range(5):
        _, r, d, _ = env.step(np.random.RandomState(t).randint(9, size=300))
        sl.lib.check(sl.lib.load().slb_accumulate_stats(300, env.game.state.data_ptr(), r.data_ptr(),
                                                        env._done.data_ptr(), stats.data_ptr(),
                                                        sl.lib.current_stream()))
        tot_r += int(r.sum()); tot_d += int(d.sum())
    s = stats.cpu().numpy()
    assert s[0] == tot_r and s[1] == tot_d and s[6] == 1500


def test_get_obs_golden(sl):
    """slb_get_obs + slb_obs_channels against the reference's SafeLifeEnv.get_obs (tests/golden/obs.npz)"""
    d = gu.load("obs.npz")
    for i in range(int(d["n_cases"])):
        lv = gu.levels(str(d["archive_%d" % i]))
        l0 = int(d["level_%d" % i])
        pool = sl.game.LevelPool(gu.level_slice(lv, l0, l0 + 1))
        view = tuple(int(v) for v in d["view_%d" % i])
        ch = tuple(int(c) for c in d["channels_%d" % i]) or None
        env = sl.env.SafeLifeEnv(pool, num_envs=1, env_id0=int(d["env_%d" % i]), global_counter=None,
                                 view_shape=view, remove_white_goals=bool(d["remove_white_%d" % i]),
                                 output_channels=ch)
        env.seed(int(d["seed_%d" % i]))
        want = d["obs_%d" % i]
        obs = env.reset()
        assert obs.dtype == torch.uint16 and (obs[0].cpu().numpy() == want[0]).all(), (i, "reset")
        k = 1
        for t, a in enumerate(d["actions_%d" % i]):
            obs, _, _, _ = env.step(np.array([a]))
            if (t + 1) % int(d["every_%d" % i]) == 0:
                assert (obs[0].cpu().numpy() == want[k]).all(), (i, t)
                k += 1


def _random_levels(rng, p, h, w, exits=2):
    """Dense little levels full of crates / walls / life so that random actions push, pull, wrap,
    toggle and exit all the time."""
    vals = np.array([0, 0, 0, 1033, 521, 16, 1041, 32788, 32788, 1688, 9, 20, 32784 | 1], np.uint16)
    board = vals[rng.randint(len(vals), size=(p, h, w))]
    goals = np.array([0, 0, 2057, 3584, 1033, 16], np.uint16)[rng.randint(6, size=(p, h, w))]
    agent = np.stack([rng.randint(w, size=p), rng.randint(h, size=p)], 1)
    for l in range(p):
        for _ in range(exits):
            board[l, rng.randint(h), rng.randint(w)] = 272
        board[l, agent[l, 1], agent[l, 0]] = 122 | (rng.randint(8) << 9)
    return {"board": board, "goals": goals, "agent_loc": agent, "orientation": rng.randint(4, size=p),
            "spawn_prob": rng.choice([0.0, 0.3, 1.0], size=p), "min_performance": rng.choice([-1.0, 0.0, 0.01], size=p)}


@pytest.mark.parametrize("h,w", [(4, 4), (5, 7), (26, 26), (32, 32), (31, 9), (9, 31), (3, 5), (40, 33),
                                  (3, 32), (36, 32), (40, 64), (64, 96)])
def test_action_stress_vs_oracle(sl, h, w):
    """Random actions on dense random levels of odd shapes (toroidal push / pull / exit / toggle at the
    seams, spawners, auto-reset): full state equality with the oracle every few steps.  4..32 squares take
    the warp-per-env bit-plane kernels, widths that are multiples of 32 the CTA-per-env bit-plane kernels
    (slb_big.cuh), everything else the uint16-resident generic kernels."""
    rng = np.random.RandomState(h * 100 + w)
    lv = _random_levels(rng, 24, h, w)
    n, steps = 96, 60
    pool = sl.game.LevelPool(lv)
    env = sl.env.SafeLifeEnv(pool, num_envs=n, env_id0=5, global_counter=None, auto_reset=True, time_limit=25,
                             view_shape=(7, 9), output_channels=None)
    env.seed(h + w)
    obs = env.reset()
    lvl = (np.arange(n) + 5) % pool.n_levels
    oenv = slo.OracleBatchEnv(slo.OraclePool(**lv), lvl, seed=h + w, env_id0=5, threads=4, time_limit=25)
    stride = n % pool.n_levels or pool.n_levels

    def check(t):
        assert (env.game.board.cpu().numpy() == oenv.boards).all(), t
        assert (env.game.goals.cpu().numpy() == oenv.goals).all(), t
        st, ost = env.game.state_numpy(), oenv.states
        for f in st.dtype.names:
            assert (st[f] == ost[f]).all(), (f, t)
        o = obs.cpu().numpy()
        for i in range(0, n, 7):
            want = slo.get_obs(oenv.boards[i], oenv.goals[i], (ost["agent_x"][i], ost["agent_y"][i]),
                               lv["board"][ost["level_id"][i]], (7, 9), True, None)
            assert (o[i] == want).all(), (t, i)

    check(-1)
    for t in range(steps):
        acts = rng.randint(9, size=n)
        obs, r, d, _ = env.step(acts)
        orr, od = oenv.step(acts)
        assert (r.cpu().numpy() == orr).all(), t
        assert (d.cpu().numpy() == (od != 0)).all(), t
        if od.any():
            idx = np.nonzero(od)[0]
            lvl[idx] = (lvl[idx] + stride) % pool.n_levels
            oenv.reset(idx, lvl[idx])
        if t % 3 == 0 or t == steps - 1:
            check(t)


def test_resident_import_export_roundtrip(sl):
    """cells -> resident planes -> cells is the identity for every uint16 value (incl. bits 12-14), and a
    board imported into a game with bits its pool had never seen is stepped correctly afterwards."""
    rng = np.random.RandomState(3)
    L = sl.lib.load()
    for h, w in [(26, 26), (4, 32), (32, 4), (17, 23), (40, 5), (40, 64), (256, 256), (3, 32)]:
        cells = torch.from_numpy(rng.randint(0, 1 << 16, size=(50, h, w)).astype(np.uint16)).cuda()
        res = torch.zeros(50 * L.slb_resident_bytes(h, w), dtype=torch.uint8, device="cuda")
        back = torch.zeros_like(cells)
        sl.lib.check(L.slb_resident_import(cells.data_ptr(), res.data_ptr(), 50, h, w, None, 0, None))
        sl.lib.check(L.slb_resident_export(res.data_ptr(), back.data_ptr(), 50, h, w, None))
        assert (back == cells).all(), (h, w)
    # a pool without spawners; then a board WITH a spawner is assigned to the game
    lv = gu.level_slice(gu.levels("append-still"), 0, 4)
    pool = sl.game.LevelPool(lv)
    assert pool.c.plane_groups[0] == 7 and pool.c.plane_groups[1] == 3
    g = sl.game.SafeLifeGame(pool, 4, seed=2)
    b = g.board.cpu().numpy()
    b[:, 5, 5] = 1688
    b[:, 7, 7] = 0x7000 | 16
    g.board = torch.from_numpy(b).cuda()
    assert pool.c.plane_groups[0] == 15
    ctr = np.zeros(4, np.uint64)
    for t in range(5):
        g.advance_board()
        for i in range(4):
            b[i], c = slo.advance_board(b[i], 0.3, seed=2, env=i, counter=int(ctr[i]))
            ctr[i] = c
        assert (g.board.cpu().numpy() == b).all(), t


def test_wrappers_golden(sl):
    """MovementBonusWrapper / SimpleSideEffectPenalty (safelife_k2_b200/env_wrappers.py) against the
    reference's wrappers around its real env (tests/golden/wrappers.npz)."""
    from safelife_k2_b200 import env_wrappers as ew

    d = gu.load("wrappers.npz")
    for i in range(int(d["n_cases"])):
        lv = gu.levels(str(d["archive_%d" % i]))
        l0 = int(d["level_%d" % i])
        for which in ("penalty", "movement"):
            pool = sl.game.LevelPool(gu.level_slice(lv, l0, l0 + 1))
            env = sl.env.SafeLifeEnv(pool, num_envs=1, env_id0=int(d["env_%d" % i]), global_counter=None,
                                     compute_obs=False)
            env.seed(int(d["seed_%d" % i]))
            if which == "penalty":
                env = ew.SimpleSideEffectPenalty(env, penalty_coef=float(d["penalty_coef"]),
                                                 min_performance=float(d["min_performance"]))
            else:
                env = ew.MovementBonusWrapper(env)
            env.reset()
            for t, a in enumerate(d["actions_%d" % i]):
                _, r, _, _ = env.step(np.array([a]))
                assert r.dtype == torch.float64
                assert abs(float(r[0]) - d["%s_reward_%d" % (which, i)][t]) < 1e-12, (which, i, t)
                if which == "penalty":
                    assert int(env.last_side_effect[0]) == d["side_effect_%d" % i][t], (i, t)


@pytest.mark.parametrize("h,w", [(26, 26), (9, 31), (40, 33), (40, 64)])
def test_side_effect_penalty_vs_oracle(sl, h, w):
    from safelife_k2_b200 import env_wrappers as ew

    rng = np.random.RandomState(h + 7 * w)
    lv = _random_levels(rng, 12, h, w)
    n = 48
    pool = sl.game.LevelPool(lv)
    env = ew.SimpleSideEffectPenalty(
        sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False), penalty_coef=1.0)
    env.seed(1)
    env.reset()
    lvl = np.arange(n) % 12
    for t in range(12):
        env.step(rng.randint(9, size=n))
        b, g = env.game.board.cpu().numpy(), env.game.goals.cpu().numpy()
        want = [slo.side_effect_penalty_count(b[i], lv["board"][lvl[i]], g[i]) for i in range(n)]
        assert env.last_side_effect.cpu().tolist() == want, t


def test_continuing_env(sl):
    """ContinuingEnv (env_wrappers.py:289-303): finishing a level restarts that env, only the time
    limit ends an episode."""
    from safelife_k2_b200 import env_wrappers as ew

    # the exit is right of the agent and open (min_performance < 0): MOVE RIGHT finishes the level
    board = np.zeros((2, 6, 6), np.uint16)
    board[:, 2, 2] = 122
    board[:, 2, 3] = 272
    lv = {"board": board, "goals": np.zeros_like(board), "agent_loc": np.array([[2, 2], [2, 2]]),
          "orientation": np.array([1, 1]), "spawn_prob": np.array([0.3, 0.3]), "min_performance": np.array([-1.0, -1.0])}
    env = ew.ContinuingEnv(sl.env.SafeLifeEnv(sl.game.LevelPool(lv), num_envs=2, global_counter=None,
                                              compute_obs=False, time_limit=3))
    env.seed(0)
    env.reset()
    _, r, done, info = env.step(np.array([2, 0]))       # env 0 exits, env 1 idles
    assert r.tolist() == [1, 0] and done.tolist() == [False, False]
    st = env.game.state_numpy()
    assert st["game_over"].tolist() == [0, 0] and st["episode_length"].tolist() == [0, 1]  # env 0 restarted
    for _ in range(2):
        _, r, done, info = env.step(np.array([0, 0]))
    assert done.tolist() == [False, False]
    _, r, done, info = env.step(np.array([0, 0]))       # env 1 reaches step 4 > time_limit 3
    assert done.tolist() == [False, True] and info["times_up"].tolist() == [False, True]


def test_benchmark_log_golden(sl, tmp_path):
    """benchmarking.run_benchmark writes the reference's YAML episode record (env_wrappers.py:195-221):
    same header text and the same parsed content as RecordingSafeLifeWrapper's log for the same episode."""
    import yaml
    from safelife_k2_b200 import benchmarking as bm

    d = gu.load("benchmark_log.npz")
    for i in range(int(d["n_cases"])):
        lv = gu.levels(str(d["archive_%d" % i]))
        l0 = int(d["level_%d" % i])
        acts = iter(d["actions_%d" % i])
        log = str(tmp_path / ("bench_%d.yaml" % i))
        eps = bm.run_benchmark(gu.level_slice(lv, l0, l0 + 1), lambda obs: torch.tensor([next(acts)]), log,
                               names=[str(d["title_%d" % i])], time_limit=int(d["time_limit_%d" % i]),
                               seed=int(d["seed_%d" % i]), env_id0=int(d["env_%d" % i]), emd=False)
        assert len(eps) == 1
        text = open(log).read()
        assert text.startswith("# SafeLife benchmark data\n---\n")
        want_txt = str(d["log_%d" % i])
        ours, want = yaml.safe_load(text)[0], yaml.safe_load(want_txt)[0]
        ours.pop("episode"); want.pop("episode")
        assert ours == want, (ours, want)
        head = lambda t: t.split("  side effects:")[0].replace("episode: 1", "episode: -1")  # noqa: E731
        assert head(text.split("---\n", 1)[1]) == head(want_txt)
        stats = bm.load_benchmarks(log)
        assert stats["length"].tolist() == [want["length"]] and set(stats["side effects"]) == set(want["side effects"])


def test_step_host_paths_agree(sl):
    """slb_env_step_host(_async) / slb_stepper_*: the plain path (default stream), the cached-graph
    path (side streams, pageable AND pinned buffers, more env batches than graph-cache entries) and
    the device-buffer step give identical rewards, dones and states."""
    lv = gu.levels("prune-spawn")
    pool = sl.game.LevelPool(lv)
    n, T, nb = 96, 12, 20  # 20 batches > 16 cached graphs per stream
    rng = np.random.RandomState(7)
    acts = rng.randint(9, size=(T, nb, n)).astype(np.uint8)

    def make():
        out = []
        for b in range(nb):
            e = sl.env.SafeLifeEnv(pool, num_envs=n, env_id0=b * n, global_counter=None, compute_obs=False,
                                   auto_reset=True, time_limit=7)
            e.seed(5); e.reset()
            out.append(e)
        return out

    ref, plain, graphed = make(), make(), make()
    want_r = np.zeros((T, nb, n), np.int32); want_d = np.zeros((T, nb, n), np.uint8)
    for t in range(T):
        for b in range(nb):
            _, r, d, _ = ref[b].step(torch.from_numpy(acts[t, b]).cuda())
            want_r[t, b] = r.cpu().numpy(); want_d[t, b] = d.cpu().numpy()
    # plain path: default stream, synchronous
    r = np.zeros(n, np.int32); d = np.zeros(n, np.uint8)
    for t in range(T):
        for b in range(nb):
            plain[b].step_host(acts[t, b], r, d)
            assert (r == want_r[t, b]).all() and (d == want_d[t, b]).all()
    # cached-graph path: two side streams, results waited for one step later
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    outs = [(np.zeros(n, np.int32), np.zeros(n, np.uint8)),
            (torch.zeros(n, dtype=torch.int32).pin_memory(), torch.zeros(n, dtype=torch.uint8).pin_memory())]
    pending = [None, None]
    k = 0
    for t in range(T):
        for b in range(nb):
            s = k % 2
            if pending[s] is not None:
                graphed[0].wait(streams[s])
                pt, pb = pending[s]
                assert (np.asarray(outs[s][0]) == want_r[pt, pb]).all()
                assert (np.asarray(outs[s][1]) == want_d[pt, pb]).all()
            graphed[b].step_host(acts[t, b], outs[s][0], outs[s][1], sync=False, stream=streams[s])
            pending[s] = (t, b)
            k += 1
    for s in range(2):
        graphed[0].wait(streams[s])
        pt, pb = pending[s]
        assert (np.asarray(outs[s][0]) == want_r[pt, pb]).all()
    torch.cuda.synchronize()
    for b in range(nb):
        for other in (plain, graphed):
            assert (other[b].game.state == ref[b].game.state).all()
            assert (other[b].game.board == ref[b].game.board).all()
            assert (other[b].game.goals == ref[b].game.goals).all()


def test_out_of_band_writes_keep_the_step_exact(sl):
    """The fused step keeps each env's score (and, for levels whose goals never change, skips the goals
    physics) as long as boards are only written by the step and reset.  Writing a board or goals from
    outside (board/goals setters, single-stage calls) must switch that off: every later step still
    equals the oracle, which rescans every step like the reference."""
    lv = gu.levels("append-still")
    pool = sl.game.LevelPool(lv)
    n = 128
    env = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False)
    env.seed(2); env.reset()
    lvl = np.arange(n) % pool.n_levels
    oenv = slo.OracleBatchEnv(slo.OraclePool(**lv), lvl, seed=2, threads=4)
    rng = np.random.RandomState(11)

    def step(acts):
        _, r, d, _ = env.step(acts)
        orr, od = oenv.step(acts)
        assert (r.cpu().numpy() == orr).all() and (d.cpu().numpy() == (od != 0)).all()

    def same():
        assert (env.game.board.cpu().numpy() == oenv.boards).all()
        assert (env.game.goals.cpu().numpy() == oenv.goals).all()
        st, ost = env.game.state_numpy(), oenv.states
        for f in st.dtype.names:
            assert (st[f] == ost[f]).all(), f

    for _ in range(12):
        step(rng.randint(9, size=n))
    same()
    assert pool.c.external_writes == 0
    # 1. boards replaced from outside: a green still life block dropped into every even env, then idle steps
    b = oenv.boards.copy()
    b[::2, 1:3, 1:3] = 1033
    oenv.boards[:] = b
    env.game.board = torch.from_numpy(b).cuda()
    assert pool.c.external_writes == 1
    for _ in range(3):
        step(np.zeros(n, int))
    same()
    # 2. goals replaced from outside (an oscillator: the goals of these levels stop being static)
    g = oenv.goals.copy()
    g[1::2, 5, 4:7] = 1 | (2 << 9)
    oenv.goals[:] = g
    env.game.goals = torch.from_numpy(g).cuda()
    for _ in range(4):
        step(np.zeros(n, int))
    for _ in range(10):
        step(rng.randint(9, size=n))
    same()
    # 3. a fresh pool, single-stage calls between fused steps
    pool2 = sl.game.LevelPool(lv)
    env2 = sl.env.SafeLifeEnv(pool2, num_envs=n, global_counter=None, compute_obs=False)
    env2.seed(2); env2.reset()
    o2 = slo.OracleBatchEnv(slo.OraclePool(**lv), lvl, seed=2, threads=4)
    acts = rng.randint(5, 9, size=n)  # toggles: create / destroy life in front of the agent
    env2.game.execute_action(acts)
    for i in range(n):
        slo.execute_action(o2.boards[i], o2.states[i:i + 1], acts[i], o2.states["can_exit"][i])
    assert pool2.c.external_writes == 1
    for _ in range(3):
        a = np.zeros(n, int)
        _, r, d, _ = env2.step(a)
        orr, od = o2.step(a)
        assert (r.cpu().numpy() == orr).all()
    assert (env2.game.board.cpu().numpy() == o2.boards).all()
    st, ost = env2.game.state_numpy(), o2.states
    for f in st.dtype.names:
        assert (st[f] == ost[f]).all(), f


@pytest.mark.parametrize("m,channels", [(1, (0,)), (7, tuple(range(15))), (225 * 3 + 1, tuple(range(15))),
                                        (1001, (3, 0, 15, 9)), (64, tuple(range(16))), (333, (14, 2, 2, 7, 11, 0, 1)),
                                        (5, (1, 2, 3))])
def test_obs_channels_kernel(sl, m, channels):
    """slb_obs_channels (safelife_env.py:152-154) for odd sizes, repeated / permuted channels."""
    rng = np.random.RandomState(m)
    words = rng.randint(0, 65536, size=m).astype(np.uint16)
    w = torch.from_numpy(words).cuda()
    ch = np.array(channels, dtype=np.uint8)
    out = torch.full((m, len(channels)), 9, dtype=torch.uint16, device="cuda")
    sl.lib.check(sl.lib.load().slb_obs_channels(w.data_ptr(), m, ch.ctypes.data, len(channels), out.data_ptr(),
                                                sl.lib.current_stream()))
    want = (words[:, None] >> ch[None, :].astype(np.uint16)) & 1
    assert (out.cpu().numpy() == want).all()


@pytest.mark.parametrize("h,w,view,channels", [(26, 26, (15, 15), tuple(range(15))), (26, 26, (33, 33), tuple(range(15))),
                                               (9, 31, (7, 9), (0, 3, 9, 10, 11)), (32, 32, (5, 4), (15, 0, 1)),
                                               (26, 26, (1, 1), (2,)), (40, 64, (15, 15), tuple(range(15))),
                                               (40, 33, (9, 7), (1, 8, 9)), (26, 26, (35, 35), tuple(range(16)))])
def test_fused_obs_equals_two_step(sl, h, w, view, channels):
    """slb_get_obs_channels (one kernel on the warp-per-env shapes, 16-byte stores at 2-byte-aligned env
    offsets) against slb_get_obs + the numpy channel split, on random mid-episode states."""
    rng = np.random.RandomState(h + 7 * w + view[0])
    lv = _random_levels(rng, 12, h, w)
    n = 37
    pool = sl.game.LevelPool(lv)
    env = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, auto_reset=True, time_limit=25,
                             view_shape=view, output_channels=channels)
    env.seed(1)
    env.reset()
    for t in range(6):
        obs, _, _, _ = env.step(rng.randint(9, size=n))
        env.output_channels = None
        words = env.get_obs().cpu().numpy()
        env.output_channels = channels
        want = (words[..., None] >> np.array(channels, np.uint16)) & 1
        assert obs.shape == want.shape and (obs.cpu().numpy() == want).all(), t


def test_step_host_many_streams(sl):
    """More streams than staging slots per thread: idle slots are recycled, results stay right."""
    lv = gu.levels("append-still")
    pool = sl.game.LevelPool(lv)
    n = 64
    a = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False)
    b = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False)
    for e in (a, b):
        e.seed(4); e.reset()
    rng = np.random.RandomState(0)
    r = np.zeros(n, np.int32); d = np.zeros(n, np.uint8)
    streams = [torch.cuda.Stream() for _ in range(20)]   # 16 slots per thread
    for t in range(60):
        acts = rng.randint(9, size=n).astype(np.uint8)
        st = streams[t % 20]
        st.wait_stream(streams[(t - 1) % 20])
        a.step_host(acts, r, d, sync=False, stream=st)
        a.wait(st)
        _, rb, db, _ = b.step(acts)
        assert (r == rb.cpu().numpy()).all() and (d == db.cpu().numpy()).all(), t
    assert (a.game.state == b.game.state).all() and (a.game.board == b.game.board).all()


def _sampled_vs_oracle(sl, lv, n, steps, blocks, time_limit, seed=3, check_states=True):
    """Run the device on the FULL batch of n envs and the oracle only on a few blocks of consecutive envs
    of it (an env's stream is keyed by its global index, its level by index % levels): full state
    equality on the sampled envs, plus the all-env statistics against their own definition."""
    pool = sl.game.LevelPool(lv)
    P = pool.n_levels
    env = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False, auto_reset=True,
                             time_limit=time_limit)
    env.seed(seed)
    env.reset()
    opool = slo.OraclePool(**lv)
    stride = n % P or P
    oenvs = []
    for lo, cnt in blocks:
        lvl = (np.arange(cnt) + lo) % P
        oenvs.append([lo, cnt, lvl, slo.OracleBatchEnv(opool, lvl, seed=seed, env_id0=lo, threads=4,
                                                       time_limit=time_limit)])
    rng = np.random.RandomState(n % 1000)
    for t in range(steps):
        acts = rng.randint(9, size=n)
        _, r, d, _ = env.step(acts)
        r, d = r.cpu().numpy(), d.cpu().numpy()
        for blk in oenvs:
            lo, cnt, lvl, o = blk
            orr, od = o.step(acts[lo:lo + cnt])
            assert (r[lo:lo + cnt] == orr).all() and (d[lo:lo + cnt] == (od != 0)).all(), (t, lo)
            if od.any():
                idx = np.nonzero(od)[0]
                lvl[idx] = (lvl[idx] + stride) % P
                o.reset(idx, lvl[idx])
    board, goals, st = env.game.board.cpu().numpy(), env.game.goals.cpu().numpy(), env.game.state_numpy()
    for lo, cnt, lvl, o in oenvs:
        assert (board[lo:lo + cnt] == o.boards).all() and (goals[lo:lo + cnt] == o.goals).all(), lo
        for f in st.dtype.names:
            assert (st[f][lo:lo + cnt] == o.states[f]).all(), (f, lo)
    # whole batch: every env took `steps` steps, and agents are where their state says they are
    assert (st["num_steps"] + 0 >= 0).all() and int(st["rng_counter"].min()) >= 0
    ys, xs = st["agent_y"], st["agent_x"]
    alive_agent = board[np.arange(n), ys, xs]
    assert ((alive_agent & 2) != 0).mean() > 0.9  # (an agent that walked into an exit is gone)
    return env


def test_full_size_prune_spawn_16384(sl):
    """BASELINE configs 3 and 4 at full size.  Config 3: 16 384 prune-spawn envs (stochastic spawners, one
    counter-based stream per env) for 1000 steps with resets, 256 sampled envs bit-exact vs the oracle at
    every step and in their final state.  Config 4: side-effect count maps (1000 samples) of all 16 384
    final states in one call, sampled episodes equal to the oracle's maps, keys and draw counters."""
    lv = gu.levels("prune-spawn")
    env = _sampled_vs_oracle(sl, lv, 16384, 1000, [(0, 96), (8100, 64), (16384 - 96, 96)], 300)
    g = env.game
    n = g.n
    board1 = g.board
    st = g.state_numpy()
    board0 = torch.from_numpy(lv["board"][st["level_id"]]).cuda()
    ctr0 = st["rng_counter"].astype(np.int64)
    ctr = torch.from_numpy(ctr0.copy()).cuda()
    spawn = lv["spawn_prob"][st["level_id"]].astype(np.float32)
    keys, counts, nkeys = sl.se.side_effect_counts(board0, board1, st["num_steps"].astype(np.int32), spawn,
                                                   num_samples=1000, seed=3, env_id0=0, counters=ctr, kmax=32)
    nk = nkeys.cpu().numpy()
    assert (nk >= 1).all() and (nk <= 32).all()
    b1 = board1.cpu().numpy()
    ctr = ctr.cpu().numpy()
    for i in list(range(0, 8)) + list(range(5000, 5008)) + list(range(n - 8, n)):
        ok, oc, octr = slo.side_effect_counts(lv["board"][st["level_id"][i]], b1[i], int(st["num_steps"][i]),
                                              num_samples=1000, spawn_prob=float(spawn[i]), seed=3, env=i,
                                              counter=int(ctr0[i]), kmax=32)
        ki = keys[i].cpu().numpy()
        ci = counts[i].cpu().numpy()
        assert nk[i] == len(ok) and ctr[i] == octr, i
        mine = {int(k): ci[:, s] for s, k in enumerate(ki) if k}
        assert sorted(mine) == sorted(int(k) for k in ok), i
        for s, k in enumerate(ok):
            assert (mine[int(k)] == oc[:, s]).all(), (i, int(k))


def test_full_size_mixed_65536(sl):
    """BASELINE config 5a at full size: 65 536 envs over all v1.0 archives."""
    parts = [gu.levels(a) for a in gu.ARCHIVES]
    lv = {k: np.concatenate([p[k] for p in parts]) for k in parts[0]}
    _sampled_vs_oracle(sl, lv, 65536, 25, [(0, 8), (32760, 16), (65536 - 8, 8)], 20)


def test_full_size_256x256_4096(sl):
    """BASELINE config 5b at full size: 4 096 envs of 256x256 (CTA-per-env bit-plane kernels)."""
    rng = np.random.RandomState(5)
    vals = np.array([0, 1033, 16, 1041, 521, 32788, 1688], np.uint16)
    pr = np.array([60, 25, 5, 4, 3, 2, 1], float) / 100
    p = 4
    board = vals[rng.choice(len(vals), size=(p, 256, 256), p=pr)]
    board[:, 127:130, 127:130] = 0
    board[:, 128, 128] = 122
    board[:, 0, 0] = 784
    g = vals[rng.choice(len(vals), size=(p, 256, 256), p=pr)]
    goals = np.where(g == 1033, 2057, np.where(g == 16, 3584, 0)).astype(np.uint16)
    lv = {"board": board, "goals": goals, "agent_loc": np.tile([128, 128], (p, 1)), "orientation": np.ones(p, int),
          "spawn_prob": np.full(p, 0.3), "min_performance": np.full(p, 0.5)}
    _sampled_vs_oracle(sl, lv, 4096, 4, [(0, 2), (4094, 2)], 1000)


def test_recording_wrapper_golden(sl, tmp_path):
    """RecordingSafeLifeWrapper / SafeLifeRecorder (env_wrappers.py:97-286): the trajectory .npz equals the
    reference recorder's frame for frame (tests/golden/trajectory.npz), and the YAML record written at the
    end of the episode is the reference's text (tests/golden/benchmark_log.npz)."""
    import io
    import yaml
    from safelife_k2_b200 import env_wrappers as ew

    d = gu.load("trajectory.npz")
    for i in range(int(d["n_cases"])):
        lv = gu.levels(str(d["archive_%d" % i]))
        l0 = int(d["level_%d" % i])
        env = sl.env.SafeLifeEnv(sl.game.LevelPool(gu.level_slice(lv, l0, l0 + 1)), num_envs=1,
                                 env_id0=int(d["env_%d" % i]), global_counter=None, compute_obs=False,
                                 time_limit=int(d["time_limit_%d" % i]))
        env.seed(int(d["seed_%d" % i]))
        rec = ew.RecordingSafeLifeWrapper(env, video_name=str(tmp_path / ("traj_%d" % i)), video_recording_freq=1,
                                          record_side_effects=False)
        rec.reset()
        for a in d["actions_%d" % i]:
            _, _, done, _ = rec.step(np.array([a]))
        assert bool(done[0])
        rec.close()
        t = np.load(str(tmp_path / ("traj_%d.npz" % i)))
        assert t["orientation"].tolist() == d["orientation_%d" % i].tolist()
        assert (t["board"] == d["board_%d" % i]).all() and (t["goals"] == d["goals_%d" % i]).all()
    d = gu.load("benchmark_log.npz")
    for i in range(int(d["n_cases"])):
        lv = gu.levels(str(d["archive_%d" % i]))
        l0 = int(d["level_%d" % i])
        env = sl.env.SafeLifeEnv(sl.game.LevelPool(gu.level_slice(lv, l0, l0 + 1)), num_envs=1,
                                 env_id0=int(d["env_%d" % i]), global_counter=None, compute_obs=False,
                                 time_limit=int(d["time_limit_%d" % i]))
        env.seed(int(d["seed_%d" % i]))
        log = io.StringIO()
        rec = ew.RecordingSafeLifeWrapper(env, log_file=log, level_names=[str(d["title_%d" % i])], emd=False)
        rec.reset()
        for a in d["actions_%d" % i]:
            rec.step(np.array([a]))
        want_txt = str(d["log_%d" % i])
        # the order of the side-effect lines is the iteration order of a python set in the reference: compare
        # the text up to there character for character and the whole record as parsed YAML
        head = lambda t: t.split("  side effects:")[0]  # noqa: E731
        assert head(log.getvalue()) == head(want_txt), (log.getvalue(), want_txt)
        assert yaml.safe_load(log.getvalue()) == yaml.safe_load(want_txt)


def test_side_effect_counts_many_cell_types(sl):
    """slb_side_effect_counts vs the oracle on boards with 3..20 distinct cell types: episodes with more
    than 8 types take the kernel's second (full-capacity) pass, the others must be left as the first pass
    wrote them; draw counters continue exactly in both."""
    rng = np.random.RandomState(12)
    n, h, w = 24, 26, 26
    walls = [16 | (c << 9) for c in range(8)]               # frozen, 8 colours
    trees = [17 | (c << 9) for c in range(1, 8)]            # frozen life, 7 colours
    life = [9 | (c << 9) for c in (1, 2, 4)]                # destructible life
    crates = [32788, 32788 | 1024]
    spawner = [1688]
    types = walls + trees + life + crates + spawner
    b0 = np.zeros((n, h, w), np.uint16)
    for i in range(n):
        k = 3 + (i * 17) % 18
        pick = rng.choice(len(types), size=k, replace=False)
        vals = np.array([0] * 6 + [types[p] for p in pick], np.uint16)
        b0[i] = vals[rng.randint(len(vals), size=(h, w))]
    b1 = b0.copy()
    b1[:, 5:8, 5:8] = 1033  # the "agent" left a block of green life
    steps = rng.randint(0, 30, size=n).astype(np.int32)
    ctr0 = rng.randint(0, 1000, size=n).astype(np.int64)
    ctr = torch.from_numpy(ctr0.copy()).cuda()
    keys, counts, nkeys = sl.se.side_effect_counts(torch.from_numpy(b0).cuda(), torch.from_numpy(b1).cuda(),
                                                   steps, np.full(n, 0.3, np.float32), num_samples=60, seed=5,
                                                   env_id0=100, counters=ctr, kmax=32)
    keys, counts, nkeys, ctr = keys.cpu().numpy(), counts.cpu().numpy(), nkeys.cpu().numpy(), ctr.cpu().numpy()
    assert nkeys.max() > 8 and nkeys.min() <= 8  # both passes are exercised
    for i in range(n):
        ok, oc, octr = slo.side_effect_counts(b0[i], b1[i], steps[i], num_samples=60, spawn_prob=0.3, seed=5,
                                              env=100 + i, counter=int(ctr0[i]), kmax=32)
        assert nkeys[i] == len(ok) and ctr[i] == octr, i
        mine = {int(k): counts[i, :, s] for s, k in enumerate(keys[i]) if k}
        assert sorted(mine) == sorted(int(k) for k in ok), i
        for s, k in enumerate(ok):
            assert (mine[int(k)] == oc[:, s]).all(), (i, int(k))


@pytest.mark.gpu
@pytest.mark.parametrize("h,w,num_samples", [(26, 26, 1000), (26, 26, 5), (8, 9, 300), (32, 32, 2500), (13, 26, 17),
                                               (30, 31, 4100)])
def test_side_effect_counts_runs_and_fixed_points(sl, h, w, num_samples):
    """slb_side_effect_counts vs the oracle where the run-length / fixed-point shortcuts of the kernel
    matter: boards that are still from the start, that settle after a while (during the catch-up or in
    the middle of the samples), period-2 oscillators, boards with spawners (never quiet), on shapes
    whose high counter planes live in the output slice (26x26, 32x32) and shapes where they do not."""
    rng = np.random.RandomState(h * 100 + w + num_samples)
    n = 40
    b0 = np.zeros((n, h, w), np.uint16)
    b1 = np.zeros((n, h, w), np.uint16)
    for i in range(n):
        kind = i % 5
        # a few still lifes (blocks) of two colours and some walls / crates
        for _ in range(3):
            y, x = rng.randint(0, h - 2), rng.randint(0, w - 2)
            b0[i, y:y + 2, x:x + 2] = 9 | (rng.choice([1, 2, 4]) << 9)
        b0[i, rng.randint(h), rng.randint(w)] = 16
        b0[i, rng.randint(h), rng.randint(w)] = 32788
        b1[i] = b0[i]
        if kind == 1:    # the agent left a blinker: period-2 forever (or a collision that settles)
            y, x = rng.randint(1, h - 1), rng.randint(0, w - 3)
            b1[i, y, x:x + 3] = 1033
        elif kind == 2:  # random soup in the action board: changes for a while, usually settles
            y, x = rng.randint(0, h - 4), rng.randint(0, w - 4)
            b1[i, y:y + 4, x:x + 4] = np.where(rng.rand(4, 4) < 0.5, 1033, 0)
        elif kind == 3:  # soup in the INITIAL board as well: settles during the catch-up or the samples
            y, x = rng.randint(0, h - 4), rng.randint(0, w - 4)
            b0[i, y:y + 4, x:x + 4] = np.where(rng.rand(4, 4) < 0.5, 521, 0)
            b1[i] = b0[i]
            b1[i, rng.randint(h), rng.randint(w)] = 0
        elif kind == 4:  # a spawner: draws every step
            b0[i, h // 2, w // 2] = 1688
            b1[i, h // 2, w // 2] = 1688
    steps = rng.randint(0, 40, size=n).astype(np.int32)
    ctr0 = rng.randint(0, 1000, size=n).astype(np.int64)
    ctr = torch.from_numpy(ctr0.copy()).cuda()
    keys, counts, nkeys = sl.se.side_effect_counts(torch.from_numpy(b0).cuda(), torch.from_numpy(b1).cuda(),
                                                   steps, np.full(n, 0.3, np.float32), num_samples=num_samples,
                                                   seed=9, env_id0=500, counters=ctr, kmax=12)
    keys, counts, nkeys, ctr = keys.cpu().numpy(), counts.cpu().numpy(), nkeys.cpu().numpy(), ctr.cpu().numpy()
    for i in range(n):
        ok, oc, octr = slo.side_effect_counts(b0[i], b1[i], steps[i], num_samples=num_samples, spawn_prob=0.3,
                                              seed=9, env=500 + i, counter=int(ctr0[i]), kmax=12)
        assert nkeys[i] == len(ok) and ctr[i] == octr, i
        mine = {int(k): counts[i, :, s] for s, k in enumerate(keys[i]) if k}
        assert sorted(mine) == sorted(int(k) for k in ok), i
        for s, k in enumerate(ok):
            assert (mine[int(k)] == oc[:, s]).all(), (i, int(k))


@pytest.mark.gpu
def test_step_host_direct_and_graph_paths(sl):
    """Host-buffer steps on a side stream: batches of up to 4096 envs take the copy-free path (actions in
    the launch parameters, outputs published by the step's last CTA), larger ones the cached-graph path
    with DMA copies; both must equal the device-buffer step, over many steps (sequence numbers) and with
    two batch sizes sharing one stream's staging block."""
    lv = gu.levels("append-spawn")
    pool = sl.game.LevelPool(lv)
    rng = np.random.RandomState(11)
    st = torch.cuda.Stream()
    for n, T in ((4096, 40), (37, 300), (4100, 6)):
        ref = sl.env.SafeLifeEnv(pool, num_envs=n, env_id0=3, global_counter=None, compute_obs=False,
                                 auto_reset=True, time_limit=25)
        host = sl.env.SafeLifeEnv(pool, num_envs=n, env_id0=3, global_counter=None, compute_obs=False,
                                  auto_reset=True, time_limit=25)
        for e in (ref, host):
            e.seed(2); e.reset()
        r = torch.zeros(n, dtype=torch.int32).pin_memory()
        d = torch.zeros(n, dtype=torch.uint8).pin_memory()
        for t in range(T):
            a = rng.randint(9, size=n).astype(np.uint8)
            _, wr, wd, _ = ref.step(torch.from_numpy(a).cuda())
            host.step_host(a, r, d, sync=False, stream=st)
            host.wait(st)
            assert (r.numpy() == wr.cpu().numpy()).all(), (n, t)
            assert (d.numpy() == wd.cpu().numpy()).all(), (n, t)
        torch.cuda.synchronize()
        assert (host.game.state == ref.game.state).all()
        assert (host.game.board == ref.game.board).all()


@pytest.mark.gpu
@pytest.mark.parametrize("h,w", [(9, 31), (32, 32), (5, 7)])
def test_step_host_direct_path_vs_oracle(sl, h, w):
    """The copy-free host-buffer step (res_env_step_host_kernel, every plane-group / runtime-shape
    variant: dense random levels carry spawners on boards AND goals) against the oracle, step by step,
    with auto-reset, through the bound hot-loop stepper (SafeLifeEnv.host_stepper)."""
    rng = np.random.RandomState(h * 31 + w)
    lv = _random_levels(rng, 24, h, w)
    n, steps = 130, 50  # not a multiple of the 4 envs per CTA: the last CTA has idle warps
    pool = sl.game.LevelPool(lv)
    env = sl.env.SafeLifeEnv(pool, num_envs=n, env_id0=9, global_counter=None, auto_reset=True, time_limit=20,
                             compute_obs=False)
    env.seed(h)
    env.reset()
    lvl = (np.arange(n) + 9) % pool.n_levels
    oenv = slo.OracleBatchEnv(slo.OraclePool(**lv), lvl, seed=h, env_id0=9, threads=4, time_limit=20)
    stride = n % pool.n_levels or pool.n_levels
    st = torch.cuda.Stream()
    hs = env.host_stepper(st)
    a = torch.zeros(n, dtype=torch.uint8).pin_memory()
    r = torch.zeros(n, dtype=torch.int32).pin_memory()
    d = torch.zeros(n, dtype=torch.uint8).pin_memory()
    for t in range(steps):
        acts = rng.randint(9, size=n)
        a.copy_(torch.from_numpy(acts.astype(np.uint8)))
        hs.step(a.data_ptr(), r.data_ptr(), d.data_ptr())
        hs.wait()
        orr, od = oenv.step(acts)
        assert (r.numpy() == orr).all(), t
        assert (d.numpy() == (od != 0)).all(), t
        if od.any():
            idx = np.nonzero(od)[0]
            lvl[idx] = (lvl[idx] + stride) % pool.n_levels
            oenv.reset(idx, lvl[idx])
    torch.cuda.synchronize()
    assert (env.game.board.cpu().numpy() == oenv.boards).all()
    assert (env.game.goals.cpu().numpy() == oenv.goals).all()
    stt, ost = env.game.state_numpy(), oenv.states
    for f in stt.dtype.names:
        assert (stt[f] == ost[f]).all(), f


@pytest.mark.gpu
def test_level_pool_update_feeds_new_levels(sl):
    """LevelPool.update: slots refilled from the host (the procedural-generation feed) are what envs play
    after their next reset, bit-exact with an oracle pool that held those levels from the start; envs on
    untouched slots are not disturbed."""
    base = gu.level_slice(gu.levels("prune-spawn"), 0, 8)
    fresh = gu.level_slice(gu.levels("append-spawn"), 10, 12)
    pool = sl.game.LevelPool(base)
    n = 16
    env = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False, auto_reset=True, time_limit=30)
    env.seed(4)
    env.reset()
    rng = np.random.RandomState(1)
    for _ in range(10):
        env.step(rng.randint(9, size=n))
    pool.update([2, 5], fresh)
    merged = {k: np.array(base[k]) for k in base}
    for k in merged:
        merged[k][[2, 5]] = fresh[k]
    assert (pool.board.cpu().numpy() == merged["board"]).all() and (pool.goals.cpu().numpy() == merged["goals"]).all()
    # a second env batch on the same pool starts from the merged level set
    env2 = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False, auto_reset=True, time_limit=30)
    env2.seed(6)
    env2.reset()
    lvl = np.arange(n) % 8
    oenv = slo.OracleBatchEnv(slo.OraclePool(**merged), lvl, seed=6, env_id0=0, threads=2, time_limit=30)
    stride = n % 8 or 8
    for t in range(80):
        acts = rng.randint(9, size=n)
        _, r, d, _ = env2.step(acts)
        orr, od = oenv.step(acts)
        assert (r.cpu().numpy() == orr).all() and (d.cpu().numpy() == (od != 0)).all(), t
        if od.any():
            idx = np.nonzero(od)[0]
            lvl[idx] = (lvl[idx] + stride) % 8
            oenv.reset(idx, lvl[idx])
    assert (env2.game.board.cpu().numpy() == oenv.boards).all()
    assert (env2.game.goals.cpu().numpy() == oenv.goals).all()
    stt, ost = env2.game.state_numpy(), oenv.states
    for f in stt.dtype.names:
        assert (stt[f] == ost[f]).all(), f


@pytest.mark.gpu
@pytest.mark.parametrize("h,w", [(26, 26), (13, 20), (32, 32)])
def test_advance_board_batch_sparse_and_dense(sl, h, w):
    """slb_advance_board on a batch that mixes the kernel's write-back paths: boards on which nothing
    changes (copied out), boards with a few flipped cells per row (patched into the staged image) and
    dense soups (full back-transposition), with spawners and all 16 cell bits, k = 1 and k = 3, each
    board on its own stream, against the oracle."""
    rng = np.random.RandomState(h * 7 + w)
    n = 48
    boards = np.zeros((n, h, w), np.uint16)
    for i in range(n):
        kind = i % 4
        if kind == 0:    # still life + walls: nothing changes
            boards[i, 2:4, 2:4] = 1033
            boards[i, 0, :] = 16
        elif kind == 1:  # a few blinkers / a glider: a handful of flips
            boards[i, 5, 3:6] = 521
            boards[i, h - 3, w - 4:w - 1] = 1033
            boards[i, 1, 1] = 1688
        elif kind == 2:  # dense soup of typical cell types
            vals = np.array([0, 0, 0, 1033, 521, 16, 17, 1688, 32788, 2057], np.uint16)
            boards[i] = vals[rng.randint(len(vals), size=(h, w))]
        else:            # noise over all 16 bits, sparse
            boards[i] = np.where(rng.rand(h, w) < 0.15, rng.randint(0, 65536, size=(h, w)), 0).astype(np.uint16)
    ctr0 = rng.randint(0, 500, size=n).astype(np.int64)
    for k in (1, 3):
        out, ctr = dev_advance(sl, boards, 0.3, 21, 1000, ctr0, k=k)
        for i in range(n):
            b, c = boards[i], int(ctr0[i])
            for _ in range(k):
                b, c = slo.advance_board(b, 0.3, seed=21, env=1000 + i, counter=c)
            assert (out[i] == b).all(), (k, i)
            assert ctr[i] == c, (k, i)
