"""CPU: the oracle (oracle/slo_oracle.c) against the golden vectors generated from the reference
(tests/golden/make_golden.py).  This is what pins the oracle on machines without /root/reference."""
import numpy as np
import pytest

from oracle import slo
from tests import golden_util as gu


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32 10 rounds
    assert [hex(x) for x in slo.philox([0, 0, 0, 0], [0, 0])] == \
        ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    assert [hex(x) for x in slo.philox([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2)] == \
        ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    assert [hex(x) for x in slo.philox([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344],
                                       [0xA4093822, 0x299F31D0])] == \
        ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]


def test_threshold_matches_double_compare():
    rng = np.random.RandomState(0)
    for p in [0.0, 1.0, 0.3, 1e-9, 0.999, 0.5, float(np.float32(1) - np.float32(2 ** -24))]:
        T = slo.threshold_u53(p)
        pf = float(np.float32(p))
        us = np.concatenate([rng.randint(0, 1 << 53, size=200, dtype=np.int64),
                             np.clip(np.array([T - 2, T - 1, T, T + 1]), 0, (1 << 53) - 1)])
        for u in us:
            assert (float(u) / 9007199254740992.0 < pf) == (int(u) < T)


@pytest.mark.parametrize("use_ref", [False, True])
def test_advance_board_known_answers(use_ref):
    if use_ref and not slo.have_ref():
        pytest.skip("oracle/_ref not built")
    n = 0
    for name, env, seed, board, outs, p in gu.kat_cases():
        if use_ref and board.size > 40000:
            continue
        b, ctr = board, 0
        for t in range(outs.shape[0]):
            b, ctr = slo.advance_board(b, p, seed=seed, env=env, counter=ctr, use_ref=use_ref)
            assert (b == outs[t]).all(), "%s step %d" % (name, t)
        n += 1
    assert n > 80


@pytest.mark.parametrize("archive", ["append_still", "prune_spawn", "append_spawn", "navigation"])
def test_env_trace(archive):
    tr = gu.load("trace_%s.npz" % archive)
    lv = gu.levels(str(tr["archive"]))
    l0 = int(tr["level0"])
    pool = slo.OraclePool(**gu.level_slice(lv, l0, l0 + 2))
    env = slo.OracleBatchEnv(pool, [0], seed=int(tr["seed"]), env_id0=int(tr["env_id"]))
    level = 0
    for t, a in enumerate(tr["actions"]):
        r, d = env.step([a])
        st = env.states[0]
        assert (env.boards[0] == tr["board"][t]).all(), t
        assert (env.goals[0] == tr["goals"][t]).all(), t
        assert r[0] == tr["reward"][t] and d[0] == tr["done"][t]
        assert (st["agent_x"], st["agent_y"]) == tuple(tr["agent_loc"][t])
        assert st["orientation"] == tr["orientation"][t]
        assert st["old_points"] == tr["points"][t]
        if d[0]:
            level = (level + 1) % 2
            env.reset([0], [level])
    assert tr["done"].sum() >= 1 or len(tr["actions"]) < 1001


def test_scenarios():
    d = gu.load("scenarios.npz")
    for i, name in enumerate(d["names"]):
        b0, g0 = d["board0_%d" % i], d["goals0_%d" % i]
        ax, ay, orient = d["agent_%d" % i]
        minperf = float(d["minperf_%d" % i])
        base = slo.baseline_score(b0, g0)
        board, goals = d["start_%d" % i].copy(), g0.copy()
        st = np.zeros(1, slo.ENV_STATE_DTYPE)
        st["agent_x"], st["agent_y"], st["orientation"] = ax, ay, orient
        ctr = 0
        for t, a in enumerate(d["actions_%d" % i]):
            c, p = slo.performance_ratio(board, goals, base)
            ce = slo.can_exit(minperf, c, p)
            meta = d["meta_%d" % i][t]
            assert ce == bool(meta[5]), (name, t)
            r = slo.execute_action(board, st, int(a), ce)
            assert (board == d["boards_%d" % i][t, 0]).all(), (name, t)
            assert r == meta[0] and st["agent_x"][0] == meta[1] and st["agent_y"][0] == meta[2]
            assert st["orientation"][0] == meta[3] and st["game_over"][0] == meta[4]
            board, ctr = slo.advance_board(board, 0.3, int(d["seed"]), 1000 + i, ctr)
            goals, ctr = slo.advance_board(goals, 0.3, int(d["seed"]), 1000 + i, ctr)
            assert slo.current_points(board, goals) == meta[6]
            c, p = slo.performance_ratio(board, goals, base)
            assert c == meta[7]
            slo.lib().slo_update_exit_colors(board.ctypes.data, b0.ctypes.data, board.shape[0],
                                             board.shape[1], int(slo.can_exit(minperf, c, p)))
            assert (board == d["boards_%d" % i][t, 1]).all(), (name, t)
            assert (goals == d["goals_%d" % i][t]).all(), (name, t)


def test_side_effect_counts():
    d = gu.load("side_effects.npz")
    for i in range(int(d["n_cases"])):
        keys, counts, _ = slo.side_effect_counts(
            d["board0_%d" % i], d["final_board_%d" % i], int(d["num_steps_%d" % i]),
            int(d["num_samples_%d" % i]), 0.3, seed=int(d["seed"]), env=int(d["env_%d" % i]),
            counter=_counter_after_episode(d, i))
        assert keys.tolist() == d["keys_%d" % i].tolist()
        assert (counts == d["counts_%d" % i]).all()
        n = int(d["num_samples_%d" % i])
        sums = [np.sum(counts[0, k] / n) for k in range(len(keys))]
        assert sums == d["sum_inaction_%d" % i].tolist()


def _counter_after_episode(d, i):
    """Replays the episode with the oracle env to learn how many draws it consumed."""
    lv = gu.levels(str(d["archive_%d" % i]))
    l0 = int(d["level_%d" % i])
    pool = slo.OraclePool(**gu.level_slice(lv, l0, l0 + 1))
    env = slo.OracleBatchEnv(pool, [0], seed=int(d["seed"]), env_id0=int(d["env_%d" % i]))
    for a in d["actions_%d" % i]:
        env.step([a])
    assert (env.boards[0] == d["final_board_%d" % i]).all()
    return int(env.states[0]["rng_counter"])


def _obs_cases():
    d = gu.load("obs.npz")
    for i in range(int(d["n_cases"])):
        yield i, d


def test_get_obs_golden():
    """oracle get_obs (numpy) replayed along the reference traces of tests/golden/obs.npz"""
    for i, d in _obs_cases():
        lv = gu.levels(str(d["archive_%d" % i]))
        l0 = int(d["level_%d" % i])
        one = gu.level_slice(lv, l0, l0 + 1)
        env = slo.OracleBatchEnv(slo.OraclePool(**one), [0], seed=int(d["seed_%d" % i]),
                                 env_id0=int(d["env_%d" % i]))
        view = tuple(int(v) for v in d["view_%d" % i])
        ch = tuple(int(c) for c in d["channels_%d" % i]) or None
        kw = dict(view_shape=view, remove_white_goals=bool(d["remove_white_%d" % i]), output_channels=ch)

        def obs():
            st = env.states[0]
            return slo.get_obs(env.boards[0], env.goals[0], (st["agent_x"], st["agent_y"]),
                               one["board"][0], **kw)

        want = d["obs_%d" % i]
        assert (obs() == want[0]).all(), (i, "reset")
        k = 1
        for t, a in enumerate(d["actions_%d" % i]):
            env.step([a])
            if (t + 1) % int(d["every_%d" % i]) == 0:
                assert (obs() == want[k]).all(), (i, t)
                k += 1
        assert k == len(want)


def test_side_effect_penalty_golden():
    """numpy restatement of SimpleSideEffectPenalty's count along the reference traces (wrappers.npz)"""
    d = gu.load("wrappers.npz")
    for i in range(int(d["n_cases"])):
        lv = gu.levels(str(d["archive_%d" % i]))
        l0 = int(d["level_%d" % i])
        one = gu.level_slice(lv, l0, l0 + 1)
        one["min_performance"] = np.array([float(d["min_performance"])])  # the wrapper overrides it at reset
        env = slo.OracleBatchEnv(slo.OraclePool(**one), [0], seed=int(d["seed_%d" % i]),
                                 env_id0=int(d["env_%d" % i]))
        last, coef = 0, float(d["penalty_coef"])
        for t, a in enumerate(d["actions_%d" % i]):
            r, _ = env.step([a])
            side = slo.side_effect_penalty_count(env.boards[0], one["board"][0], env.goals[0])
            assert side == d["side_effect_%d" % i][t], (i, t)
            assert abs((float(r[0]) - (side - last) * coef) - d["penalty_reward_%d" % i][t]) < 1e-12, (i, t)
            last = side
