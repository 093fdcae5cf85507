"""
Generates every fixture under tests/golden/ by RUNNING THE UNMODIFIED REFERENCE (the compiled
extension + the reference's Python, via oracle/reference_harness.py).  Runs only in the build
container (needs /root/reference).  The reference ships no tests or golden vectors of its own
(SURVEY.md §4), so these are the pins.

    python tests/golden/make_golden.py

Outputs (all np.savez_compressed):
    levels_v1.npz     the 8 x 100 benchmarks/v1.0 levels re-packed as plain arrays (level pool input)
    kat_advance.npz   advance_board known answers: hand cases, Life patterns, example levels, random
                      boards of odd sizes incl. all-16-bit noise, 64x64 and 256x256, with spawners
    trace_*.npz       full per-step SafeLifeEnv traces (config 1 append-still; prune-spawn stochastic)
    scenarios.npz     scripted execute_action cases on real SafeLifeGame objects (push/pull/exit/toggle)
    side_effects.npz  pre-EMD density counts of side_effect_score for finished episodes
    obs.npz           SafeLifeEnv.get_obs along short traces (several view shapes / channel sets)
    wrappers.npz      rewards of MovementBonusWrapper / SimpleSideEffectPenalty around the real env
    benchmark_log.npz the YAML episode records RecordingSafeLifeWrapper writes (EMD stubbed to 0)
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import reference_harness as rh  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
ARCHIVES = ["append-still", "prune-still", "append-dynamic", "prune-dynamic", "append-spawn",
            "prune-spawn", "navigation", "prune-still-hard"]
ACTION_NAMES = ["NULL", "MOVE UP", "MOVE RIGHT", "MOVE DOWN", "MOVE LEFT",
                "TOGGLE UP", "TOGGLE RIGHT", "TOGGLE DOWN", "TOGGLE LEFT"]
SEED = 20261104  # stream seed used for all KATs


def save(name, **arrays):
    path = os.path.join(OUT, name)
    np.savez_compressed(path, **arrays)
    print("%-22s %8.1f KiB" % (name, os.path.getsize(path) / 1024))


# -------------------------------------------------------------------------------------------------
def make_levels():
    out = {}
    for a in ARCHIVES:
        lv = rh.load_archive("benchmarks/v1.0/" + a)
        key = a.replace("-", "_")
        out[key + "__board"] = lv["board"].astype(np.uint16)
        out[key + "__goals"] = lv["goals"].astype(np.uint16)
        out[key + "__agent_loc"] = lv["agent_loc"].astype(np.int32)  # (x, y)
        out[key + "__orientation"] = lv["orientation"].astype(np.int32)
        out[key + "__spawn_prob"] = lv["spawn_prob"].astype(np.float64)
        out[key + "__min_performance"] = lv["min_performance"].astype(np.float64)
    save("levels_v1.npz", **out)


# -------------------------------------------------------------------------------------------------
def random_board(rng, h, w, kind):
    """Synthetic boards. kind: 'life' (colours + walls), 'mixed' (all cell kinds incl. spawners),
    'noise' (uniform random uint16: every bit incl. the unused 12-14)."""
    if kind == "noise":
        return rng.randint(0, 1 << 16, size=(h, w)).astype(np.uint16)
    vals = [0, 1033, 16, 1041, 521, 32788, 1688, 1545, 513, 9, 2057, 3584 | 1, 272, 784, 1 | 256,
            152 | 2048, 144 | 512, 48, 80, 122, 17 | 4, 1 | 32 | 4 | 16, 1 | 64 | 4 | 16]
    if kind == "life":
        p = np.array([60, 25, 5, 4, 3, 0, 0, 3] + [0] * 15, dtype=float)
    else:
        p = np.array([50, 15, 4, 3, 5, 2, 2, 3, 2, 2, 2, 1, 1, 1, 1, 1, 1, 1, 1, 0.5, 0.5, 0.5, 0.5])
    p = p / p.sum()
    return np.array(vals, dtype=np.uint16)[rng.choice(len(vals), size=(h, w), p=p)]


def advance_seq(board, steps, env, spawn_prob=0.3):
    """-> list of successive boards from the real reference on stream (SEED, env)."""
    seq = []
    ctx = rh.fed_reference(SEED, env)
    with ctx:
        b = board
        for _ in range(steps):
            b = rh.advance_board(b, spawn_prob)
            seq.append(b.copy())
    return seq


def make_kat_advance():
    R, G, B_ = 512, 1024, 2048
    cases = []  # (name, board, steps, spawn_prob)

    def blank(h, w):
        return np.zeros((h, w), np.uint16)

    b = blank(7, 7); b[3, 2:5] = 9 | R
    cases.append(("blinker_red", b, 4, 0.3))
    b = blank(7, 7); b[3, 2] = 1 | 8 | R; b[3, 3] = 1 | R | G; b[3, 4] = 1 | 8 | G
    cases.append(("mixed_colour_birth", b, 3, 0.3))
    b = blank(7, 7); b[3, 2] = 1 | 256; b[3, 3] = 1 | 256; b[3, 4] = 1
    cases.append(("exit_bit_parents", b, 3, 0.3))
    b = blank(9, 9); b[4, 3:6] = 9 | G; b[3, 6] = 122
    cases.append(("player_freeze", b, 4, 0.3))
    b = blank(5, 5); b[2, 2] = 16 | 128 | 8 | G
    cases.append(("lone_spawner", b, 12, 0.3))
    b = blank(6, 6); b[0, 0] = 16 | 128 | R; b[5, 5] = 16 | 128 | B_ | 8; b[2, 3] = 9 | G
    cases.append(("corner_spawners_wrap", b, 12, 0.3))
    b = blank(5, 5); b[2, 2] = 16 | 128 | 8 | G
    cases.append(("spawner_p0", b, 4, 0.0))
    cases.append(("spawner_p1", b, 4, 1.0))
    cases.append(("spawner_p_tiny", b, 4, 1e-9))
    cases.append(("spawner_p_0.999", b, 4, 0.999))
    b = blank(8, 8); b[1:4, 1:4] = 9 | R; b[5, 5] = 32 | 16; b[4, 4:7] = 1 | B_
    cases.append(("preserving_fountain", b, 5, 0.3))
    b = blank(8, 8); b[2, 1:4] = 9 | R; b[3, 4] = 64 | 16
    cases.append(("inhibitor", b, 5, 0.3))
    b = blank(6, 6); b[1, 1] = 9 | R | 32768 | 4096 | 2; b[1, 2] = 9; b[2, 1] = 9; b[4, 4] = 32768 | 8192
    cases.append(("high_bits_passthrough", b, 4, 0.3))
    for h, w in [(3, 3), (3, 4), (4, 3), (3, 31), (31, 3)]:
        rng = np.random.RandomState(h * 100 + w)
        cases.append(("tiny_%dx%d" % (h, w), random_board(rng, h, w, "mixed"), 6, 0.3))

    for nm, d in rh.load_level_files("patterns"):
        cases.append(("pattern_" + nm, d["board"].astype(np.uint16), 40, 0.3))
    for nm, d in rh.load_level_files("examples"):
        cases.append(("example_" + nm.replace(" ", "_"), d["board"].astype(np.uint16), 25,
                      float(d["spawn_prob"]) if "spawn_prob" in d else 0.3))
        if "goals" in d:
            cases.append(("example_goals_" + nm.replace(" ", "_"), d["goals"].astype(np.uint16), 10, 0.3))
    for nm, d in rh.load_level_files("benchmarks/v0.1")[:12]:
        cases.append(("v01_" + nm, d["board"].astype(np.uint16), 25, float(d["spawn_prob"])))

    sizes = [(5, 5), (7, 9), (13, 8), (16, 16), (25, 25), (26, 26), (26, 26), (26, 26), (31, 32),
             (32, 32), (32, 31), (33, 40), (40, 33), (26, 64), (64, 26), (64, 64), (100, 37)]
    for i, (h, w) in enumerate(sizes):
        for kind in ("life", "mixed", "noise"):
            rng = np.random.RandomState(1000 + 10 * i + len(kind))
            steps = 8 if kind != "noise" else 4
            cases.append(("rand_%s_%dx%d_%d" % (kind, h, w, i), random_board(rng, h, w, kind), steps, 0.3))
    rng = np.random.RandomState(5)
    cases.append(("big_256_mixed", random_board(rng, 256, 256, "mixed"), 3, 0.3))
    cases.append(("big_256_life", random_board(rng, 256, 256, "life"), 3, 0.3))
    cases.append(("big_128x96_noise", random_board(rng, 128, 96, "noise"), 2, 0.3))

    out = {"names": np.array([c[0] for c in cases])}
    for env, (name, board, steps, p) in enumerate(cases):
        seq = advance_seq(board, steps, env, p)
        out["in_%d" % env] = board
        out["out_%d" % env] = np.stack(seq)
        out["p_%d" % env] = np.float64(p)
    out["seed"] = np.uint64(SEED)
    save("kat_advance.npz", **out)


# -------------------------------------------------------------------------------------------------
def make_traces():
    for archive, level, s, a, env_id, T in [("append-still", 0, 1234, 99, 0, 1003),
                                            ("prune-spawn", 0, 7, 3, 5, 1003),
                                            ("append-spawn", 17, 21, 22, 123456789012, 400),
                                            ("navigation", 3, 31, 32, 7, 400)]:
        lv = rh.load_archive("benchmarks/v1.0/" + archive)
        actions = np.array([np.random.RandomState(a).randint(9) for _ in range(1)])  # placeholder
        rng = np.random.RandomState(a)
        actions = np.array([int(rng.randint(9)) for _ in range(T)], dtype=np.int32)
        levels = lv[level:level + 2]
        tr = rh.run_env_trace(levels, actions, seed=s, env_id=env_id)
        h = hashlib.sha256()
        for t in range(min(T, 1000)):
            h.update(tr["board"][t].tobytes()); h.update(tr["goals"][t].tobytes())
            h.update(np.int64(tr["reward"][t]).tobytes())
        print("  %s level %d digest(1000) %s  sum_r %d" % (archive, level, h.hexdigest(),
                                                            tr["reward"][:1000].sum()))
        tr.update(actions=actions, seed=np.uint64(s), env_id=np.uint64(env_id),
                  level0=np.int32(level), archive=np.array(archive), digest1000=np.array(h.hexdigest()))
        save("trace_%s.npz" % archive.replace("-", "_"), **tr)


# -------------------------------------------------------------------------------------------------
def make_scenarios():
    """Scripted actions on real SafeLifeGame objects built from dict records."""
    rh._import()
    from safelife.safelife_game import SafeLifeGame

    PLAYER, CRATE, EXIT, WALL = 122, 32788, 272, 16
    scen = []

    def level(h, w, agent, cells, goals=None, min_perf=-1.0, orientation=1):
        b = np.zeros((h, w), np.uint16)
        for (x, y), v in cells.items():
            b[y, x] = v
        b[agent[1], agent[0]] = PLAYER
        g = np.zeros((h, w), np.uint16)
        for (x, y), v in (goals or {}).items():
            g[y, x] = v
        return {"board": b, "goals": g, "agent_loc": np.array(agent), "orientation": orientation,
                "spawn_prob": 0.3, "min_performance": min_perf,
                "class": "safelife.safelife_game.SafeLifeGame"}

    U, R_, D, L, TU, TR, TD, TL, N = 1, 2, 3, 4, 5, 6, 7, 8, 0
    scen.append(("push_crate_right", level(8, 8, (2, 3), {(3, 3): CRATE}), [R_, R_, R_, R_, R_, R_, R_]))
    scen.append(("push_crate_blocked", level(8, 8, (2, 3), {(3, 3): CRATE, (4, 3): WALL}), [R_, R_, U, D]))
    scen.append(("pull_crate", level(8, 8, (3, 3), {(2, 3): CRATE}), [R_, R_, D, U, U, L, L]))
    scen.append(("push_and_pull_train", level(9, 9, (4, 4), {(3, 4): CRATE, (5, 4): CRATE}), [R_, R_, L, L, L]))
    scen.append(("push_crate_into_exit", level(8, 8, (2, 3), {(3, 3): CRATE, (4, 3): EXIT}), [R_, R_, R_]))
    scen.append(("walk_into_exit_open", level(8, 8, (2, 3), {(4, 3): EXIT}), [R_, R_, R_, L, TU]))
    scen.append(("exit_closed_then_open",
                 level(8, 8, (1, 1), {(3, 1): EXIT}, goals={(5, 5): 2057, (6, 5): 2057}, min_perf=0.5),
                 [R_, R_, D, D, D, D, R_, R_, R_, TR, U, U, U, U, L, L, L, U, R_, R_, R_]))
    scen.append(("toggle_create_destroy", level(8, 8, (3, 3), {(3, 2): 1033, (4, 3): WALL, (3, 4): 1688}),
                 [TU, TU, TR, TD, TL, TL, N]))
    scen.append(("toggle_coloured_agent", level(8, 8, (3, 3), {}), [TU, TR, TD, TL, N, N]))
    lv = level(8, 8, (3, 3), {}); lv["board"][3, 3] = PLAYER | 1024
    scen.append(("toggle_green_agent", lv, [TU, TR, TD, TL, N, N]))
    scen.append(("wrap_moves", level(5, 6, (0, 0), {(5, 0): CRATE}), [L, L, U, U, R_, D, L]))
    scen.append(("w3_alias_push_pull", level(3, 3, (1, 1), {(2, 1): CRATE}), [R_, R_, L, U, D]))
    scen.append(("w3_alias_vertical", level(3, 4, (1, 1), {(1, 0): CRATE, (1, 2): 32788}), [U, D, D, U]))
    scen.append(("h4_push_pull", level(4, 4, (1, 1), {(1, 0): CRATE, (1, 2): CRATE}), [U, U, D, D, D]))
    scen.append(("push_pushable_exit_cell", level(8, 8, (2, 3), {(3, 3): 272 | 4}), [R_, R_, U]))
    scen.append(("move_into_life", level(8, 8, (2, 3), {(3, 3): 1033, (3, 2): 1033, (4, 3): 1033}), [R_, TR, R_, R_]))
    scen.append(("pull_plant_and_weed", level(9, 9, (4, 4), {(3, 4): 16 | 1 | 4 | 32768 | 1024, (4, 5): 32 | 1 | 4 | 16}),
                 [R_, R_, U, U, D]))

    out = {"names": np.array([s[0] for s in scen])}
    for i, (name, rec, acts) in enumerate(scen):
        ctx = rh.fed_reference(SEED, 1000 + i)
        with ctx:
            g = SafeLifeGame.loaddata(rec)
            T = len(acts)
            H, W = g.board.shape
            boards = np.zeros((T, 2, H, W), np.uint16)  # [t,0]=after action, [t,1]=after physics+exit colours
            goals = np.zeros((T, H, W), np.uint16)
            meta = np.zeros((T, 8), np.int64)  # reward, x, y, orient, game_over, can_exit_before, points, completed
            g.update_exit_colors()
            start_board = g.board.copy()
            for t, a in enumerate(acts):
                ce = bool(g.can_exit())
                r = g.execute_action(ACTION_NAMES[a])
                boards[t, 0] = g.board
                g.advance_board()
                pts = g.current_points()
                g.update_exit_colors()
                boards[t, 1] = g.board
                goals[t] = g.goals
                meta[t] = [r, g.agent_loc[0], g.agent_loc[1], g.orientation, bool(g.game_over), ce,
                           pts, g.performance_ratio()[0]]
        out["board0_%d" % i] = rec["board"]
        out["goals0_%d" % i] = rec["goals"]
        out["start_%d" % i] = start_board
        out["agent_%d" % i] = np.array(list(rec["agent_loc"]) + [rec["orientation"]], np.int32)
        out["minperf_%d" % i] = np.float64(rec["min_performance"])
        out["actions_%d" % i] = np.array(acts, np.int32)
        out["boards_%d" % i] = boards
        out["goals_%d" % i] = goals
        out["meta_%d" % i] = meta
    out["seed"] = np.uint64(SEED)
    save("scenarios.npz", **out)


# -------------------------------------------------------------------------------------------------
def make_side_effects():
    rh._import()
    out = {"seed": np.uint64(SEED)}
    cases = [("prune-spawn", 0, 7, 3, 1001, 1000), ("append-still", 4, 1, 2, 1001, 1000),
             ("append-spawn", 9, 3, 4, 250, 300), ("prune-dynamic", 2, 5, 6, 120, 200),
             ("navigation", 1, 8, 9, 60, 100)]
    for i, (archive, level, s, a, n_steps, n_samples) in enumerate(cases):
        lv = rh.load_archive("benchmarks/v1.0/" + archive)
        ctx = rh.fed_reference(SEED, 2000 + i)
        with ctx:
            env = rh.make_env(levels=lv[level:level + 1], global_counter=None)
            env.reset()
            ctx.rewind(0)
            rng = np.random.RandomState(a)
            acts = np.array([int(rng.randint(9)) for _ in range(n_steps)], np.int32)
            for act in acts:
                env.step(int(act))
            game = env.game
            final_board = game.board.copy()
            res = rh.side_effect_pre_emd(game, num_samples=n_samples)
        keys = np.array(sorted(res.keys()), np.uint16)
        H, W = final_board.shape
        counts = np.zeros((2, len(keys), H, W), np.uint16)
        sums = np.zeros(len(keys), np.float64)
        for k, key in enumerate(keys):
            ina, act_, s_ = res[int(key)]
            c0, c1 = np.rint(ina * n_samples), np.rint(act_ * n_samples)
            assert (c0 / n_samples == ina).all() and (c1 / n_samples == act_).all()
            counts[0, k], counts[1, k], sums[k] = c0, c1, s_
        out.update({
            "archive_%d" % i: np.array(archive), "level_%d" % i: np.int32(level),
            "actions_%d" % i: acts, "env_%d" % i: np.uint64(2000 + i),
            "num_steps_%d" % i: np.int32(game.num_steps), "num_samples_%d" % i: np.int32(n_samples),
            "board0_%d" % i: lv[level]["board"].astype(np.uint16),
            "goals0_%d" % i: lv[level]["goals"].astype(np.uint16),
            "final_board_%d" % i: final_board, "keys_%d" % i: keys, "counts_%d" % i: counts,
            "sum_inaction_%d" % i: sums,
        })
        print("  side effects %s: keys %s sums %s" % (archive, keys.tolist(), np.round(sums, 3).tolist()))
    out["n_cases"] = np.int32(len(cases))
    save("side_effects.npz", **out)


# -------------------------------------------------------------------------------------------------
def make_obs():
    """SafeLifeEnv.get_obs (safelife_env.py:125-155) of the real reference env along short traces:
    default 15x15x15 channels, packed words, a view larger than the board (33x33, the training
    view), an even-sized view, white goals kept."""
    cases = [
        ("append-still", 4, 11, 12, 3, (15, 15), True, tuple(range(15))),
        ("prune-spawn", 9, 13, 14, 4, (33, 33), True, None),
        ("navigation", 2, 15, 16, 5, (9, 12), False, (0, 1, 4, 8, 9, 10, 11, 12, 13, 14)),
        ("append-dynamic", 30, 17, 18, 6, (15, 15), False, None),
        ("prune-still-hard", 7, 19, 20, 7, (26, 26), True, None),
    ]
    out = {"n_cases": np.int32(len(cases))}
    T, every = 80, 4
    for i, (archive, level, s, a, env_id, view, rw, ch) in enumerate(cases):
        lv = rh.load_archive("benchmarks/v1.0/" + archive)
        rng = np.random.RandomState(a)
        actions = np.array([int(rng.randint(9)) for _ in range(T)], dtype=np.int32)
        ctx = rh.fed_reference(s, env_id)
        with ctx:
            env = rh.make_env(levels=lv[level:level + 1], global_counter=None, view_shape=view,
                              remove_white_goals=rw, output_channels=ch)
            obs = [np.asarray(env.reset()).astype(np.uint16)]
            ctx.rewind(0)
            for t, act in enumerate(actions):
                o, r, done, info = env.step(int(act))
                if (t + 1) % every == 0:
                    obs.append(np.asarray(o).astype(np.uint16))
        out.update({
            "archive_%d" % i: np.array(archive), "level_%d" % i: np.int32(level), "seed_%d" % i: np.uint64(s),
            "env_%d" % i: np.uint64(env_id), "actions_%d" % i: actions, "view_%d" % i: np.array(view, np.int32),
            "remove_white_%d" % i: np.int32(rw), "channels_%d" % i: np.array(ch if ch else [], np.int32),
            "every_%d" % i: np.int32(every), "obs_%d" % i: np.stack(obs),
        })
        print("  obs %s view %s -> %s" % (archive, view, np.stack(obs).shape))
    save("obs.npz", **out)


# -------------------------------------------------------------------------------------------------
def make_wrappers():
    """Rewards of the reference's MovementBonusWrapper and SimpleSideEffectPenalty (env_wrappers.py)
    around the real SafeLifeEnv, plus the penalty's side-effect count per step."""
    rh._import()
    from safelife import env_wrappers as ew
    cases = [("prune-spawn", 3, 41, 42, 9), ("append-still", 5, 43, 44, 10), ("navigation", 6, 45, 46, 11)]
    out = {"n_cases": np.int32(len(cases)), "penalty_coef": np.float64(0.7), "min_performance": np.float64(0.3)}
    T = 160
    for i, (archive, level, s, a, env_id) in enumerate(cases):
        lv = rh.load_archive("benchmarks/v1.0/" + archive)
        rng = np.random.RandomState(a)
        actions = np.array([int(rng.randint(9)) for _ in range(T)], dtype=np.int32)
        res = {}
        for which in ("penalty", "movement"):
            ctx = rh.fed_reference(s, env_id)
            with ctx:
                env = rh.make_env(levels=lv[level:level + 1], global_counter=None)
                if which == "penalty":
                    env = ew.SimpleSideEffectPenalty(env, penalty_coef=0.7, min_performance=0.3)
                else:
                    env = ew.MovementBonusWrapper(env)
                env.reset()
                ctx.rewind(0)
                rew, side = [], []
                for act in actions:
                    o, r, done, info = env.step(int(act))
                    rew.append(float(r))
                    if which == "penalty":
                        side.append(int(env.last_side_effect))
                res[which] = (np.array(rew, np.float64), np.array(side, np.int64))
        out.update({"archive_%d" % i: np.array(archive), "level_%d" % i: np.int32(level), "seed_%d" % i: np.uint64(s),
                    "env_%d" % i: np.uint64(env_id), "actions_%d" % i: actions,
                    "penalty_reward_%d" % i: res["penalty"][0], "side_effect_%d" % i: res["penalty"][1],
                    "movement_reward_%d" % i: res["movement"][0]})
        print("  wrappers %s: final side effect %d, sum rewards %.3f / %.3f" % (
            archive, res["penalty"][1][-1], res["penalty"][0].sum(), res["movement"][0].sum()))
    save("wrappers.npz", **out)


# -------------------------------------------------------------------------------------------------
def make_benchmark_log():
    """The YAML episode record RecordingSafeLifeWrapper.log_episode writes (env_wrappers.py:172-231)
    for short episodes of the real env (EMD stubbed to 0, pyemd being absent)."""
    import io
    rh._import()
    from safelife import env_wrappers as ew, side_effects
    cases = [("prune-spawn", 2, 51, 52, 20, 35), ("append-still", 8, 53, 54, 21, 25)]
    out = {"n_cases": np.int32(len(cases))}
    orig = side_effects.earth_mover_distance
    side_effects.earth_mover_distance = lambda a, b, *args, **kw: 0.0
    try:
        for i, (archive, level, s, a, env_id, tl) in enumerate(cases):
            lv = rh.load_archive("benchmarks/v1.0/" + archive)
            rng = np.random.RandomState(a)
            actions = np.array([int(rng.randint(9)) for _ in range(tl + 1)], dtype=np.int32)
            ctx = rh.fed_reference(s, env_id)
            log = io.StringIO()
            with ctx:
                env = rh.make_env(levels=lv[level:level + 1], global_counter=None, time_limit=tl)
                env = ew.RecordingSafeLifeWrapper(env, log_file=log, video_name=None)
                env.reset()
                ctx.rewind(0)
                for act in actions:
                    o, r, done, info = env.step(int(act))
                assert done
                title = env.game.title
            out.update({"archive_%d" % i: np.array(archive), "level_%d" % i: np.int32(level),
                        "seed_%d" % i: np.uint64(s), "env_%d" % i: np.uint64(env_id), "actions_%d" % i: actions,
                        "time_limit_%d" % i: np.int32(tl), "title_%d" % i: np.array(str(title)),
                        "log_%d" % i: np.array(log.getvalue())})
            print(log.getvalue())
    finally:
        side_effects.earth_mover_distance = orig
    save("benchmark_log.npz", **out)


# -------------------------------------------------------------------------------------------------
def make_trajectory():
    """The trajectory .npz that the reference's SafeLifeRecorder writes (env_wrappers.py:97-136) when
    RecordingSafeLifeWrapper records an episode (frames rendered by nobody: gym's VideoRecorder is a
    no-op stand-in, oracle/refpkg)."""
    import tempfile
    rh._import()
    from safelife import env_wrappers as ew
    cases = [("prune-spawn", 4, 61, 62, 30, 40), ("navigation", 1, 63, 64, 31, 30)]
    out = {"n_cases": np.int32(len(cases))}
    for i, (archive, level, s, a, env_id, tl) in enumerate(cases):
        lv = rh.load_archive("benchmarks/v1.0/" + archive)
        rng = np.random.RandomState(a)
        actions = np.array([int(rng.randint(9)) for _ in range(tl + 1)], dtype=np.int32)
        tmp = tempfile.mkdtemp()
        ctx = rh.fed_reference(s, env_id)
        with ctx:
            env = rh.make_env(levels=lv[level:level + 1], global_counter=None, time_limit=tl)
            env = ew.RecordingSafeLifeWrapper(env, log_file=None, record_side_effects=False,
                                              video_name=os.path.join(tmp, "traj"), video_recording_freq=1)
            env.reset()
            ctx.rewind(0)
            for act in actions:
                o, r, done, info = env.step(int(act))
            assert done
            env.close()
        t = np.load(os.path.join(tmp, "traj.npz"))
        out.update({"archive_%d" % i: np.array(archive), "level_%d" % i: np.int32(level),
                    "seed_%d" % i: np.uint64(s), "env_%d" % i: np.uint64(env_id), "actions_%d" % i: actions,
                    "time_limit_%d" % i: np.int32(tl), "orientation_%d" % i: np.asarray(t["orientation"]),
                    "board_%d" % i: np.asarray(t["board"]).astype(np.uint16),
                    "goals_%d" % i: np.asarray(t["goals"]).astype(np.uint16)})
        print("  trajectory %s: %d frames" % (archive, len(t["orientation"])))
    save("trajectory.npz", **out)


if __name__ == "__main__":
    assert rh.available(), "needs /root/reference and `make -C oracle ref`"
    which = sys.argv[1:] or ["levels", "kat", "traces", "scenarios", "side", "obs", "wrappers", "log", "trajectory"]
    if "levels" in which:
        make_levels()
    if "kat" in which:
        make_kat_advance()
    if "traces" in which:
        make_traces()
    if "scenarios" in which:
        make_scenarios()
    if "side" in which:
        make_side_effects()
    if "obs" in which:
        make_obs()
    if "wrappers" in which:
        make_wrappers()
    if "log" in which:
        make_benchmark_log()
    if "trajectory" in which:
        make_trajectory()
