"""Times the UNMODIFIED reference (its own Python + its compiled extension) in THIS container, single
core: tiers T0 / T1 / T2 of SURVEY.md §8(d).  Needs /root/reference (cannot run on the GPU box, where
bench.py times oracle/_ref's advance_board.c + the C port of the rule code instead).

    python oracle/time_reference_here.py
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import reference_harness as rh  # noqa: E402

ACTIONS = ["NULL", "MOVE UP", "MOVE RIGHT", "MOVE DOWN", "MOVE LEFT", "TOGGLE UP", "TOGGLE RIGHT",
           "TOGGLE DOWN", "TOGGLE LEFT"]


def best_of(fn, reps=3):
    return min(fn() for _ in range(reps))


def main():
    rh._import()
    from safelife.safelife_game import SafeLifeGame
    sp = rh.speedups()
    out = {"where": "build container, 1 core of %d" % len(os.sched_getaffinity(0))}
    for archive in ("append-still", "prune-spawn"):
        lv = rh.load_archive("benchmarks/v1.0/" + archive)
        game = SafeLifeGame.loaddata(lv[0])
        rng = np.random.RandomState(2)
        n = 20000

        def t0():
            b, g = game.board.copy(), game.goals.copy()
            t = time.perf_counter()
            for _ in range(n):
                b = sp.advance_board(b, 0.3)
                g = sp.advance_board(g, 0.3)
            return (time.perf_counter() - t) / n

        acts = rng.randint(9, size=3000)

        def t1():
            gm = SafeLifeGame.loaddata(lv[0])
            t = time.perf_counter()
            for a in acts:
                gm.execute_action(ACTIONS[a])
                gm.advance_board()
                gm.current_points()
                gm.update_exit_colors()
            return (time.perf_counter() - t) / len(acts)

        def t2():
            env = rh.make_env(levels=lv[:1], global_counter=None)
            env.reset()
            t = time.perf_counter()
            for a in acts:
                env.step(int(a))
            return (time.perf_counter() - t) / len(acts)

        out[archive] = {"T0_kernel_only_us": best_of(t0) * 1e6, "T1_game_step_us": best_of(t1) * 1e6,
                        "T2_env_step_with_obs_us": best_of(t2) * 1e6}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
