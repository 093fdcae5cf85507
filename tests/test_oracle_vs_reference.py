"""Build container only: the oracle against the UNMODIFIED reference run live (compiled extension +
the reference's own Python via oracle/reference_harness.py).  Skipped where /root/reference is
absent (the GPU box) — there the committed golden vectors (test_oracle_golden.py) are the pin."""
import numpy as np
import pytest

from oracle import reference_harness as rh
from oracle import slo

pytestmark = [pytest.mark.reference,
              pytest.mark.skipif(not rh.available(), reason="needs /root/reference + oracle/_ref")]

ARCH = ["append-still", "prune-still", "append-dynamic", "prune-dynamic", "append-spawn",
        "prune-spawn", "navigation", "prune-still-hard"]


def test_stream_three_statements_agree():
    a = slo.stream_doubles(123456789012345, 987654321, 2 ** 33 - 5, 4000)
    b = rh.stream_doubles_np(123456789012345, 987654321, 2 ** 33 - 5, 4000)
    assert (a == b).all()


def test_feeder_is_what_the_extension_consumes():
    board = np.zeros((5, 5), np.uint16)
    board[2, 2] = 16 | 128 | 8 | 1024
    with rh.fed_reference(123, 4) as feeder:
        out = rh.advance_board(board, 0.3)
        assert feeder.calls == [10000]  # one refill of RAND_BUFFER_SIZE (random.c:8)
    draws = rh.stream_doubles_np(123, 4, 0, 8)
    neigh = [(1, 1), (1, 2), (1, 3), (2, 1), (2, 3), (3, 1), (3, 2), (3, 3)]  # raster order
    for (y, x), u in zip(neigh, draws):
        assert out[y, x] == (1033 if u < float(np.float32(0.3)) else 0)


@pytest.mark.parametrize("archive", ARCH)
def test_advance_board_all_levels(archive):
    lv = rh.load_archive("benchmarks/v1.0/" + archive)
    for li in range(0, 100, 3):
        for which in ("board", "goals"):
            with rh.fed_reference(5, li):
                rb = ob = lv[li][which].copy()
                ctr = 0
                for _ in range(30):
                    rb = rh.advance_board(rb, 0.3)
                    ob, ctr = slo.advance_board(ob, 0.3, seed=5, env=li, counter=ctr)
                    assert (rb == ob).all()


@pytest.mark.parametrize("archive,seed,aseed", [("prune-spawn", 7, 3), ("append-dynamic", 2, 8),
                                                ("prune-still-hard", 4, 1)])
def test_env_trace_live(archive, seed, aseed):
    lv = rh.load_archive("benchmarks/v1.0/" + archive)[10:13]
    acts = np.random.RandomState(aseed).randint(9, size=1500)
    tr = rh.run_env_trace(lv, acts, seed=seed, env_id=42)
    pool = slo.OraclePool(lv["board"], lv["goals"], lv["agent_loc"], lv["orientation"],
                          lv["spawn_prob"], lv["min_performance"])
    env = slo.OracleBatchEnv(pool, [0], seed=seed, env_id0=42)
    level = 0
    for t, a in enumerate(acts):
        r, d = env.step([a])
        assert (env.boards[0] == tr["board"][t]).all() and (env.goals[0] == tr["goals"][t]).all()
        assert r[0] == tr["reward"][t] and d[0] == tr["done"][t]
        assert env.states[0]["completed"] == tr["completed"][t]
        assert env.states[0]["possible"] == tr["possible"][t]
        if d[0]:
            level = (level + 1) % 3
            env.reset([0], [level])
