"""GPU parity tests added in round 2: the wrapper's per-env min_performance, the device-side global counter,
speedups batch streams, the host advance_board staging paths, the 256x256 fused step at length, and a bounded
soak over every archive (what scripts/soak_*.py cover, sized for the suite).  Bit-exact throughout."""
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from oracle import slo
from tests import golden_util as gu
from tests.test_gpu_parity import _random_levels, sl  # noqa: F401  (fixture)

pytestmark = pytest.mark.gpu


def _exit_level(h, w, p=2):
    """agent at (2,2) facing right, exit right of it, one blue goal cell so that `possible` = 1 and `completed`
    = 0; the level file says min_performance = -1 (exit open)"""
    board = np.zeros((p, h, w), np.uint16)
    board[:, 2, 2] = 122
    board[:, 2, 3] = 272
    goals = np.zeros_like(board)
    goals[:, h - 1, w - 1] = 2048
    return {"board": board, "goals": goals, "agent_loc": np.tile([2, 2], (p, 1)), "orientation": np.ones(p, int),
            "spawn_prob": np.full(p, 0.3), "min_performance": np.full(p, -1.0)}


@pytest.mark.parametrize("h,w", [(26, 26), (6, 6), (6, 40), (8, 64)])
def test_wrapper_min_performance_reaches_can_exit(sl, h, w):
    """SimpleSideEffectPenalty.reset sets game.min_performance AFTER env.reset (env_wrappers.py:313-317): the
    exit was painted with the level's value, the next execute_action and update_exit_colors use the wrapper's.
    Warp-per-env (26x26, 6x6), generic (6x40) and CTA-per-env (8x64) kernels, vs the oracle and by hand."""
    from safelife_k2_b200 import env_wrappers as ew

    lv = _exit_level(h, w)
    n = 4
    pool = sl.game.LevelPool(lv)
    env = ew.SimpleSideEffectPenalty(
        sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False), min_performance=0.3)
    env.seed(0)
    env.reset()
    o = slo.OracleBatchEnv(slo.OraclePool(**lv), np.arange(n) % 2, seed=0)
    o.states["min_performance"][:] = 0.3  # what the reference wrapper assigns after the reset
    g = env.game
    assert (g.board.cpu().numpy() == o.boards).all()
    assert (g.board.cpu().numpy()[:, 2, 3] == 784).all()      # painted red by the reset (level value -1)
    assert g.min_performance.cpu().tolist() == [0.3] * n
    rng = np.random.RandomState(1)
    for t in range(8):
        acts = np.full(n, 2) if t == 0 else rng.randint(9, size=n)  # first: MOVE RIGHT into the exit
        _, r, d, _ = env.step(acts)
        orr, od = o.step(acts)
        assert (d.cpu().numpy() == (od != 0)).all(), t
        assert (g.board.cpu().numpy() == o.boards).all(), t
        st = g.state_numpy()
        for f in st.dtype.names:
            assert (st[f] == o.states[f]).all(), (f, t)
        if t == 0:
            assert not d.any() and (st["game_over"] == 0).all()      # the exit is closed at 30 %
            assert (g.board.cpu().numpy()[:, 2, 3] == 272).all()      # and no longer red
    # per-env: a masked reset only gives the scheduled value to the envs it resets
    env.min_performance = -1.0
    env.reset(np.array([True, False, False, True]))
    assert g.min_performance.cpu().tolist() == [-1.0, 0.3, 0.3, -1.0]
    _, r, d, _ = env.step(np.full(n, 2))
    st = g.state_numpy()
    assert st["game_over"][0] == 1 and st["game_over"][3] == 1 and d[0] and d[3]


@pytest.mark.parametrize("h,w,auto", [(26, 26, True), (26, 26, False), (9, 40, True), (8, 64, False)])
def test_global_counter_on_device(sl, h, w, auto):
    """The default GlobalCounter (counted by the step kernel / two small kernels, read lazily) equals the eager
    host count of safelife_env.py:169-173."""
    rng = np.random.RandomState(7)
    lv = _random_levels(rng, 8, h, w)
    n = 40
    pool = sl.game.LevelPool(lv)
    lazy = sl.env.GlobalCounter()
    eager = SimpleNamespace(episodes_started=0, episodes_completed=0, num_steps=0)
    a = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=lazy, compute_obs=(h == 26), auto_reset=auto, time_limit=9)
    b = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=eager, compute_obs=False, auto_reset=auto, time_limit=9)
    for e in (a, b):
        e.seed(3)
        e.reset()
    for t in range(30):
        acts = rng.randint(9, size=n)
        _, ra, da, _ = a.step(acts)
        _, rb, db, _ = b.step(acts)
        assert (ra == rb).all() and (da == db).all()
        if t % 7 == 6:
            assert lazy.num_steps == eager.num_steps and lazy.episodes_completed == eager.episodes_completed, t
    assert lazy.episodes_started == eager.episodes_started == n
    assert lazy.num_steps == eager.num_steps > 0 and lazy.episodes_completed == eager.episodes_completed > 0
    before = eager.num_steps
    lazy.num_steps = 5  # assignable, and keeps counting from there
    a.step(acts)
    b.step(acts)
    assert lazy.num_steps == 5 + (eager.num_steps - before)
    # the class-level default is a GlobalCounter shared by envs that are not given their own
    assert isinstance(sl.env.SafeLifeEnv.global_counter, sl.env.GlobalCounter)


def test_step_returns_fresh_tensors(sl):
    lv = gu.level_slice(gu.levels("prune-spawn"), 0, 8)
    env = sl.env.SafeLifeEnv(sl.game.LevelPool(lv), num_envs=16, global_counter=None)
    env.seed(0)
    env.reset()
    rs, ds = [], []
    rng = np.random.RandomState(0)
    for _ in range(6):
        _, r, d, _ = env.step(rng.randint(5, 9, size=16))
        rs.append(r)
        ds.append(d)
    assert len({r.data_ptr() for r in rs}) == 6 and len({d.data_ptr() for d in ds}) == 6
    assert any(int(r.abs().sum()) for r in rs) and rs[0].dtype == torch.int32


def test_speedups_batch_streams_continue(sl):
    """speedups.advance_board on a batch: board i keeps consuming env i's stream over successive calls (the
    reference's stream keeps running too); seed() rewinds.  numpy and tensor batches, vs the oracle."""
    sp = sl.speedups
    lv = gu.levels("prune-spawn")
    b0 = lv["board"][:12].copy()
    n, h, w = b0.shape
    for tensor in (False, True):
        sp.seed(41)
        ctr = np.zeros(n, dtype=np.uint64)
        cur = torch.from_numpy(b0).cuda() if tensor else b0
        ocur = b0.copy()
        firsts = []
        for k in range(4):
            cur = sp.advance_board(cur, 0.3)
            nxt = np.empty_like(ocur)
            for i in range(n):
                nxt[i], c = slo.advance_board(ocur[i], 0.3, seed=41, env=i, counter=int(ctr[i]))
                ctr[i] = c
            ocur = nxt
            got = cur.cpu().numpy() if tensor else cur
            assert (got == ocur).all(), (tensor, k)
            firsts.append(got.copy())
        assert ctr.min() > 0
        sp.seed(41)  # rewinds every board's stream
        again = sp.advance_board(torch.from_numpy(b0).cuda() if tensor else b0, 0.3)
        again = again.cpu().numpy() if tensor else again
        assert (again == firsts[0]).all()


@pytest.mark.parametrize("n,h,w,k", [(1, 26, 26, 1), (3, 26, 26, 3), (20, 26, 26, 1), (700, 26, 26, 2),
                                     (2, 9, 40, 1), (3, 9, 40, 3), (300, 40, 33, 2)])
def test_advance_board_host_paths(sl, n, h, w, k):
    """slb_advance_board_host (what speedups.advance_board(ndarray) binds): the zero-copy path (a few boards read
    from / written to the pinned block by the kernel itself), the pinned staging path and shapes that need the
    scratch board, with draw counters in and out, vs the oracle; repeated calls reuse the cached blocks."""
    rng = np.random.RandomState(n + h)
    vals = np.array([0, 0, 1033, 521, 16, 1041, 32788, 1688, 152, 9], np.uint16)
    L = sl.lib.load()
    for rep in range(3):
        b = vals[rng.randint(len(vals), size=(n, h, w))]
        p = rng.choice([0.0, 0.3, 1.0], size=n).astype(np.float32)
        ctr = rng.randint(0, 1000, size=n).astype(np.uint64)
        out = np.empty_like(b)
        c2 = ctr.copy()
        sl.lib.check(L.slb_advance_board_host(b.ctypes.data, out.ctypes.data, n, h, w, p.ctypes.data, 9, 100,
                                              c2.ctypes.data, k))
        for i in range(n):
            x, c = b[i], int(ctr[i])
            for _ in range(k):
                x, c = slo.advance_board(x, float(p[i]), seed=9, env=100 + i, counter=c)
            assert (out[i] == x).all(), (rep, i)
            assert int(c2[i]) == c, (rep, i)
    # NULL counters: streams start at draw 0
    out = np.empty_like(b)
    sl.lib.check(L.slb_advance_board_host(b.ctypes.data, out.ctypes.data, n, h, w, p.ctypes.data, 9, 0, None, 1))
    for i in range(min(n, 8)):
        assert (out[i] == slo.advance_board(b[i], float(p[i]), seed=9, env=i, counter=0)[0]).all()


def test_host_step_slot_regrows_with_a_step_in_flight(sl):
    """Two batches of different sizes sharing one stream through the async host step: the second needs a larger
    staging block while the first step's outputs are still pending (they must be delivered first)."""
    lv = gu.level_slice(gu.levels("append-spawn"), 0, 10)
    pool = sl.game.LevelPool(lv)
    st = torch.cuda.Stream()
    envs, refs = [], []
    for n in (64, 1500):
        for lst in (envs, refs):
            e = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False)
            e.seed(2)
            e.reset()
            lst.append(e)
    rng = np.random.RandomState(0)
    a0 = rng.randint(9, size=64).astype(np.uint8)
    a1 = rng.randint(9, size=1500).astype(np.uint8)
    r0, d0 = np.zeros(64, np.int32), np.zeros(64, np.uint8)
    r1, d1 = np.zeros(1500, np.int32), np.zeros(1500, np.uint8)
    envs[0].step_host(a0, r0, d0, sync=False, stream=st)      # not waited for
    envs[1].step_host(a1, r1, d1, sync=True, stream=st)       # larger block on the same slot
    _, rr0, dd0, _ = refs[0].step(a0)
    _, rr1, dd1, _ = refs[1].step(a1)
    assert (r0 == rr0.cpu().numpy()).all() and (d0 == dd0.cpu().numpy()).all()
    assert (r1 == rr1.cpu().numpy()).all() and (d1 == dd1.cpu().numpy()).all()


# ---------------------------------------------------------------------------------------------
# parity hardening
# ---------------------------------------------------------------------------------------------
def _run_vs_oracle(sl, lv, n, steps, seed, auto_reset, time_limit, env0=11, every=20, threads=8):
    pool = sl.game.LevelPool(lv)
    P = pool.n_levels
    env = sl.env.SafeLifeEnv(pool, num_envs=n, env_id0=env0, global_counter=None, compute_obs=False,
                             auto_reset=auto_reset, time_limit=time_limit)
    env.seed(seed)
    env.reset()
    lvl = (np.arange(n) + env0) % P
    o = slo.OracleBatchEnv(slo.OraclePool(**lv), lvl, seed=seed, env_id0=env0, threads=threads, time_limit=time_limit)
    rng = np.random.RandomState(seed + 1000)
    stride = n % P or P
    for t in range(steps):
        acts = rng.randint(9, size=n)
        _, r, d, _ = env.step(acts)
        orr, od = o.step(acts)
        assert (r.cpu().numpy() == orr).all() and (d.cpu().numpy() == (od != 0)).all(), t
        if auto_reset and od.any():
            idx = np.nonzero(od)[0]
            lvl[idx] = (lvl[idx] + stride) % P
            o.reset(idx, lvl[idx])
        if t % every == 0 or t == steps - 1:
            st, ost = env.game.state_numpy(), o.states
            assert (env.game.board.cpu().numpy() == o.boards).all(), t
            assert (env.game.goals.cpu().numpy() == o.goals).all(), t
            for f in st.dtype.names:
                assert (st[f] == ost[f]).all(), (f, t)
    return env, o


def test_big_board_fused_step_at_length(sl):
    """The 256x256 fused step (big_env_kernel: actions + physics + reward + exit paint) on dense random levels —
    crates, spawners, exits and life everywhere, so that pushes / pulls cross the 32-bit word seams and the torus
    edge, spawners draw every step and agents exit (auto-reset) — 24 envs x 128 steps, every env against the
    oracle: rewards and dones every step, full boards / goals / state every 16 steps."""
    rng = np.random.RandomState(256)
    lv = _random_levels(rng, 6, 256, 256, exits=40)
    # agents start on word seams / the torus edge in some levels
    for l, (x, y) in enumerate([(31, 5), (32, 255), (0, 0), (255, 128), (63, 64), (224, 31)]):
        ax, ay = lv["agent_loc"][l]
        lv["board"][l, ay, ax] = 0
        lv["board"][l, y, x] = 122 | (int(rng.randint(8)) << 9)
        lv["agent_loc"][l] = (x, y)
    _run_vs_oracle(sl, lv, 24, 128, seed=5, auto_reset=True, time_limit=50, every=16)


@pytest.mark.parametrize("h,w,spawner_goals", [(128, 128, False), (192, 128, True), (256, 64, True),
                                               (384, 64, False), (256, 256, True)])
def test_sliding_window_step_every_row_width(sl, h, w, spawner_goals):
    """The fused step of the sliding-window shapes (rows of 2 / 4 / 8 words; 2 or 3 rows per strip, so that a
    thread's share of items is not a multiple of the fix-up's batch of four), goals without spawners (score
    accumulated inside the goals' window) and with spawners (separate score pass), exits + auto-reset: rewards
    and dones every step, boards / goals / state every 8 steps, against the oracle."""
    rng = np.random.RandomState(h + w)
    lv = _random_levels(rng, 4, h, w, exits=12)
    if spawner_goals:
        m = rng.rand(*lv["goals"].shape)
        lv["goals"][m < 0.01] = 1688           # living spawner in the goals
        lv["goals"][(m >= 0.01) & (m < 0.02)] = 3712  # dead spawner, all colours
    _run_vs_oracle(sl, lv, 8, 40, seed=7, auto_reset=True, time_limit=17, every=8)


@pytest.mark.parametrize("shape", ["26x26", "256x256", "128x128"])
def test_back_to_back_launches_without_host_sync(sl, shape):
    """Dependent step launches queued back to back on one stream (no host synchronisation in between): the
    programmatic-dependent-launch kernels start before their predecessor has finished and must not read its
    outputs early.  Every step's rewards / dones and the final state against the oracle."""
    if shape == "26x26":
        lv, n, steps = gu.levels("append-spawn"), 4096, 48
    elif shape == "256x256":
        lv, n, steps = _random_levels(np.random.RandomState(31), 4, 256, 256, exits=8), 40, 16
    else:
        lv, n, steps = _random_levels(np.random.RandomState(32), 4, 128, 128, exits=8), 40, 16
    pool = sl.game.LevelPool(lv)
    P = pool.n_levels
    env = sl.env.SafeLifeEnv(pool, num_envs=n, env_id0=7, global_counter=None, compute_obs=False, auto_reset=False,
                             time_limit=1000)
    env.seed(8)
    env.reset()
    o = slo.OracleBatchEnv(slo.OraclePool(**lv), (np.arange(n) + 7) % P, seed=8, env_id0=7, threads=8, time_limit=1000)
    acts = np.random.RandomState(5).randint(9, size=(steps, n))
    dev_acts = torch.from_numpy(acts.astype(np.uint8)).cuda()
    torch.cuda.synchronize()
    rs, ds = [], []
    for t in range(steps):
        _, r, d, _ = env.step(dev_acts[t])
        rs.append(r)
        ds.append(d)
    torch.cuda.synchronize()
    got_r, got_d = torch.stack(rs).cpu().numpy(), torch.stack(ds).cpu().numpy()
    for t in range(steps):
        orr, od = o.step(acts[t])
        assert (got_r[t] == orr).all() and (got_d[t] == (od != 0)).all(), t
    assert (env.game.board.cpu().numpy() == o.boards).all()
    assert (env.game.goals.cpu().numpy() == o.goals).all()
    st = env.game.state_numpy()
    for f in st.dtype.names:
        assert (st[f] == o.states[f]).all(), f


@pytest.mark.parametrize("shape", ["26x26", "256x256"])
def test_two_batches_on_two_streams(sl, shape):
    """The actor pattern of INTEGRATION.md §4: two independent env batches stepped on two CUDA streams with their
    launches interleaved (and overlapping on the GPU); each batch equals the oracle, i.e. no kernel keeps state
    outside the buffers it is given."""
    if shape == "26x26":
        lv, n, steps = gu.levels("prune-spawn"), 2048, 40
    else:
        lv, n, steps = _random_levels(np.random.RandomState(77), 4, 256, 256, exits=8), 24, 12
    pool = sl.game.LevelPool(lv)
    P = pool.n_levels
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    envs, oracles = [], []
    for b in range(2):
        e = sl.env.SafeLifeEnv(pool, num_envs=n, env_id0=1000 * b, global_counter=None, compute_obs=False,
                               auto_reset=False, time_limit=1000)
        e.seed(3)
        e.reset()
        envs.append(e)
        oracles.append(slo.OracleBatchEnv(slo.OraclePool(**lv), (np.arange(n) + 1000 * b) % P, seed=3, env_id0=1000 * b,
                                          threads=8, time_limit=1000))
    torch.cuda.synchronize()
    rng = np.random.RandomState(9)
    acts = rng.randint(9, size=(steps, 2, n))
    dev_acts = torch.from_numpy(acts.astype(np.uint8)).cuda()
    rewards = [[], []]
    torch.cuda.synchronize()
    for t in range(steps):                      # no host synchronisation inside the loop
        for b in range(2):
            with torch.cuda.stream(streams[b]):
                _, r, d, _ = envs[b].step(dev_acts[t, b])
                rewards[b].append(r)
    torch.cuda.synchronize()
    for b in range(2):
        want = np.stack([oracles[b].step(acts[t, b])[0] for t in range(steps)])
        got = torch.stack(rewards[b]).cpu().numpy()
        assert (got == want).all(), b
        assert (envs[b].game.board.cpu().numpy() == oracles[b].boards).all(), b
        assert (envs[b].game.goals.cpu().numpy() == oracles[b].goals).all(), b
        st = envs[b].game.state_numpy()
        for f in st.dtype.names:
            assert (st[f] == oracles[b].states[f]).all(), (b, f)


def test_stepper_bound_outputs(sl):
    """slb_stepper_bind_outputs: with the outputs bound to one page-locked block the step writes them there
    directly; rewards / dones / state equal an env stepped the ordinary way, with other pointers the stepper works
    as before, and a block that is not page-locked (or not of the required shape) is refused."""
    lv = gu.levels("prune-spawn")
    pool = sl.game.LevelPool(lv)
    n = 192
    a = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False, auto_reset=True, time_limit=9)
    b = sl.env.SafeLifeEnv(pool, num_envs=n, global_counter=None, compute_obs=False, auto_reset=True, time_limit=9)
    for e in (a, b):
        e.seed(6); e.reset()
    st = torch.cuda.Stream()
    hs = a.host_stepper(st)
    out = torch.full((5 * n,), 0xAB, dtype=torch.uint8).pin_memory()
    other = np.zeros(5 * n, np.uint8)
    hs.bind_outputs(out.data_ptr(), out.data_ptr() + 4 * n)
    rng = np.random.RandomState(3)
    acts = torch.from_numpy(rng.randint(9, size=(30, n)).astype(np.uint8)).pin_memory()
    for t in range(30):
        if t % 5 == 4:   # other pointers: through the library's mirror, as before
            hs.step(acts[t].data_ptr(), other.ctypes.data, other.ctypes.data + 4 * n)
            hs.wait()
            r, d = other[:4 * n].view(np.int32), other[4 * n:]
        else:
            hs.step(acts[t].data_ptr(), out.data_ptr(), out.data_ptr() + 4 * n)
            hs.wait()
            r, d = out[:4 * n].view(torch.int32).numpy(), out[4 * n:].numpy()
        _, rb, db, _ = b.step(acts[t].numpy())
        assert (r == rb.cpu().numpy()).all() and (d == db.cpu().numpy()).all(), t
    assert (a.game.state == b.game.state).all() and (a.game.board == b.game.board).all()
    with pytest.raises(sl.lib.SlbError):
        hs.bind_outputs(other.ctypes.data, other.ctypes.data + 4 * n)              # pageable memory
    with pytest.raises(sl.lib.SlbError):
        hs.bind_outputs(out.data_ptr(), out.data_ptr() + 4 * n - 16)               # not one [reward | done] block
    hs.bind_outputs(None, None)                                                     # unbind
    hs.step(acts[0].data_ptr(), out.data_ptr(), out.data_ptr() + 4 * n)
    hs.wait()


@pytest.mark.parametrize("archive", gu.ARCHIVES)
def test_soak_every_archive(sl, archive):
    """Bounded soak: 64 envs x 300 steps of every v1.0 archive with and without auto-reset against the oracle,
    then the side-effect count maps of 8 of those episodes against the oracle's."""
    lv = gu.levels(archive)
    _run_vs_oracle(sl, lv, 64, 300, seed=1, auto_reset=True, time_limit=37, every=50)
    env, o = _run_vs_oracle(sl, lv, 64, 300, seed=2, auto_reset=False, time_limit=1000, every=50)
    g = env.game
    st = g.state_numpy()
    b0 = lv["board"][st["level_id"]]
    b1 = g.board.cpu().numpy()
    ctr0 = st["rng_counter"].astype(np.int64)
    ctr = torch.from_numpy(ctr0.copy()).cuda()
    spawn = lv["spawn_prob"][st["level_id"]].astype(np.float32)
    keys, counts, nkeys = sl.se.side_effect_counts(torch.from_numpy(b0).cuda(), g.board, st["num_steps"].astype(np.int32),
                                                   spawn, num_samples=1000, seed=2, env_id0=11, counters=ctr, kmax=32)
    keys, counts, nk, ctr = keys.cpu().numpy(), counts.cpu().numpy(), nkeys.cpu().numpy(), ctr.cpu().numpy()
    for i in range(0, 64, 8):
        ok, oc, octr = slo.side_effect_counts(b0[i], b1[i], int(st["num_steps"][i]), num_samples=1000,
                                              spawn_prob=float(spawn[i]), seed=2, env=11 + i, counter=int(ctr0[i]), kmax=32)
        mine = {int(k): counts[i, :, s] for s, k in enumerate(keys[i]) if k}
        assert nk[i] == len(ok) and ctr[i] == octr and sorted(mine) == sorted(int(k) for k in ok), i
        for s, k in enumerate(ok):
            assert (mine[int(k)] == oc[:, s]).all(), (i, int(k))


# ---------------------------------------------------------------------------------------------
# the EMD stage on the GPU (slb_emd_sizes / slb_emd_solve) against the oracle's two solvers
# ---------------------------------------------------------------------------------------------
def _emd_want(c0, c1, ns):
    from oracle import emd
    return emd.earth_mover_distance(c0.astype(np.float64) / ns, c1.astype(np.float64) / ns)


def test_emd_batch_golden_maps(sl):
    """the count maps the unmodified reference produced (tests/golden/side_effects.npz) through emd_batch:
    equal to the oracle's successive-shortest-path solver to 1e-12 (and that one equals the LP, test_emd.py)."""
    d = gu.load("side_effects.npz")
    i = 0
    while "keys_%d" % i in d:
        ks, cn, ns = d["keys_%d" % i], d["counts_%d" % i], int(d["num_samples_%d" % i])
        kmax = 8
        keys = np.zeros((1, kmax), np.uint16)
        counts = np.zeros((1, 2, kmax) + cn.shape[2:], np.uint16)
        keys[0, :len(ks)] = ks
        counts[0, :, :len(ks)] = cn
        got = sl.se.emd_batch(torch.from_numpy(keys).cuda(), torch.from_numpy(counts).cuda(), ns).cpu().numpy()
        for s in range(kmax):
            want = _emd_want(cn[0, s], cn[1, s], ns) if s < len(ks) else 0.0
            assert abs(got[0, s] - want) <= 1e-12 * max(1.0, abs(want)), (i, s, got[0, s], want)
        i += 1
    assert i >= 5


@pytest.mark.parametrize("h,w", [(26, 26), (9, 13), (32, 32)])
def test_emd_batch_random_maps(sl, h, w):
    """random count maps: sparse and dense differences (hundreds of changed cells), overlapping bins, identical
    maps, one-sided maps, empty slots — every problem against the oracle."""
    rng = np.random.RandomState(h * w)
    n, kmax, ns = 12, 5, 1000
    keys = np.zeros((n, kmax), np.uint16)
    counts = np.zeros((n, 2, kmax, h, w), np.uint16)
    for i in range(n):
        for s in range(kmax):
            kind = rng.randint(7)
            if kind == 0:
                continue  # empty slot
            keys[i, s] = 9 + s
            cells = int(rng.choice([1, 3, 10, 40, min(h * w, 300)]))
            idx = rng.choice(h * w, cells, replace=False)
            a = np.zeros(h * w, np.uint16)
            a[idx] = rng.randint(1, ns + 1, cells)
            b = a.copy()
            if kind == 1:
                pass  # identical maps
            elif kind == 2:
                b[:] = 0  # everything is surplus
            else:
                move = rng.rand(cells) < 0.7
                b[idx[move]] = 0
                j = rng.choice(h * w, int(move.sum()))
                b[j] = rng.randint(1, ns + 1, j.size)
            counts[i, 0, s] = a.reshape(h, w)
            counts[i, 1, s] = b.reshape(h, w)
    got = sl.se.emd_batch(torch.from_numpy(keys).cuda(), torch.from_numpy(counts).cuda(), ns).cpu().numpy()
    for i in range(n):
        for s in range(kmax):
            want = _emd_want(counts[i, 0, s], counts[i, 1, s], ns) if keys[i, s] else 0.0
            assert abs(got[i, s] - want) <= 1e-12 * max(1.0, abs(want)), (i, s, got[i, s], want)


def test_side_effect_score_uses_the_gpu_emd(sl):
    """side_effect_score end to end: keys and mean inaction counts as before, the EMD column from emd_batch equal
    to the oracle's value for the same count maps."""
    lv = gu.level_slice(gu.levels("prune-spawn"), 0, 6)
    env = sl.env.SafeLifeEnv(sl.game.LevelPool(lv), num_envs=6, global_counter=None, compute_obs=False)
    env.seed(4)
    env.reset()
    rng = np.random.RandomState(4)
    for _ in range(40):
        env.step(rng.randint(9, size=6))
    g = env.game
    st = g.state_numpy()
    ctr0 = st["rng_counter"].astype(np.int64)
    b1 = g.board.cpu().numpy()
    res = sl.se.side_effect_score(g, num_samples=200)
    assert len(res) == 6
    for i in range(6):
        ok, oc, _ = slo.side_effect_counts(lv["board"][st["level_id"][i]], b1[i], int(st["num_steps"][i]),
                                           num_samples=200, spawn_prob=float(np.float32(lv["spawn_prob"][st["level_id"][i]])),
                                           seed=4, env=i, counter=int(ctr0[i]), kmax=32)
        assert sorted(int(k) for k in res[i]) == sorted(int(k) for k in ok), i
        for s, k in enumerate(ok):
            want = _emd_want(oc[0, s], oc[1, s], 200)
            got, mean = res[i][np.uint16(k)]
            assert abs(got - want) <= 1e-12 * max(1.0, abs(want)), (i, int(k))
            assert mean == np.sum(oc[0, s].astype(np.float64) / 200), (i, int(k))


def test_checkpoint_resume_is_bit_exact(sl, tmp_path):
    """state_dict / load_state_dict: a batch resumed from a checkpoint continues bit-exactly (boards,
    goals, state incl. spawn-stream counters, rewards, dones, global counter); serialize / save give the
    reference's saved-game format."""
    lv = gu.levels("prune-spawn")
    pool = sl.game.LevelPool(lv)
    rng = np.random.RandomState(8)
    acts = rng.randint(9, size=(60, 200))

    def make():
        e = sl.env.SafeLifeEnv(pool, num_envs=200, env_id0=7, global_counter=sl.env.GlobalCounter(), compute_obs=False,
                               auto_reset=True, time_limit=25)
        e.seed(5)
        e.reset()
        return e

    a = make()
    for t in range(30):
        a.step(acts[t])
    ck = a.state_dict()
    torch.save(ck, str(tmp_path / "ck.pt"))
    b = make()
    b.load_state_dict(torch.load(str(tmp_path / "ck.pt"), weights_only=False))
    for t in range(30, 60):
        _, ra, da, _ = a.step(acts[t])
        _, rb, db, _ = b.step(acts[t])
        assert (ra == rb).all() and (da == db).all(), t
    assert (a.game.state == b.game.state).all() and (a.game.board_res == b.game.board_res).all()
    assert (a.game.goals_res == b.game.goals_res).all()
    assert a.global_counter.num_steps == b.global_counter.num_steps == 60 * 200
    assert a.global_counter.episodes_completed == b.global_counter.episodes_completed
    # the reference's saved-game format for one env
    f = a.game.save(str(tmp_path / "env3"), 3)
    with np.load(f) as z:
        assert (z["board"] == a.game.board[3].cpu().numpy()).all() and (z["goals"] == a.game.goals[3].cpu().numpy()).all()
        st = a.game.state_numpy()[3]
        assert tuple(z["agent_loc"]) == (st["agent_x"], st["agent_y"]) and int(z["orientation"]) == st["orientation"]
        assert float(z["min_performance"]) == st["min_performance"]


def test_sharded_env_equals_its_slice_of_the_whole_job(sl):
    """SafeLifeEnv.sharded: rank r of w plays envs shard_range(total, r, w) of the job, with the same results
    as those envs have in the unsharded batch."""
    from safelife_k2_b200 import distributed
    lv = gu.levels("append-spawn")
    pool = sl.game.LevelPool(lv)
    total = 300
    whole = sl.env.SafeLifeEnv(pool, num_envs=total, global_counter=None, compute_obs=False)
    whole.seed(9)
    whole.reset()
    parts = []
    for r in range(4):
        e = sl.env.SafeLifeEnv.sharded(pool, total, rank=r, world_size=4, global_counter=None, compute_obs=False)
        assert e.shard[:2] == distributed.shard_range(total, r, 4)
        e.seed(9)
        e.reset()
        parts.append(e)
    rng = np.random.RandomState(9)
    for t in range(25):
        acts = rng.randint(9, size=total)
        _, rw, dw, _ = whole.step(acts)
        for e in parts:
            lo, hi, _ = e.shard
            _, r, d, _ = e.step(acts[lo:hi])
            assert (r == rw[lo:hi]).all() and (d == dw[lo:hi]).all(), (t, lo)
    for e in parts:
        lo, hi, _ = e.shard
        assert (e.game.state == whole.game.state[lo:hi]).all()
        assert (e.game.board == whole.game.board[lo:hi]).all()
