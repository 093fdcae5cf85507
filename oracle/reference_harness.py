"""
Runs the UNMODIFIED reference (neale/safelife-k2 @ /root/reference) in the build container.
TEST INFRASTRUCTURE ONLY — used by tests/test_oracle_vs_reference.py and by the golden-vector
generators under tests/golden/.  /root/reference does not exist on the GPU box; nothing that
runs there may import this module.

RNG alignment: the reference extension refills its 10 000-double buffer by calling
`numpy.random.random(10000)` looked up BY NAME at every refill (speedups_src/random.c:18-20).  We
assign `numpy.random.random = feeder`, where the feeder serves the build's counter-based per-env
stream, and call `speedups.seed(x)` once to flush the old buffer.  advance_board.c stays
byte-identical.
"""
import os
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE_ROOT = os.environ.get("SAFELIFE_REFERENCE", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "safelife")) and os.path.isdir(
        os.path.join(_HERE, "_ref", "pyext")
    )


def _import():
    p = os.path.join(_HERE, "refpkg")
    if p not in sys.path:
        sys.path.insert(0, p)
    import safelife  # noqa: F401  (refpkg shim -> reference modules + compiled extension)
    return safelife


# ---------------------------------------------------------------------------------------------
# numpy statement of the stream (independent of the C oracle; cross-checked in tests)
# ---------------------------------------------------------------------------------------------
def philox4x32_10_np(c0, c1, c2, c3, k0, k1):
    c0, c1, c2, c3 = (np.asarray(x, dtype=np.uint64) & 0xFFFFFFFF for x in (c0, c1, c2, c3))
    k0 = np.uint64(k0) & np.uint64(0xFFFFFFFF)
    k1 = np.uint64(k1) & np.uint64(0xFFFFFFFF)
    M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
    W0, W1 = np.uint64(0x9E3779B9), np.uint64(0xBB67AE85)
    mask = np.uint64(0xFFFFFFFF)
    s32 = np.uint64(32)
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        n0 = (p1 >> s32) ^ c1 ^ k0
        n2 = (p0 >> s32) ^ c3 ^ k1
        c1 = p1 & mask
        c3 = p0 & mask
        c0, c2 = n0, n2
        k0 = (k0 + W0) & mask
        k1 = (k1 + W1) & mask
    return c0, c1, c2, c3


def stream_u53_np(seed, env, index):
    index = np.asarray(index, dtype=np.uint64)
    seed, env = int(seed), int(env)
    x0, x1, _, _ = philox4x32_10_np(
        index & np.uint64(0xFFFFFFFF), index >> np.uint64(32),
        np.full(index.shape, env & 0xFFFFFFFF, dtype=np.uint64),
        np.full(index.shape, env >> 32, dtype=np.uint64),
        seed & 0xFFFFFFFF, seed >> 32,
    )
    return ((x0 >> np.uint64(5)) << np.uint64(26)) | (x1 >> np.uint64(6))


def stream_doubles_np(seed, env, start, n):
    idx = np.arange(int(start), int(start) + int(n), dtype=np.uint64)
    return stream_u53_np(seed, env, idx).astype(np.float64) / 9007199254740992.0


class StreamFeeder(object):
    """Drop-in for numpy.random.random(n): serves env's stream sequentially."""

    def __init__(self, seed, env, start=0):
        self.seed, self.env, self.pos = int(seed), int(env), int(start)
        self.calls = []

    def __call__(self, size=None):
        n = 1 if size is None else int(np.prod(size))
        out = stream_doubles_np(self.seed, self.env, self.pos, n)
        self.pos += n
        self.calls.append(n)
        if size is None:
            return float(out[0])
        return np.ascontiguousarray(out.reshape(size))


class fed_reference(object):
    """Context manager: inside, the reference extension draws from (seed, env)'s stream starting
    at draw index `start`.  On exit numpy.random.random is restored."""

    def __init__(self, seed, env, start=0):
        self.feeder = StreamFeeder(seed, env, start)

    def __enter__(self):
        sl = _import()
        self._orig = np.random.random
        np.random.random = self.feeder
        sl.speedups.seed(0)  # numpy.random.seed(0) + flush -> refill through the feeder
        return self.feeder

    def __exit__(self, *exc):
        np.random.random = self._orig
        return False

    def rewind(self, start=0):
        """Re-align: the extension's next draw becomes draw index `start` (anything that called
        speedups.seed() in between — e.g. SafeLifeEnv.__init__ -> seed() — flushed the buffer)."""
        self.feeder.pos = int(start)
        _import().speedups.seed(0)


# ---------------------------------------------------------------------------------------------
# reference entry points
# ---------------------------------------------------------------------------------------------
def speedups():
    return _import().speedups


def advance_board(board, spawn_prob=0.3):
    return speedups().advance_board(board, spawn_prob)


def load_archive(name):
    """name like 'benchmarks/v1.0/append-still' -> structured array of levels (reference loader
    path resolution, file_finder.py:78-98)."""
    path = os.path.join(REFERENCE_ROOT, "safelife", "levels", name + ".npz")
    with np.load(path) as data:
        return data["levels"].copy()


def load_level_files(folder):
    """Single-level npz files of a folder (patterns/examples/mazes/...) -> list of (name, dict)."""
    d = os.path.join(REFERENCE_ROOT, "safelife", "levels", folder)
    out = []
    for fn in sorted(os.listdir(d)):
        if fn.endswith(".npz"):
            with np.load(os.path.join(d, fn)) as data:
                out.append((fn[:-4], {k: data[k] for k in data.keys()}))
    return out


def make_env(archive=None, levels=None, **env_kwargs):
    """A real reference SafeLifeEnv over a reference archive (or an iterable of level records)."""
    _import()
    from safelife.safelife_env import SafeLifeEnv
    from safelife.safelife_game import SafeLifeGame

    if levels is None:
        levels = load_archive(archive)

    def it():
        while True:
            for rec in levels:
                yield SafeLifeGame.loaddata(rec)

    return SafeLifeEnv(it(), **env_kwargs)


def run_env_trace(levels, actions, seed, env_id, time_limit=1000, record_every=1):
    """Steps ONE real reference SafeLifeEnv through `actions` on the (seed, env_id) stream.
    Returns dict of per-step arrays.  Auto-resets (next level) on the step after done, like the
    batched env's auto-reset."""
    ctx = fed_reference(seed, env_id)
    with ctx:
        env = make_env(levels=levels, time_limit=time_limit, global_counter=None)
        env.reset()
        ctx.rewind(0)  # SafeLifeEnv.__init__ called seed() -> flushed; restart the stream at 0
        T = len(actions)
        H, W = env.game.board.shape
        out = {
            "board": np.zeros((T, H, W), np.uint16), "goals": np.zeros((T, H, W), np.uint16),
            "reward": np.zeros(T, np.int64), "done": np.zeros(T, np.uint8),
            "times_up": np.zeros(T, np.uint8),
            "agent_loc": np.zeros((T, 2), np.int32), "orientation": np.zeros(T, np.int32),
            "game_over": np.zeros(T, np.uint8), "points": np.zeros(T, np.int64),
            "completed": np.zeros(T, np.int64), "possible": np.zeros(T, np.int64),
            "board0": env.game.board.copy(), "goals0": env.game.goals.copy(),
        }
        for t, a in enumerate(actions):
            obs, r, done, info = env.step(int(a))
            out["board"][t] = info["board"]
            out["goals"][t] = info["goals"]
            out["reward"][t] = r
            out["done"][t] = bool(done)
            out["times_up"][t] = bool(info["times_up"])
            out["agent_loc"][t] = env.game.agent_loc
            out["orientation"][t] = env.game.orientation
            out["game_over"][t] = bool(env.game.game_over)
            out["points"][t] = env._old_game_value
            c, p = env.game.performance_ratio()
            out["completed"][t], out["possible"][t] = c, p
            if done:
                env.reset()
        return out


def side_effect_pre_emd(game, num_samples=1000):
    """The reference's side_effect_score with the EMD call replaced by a recorder.
    Returns {key: (inaction_density[H,W] f64, action_density[H,W] f64, sum_inaction f64)}."""
    _import()
    from safelife import side_effects

    captured = []

    def recorder(a, b, *args, **kw):
        captured.append((np.array(a, dtype=float), np.array(b, dtype=float)))
        return 0.0

    orig = side_effects.earth_mover_distance
    side_effects.earth_mover_distance = recorder
    try:
        scores = side_effects.side_effect_score(game, num_samples=num_samples)
    finally:
        side_effects.earth_mover_distance = orig
    out = {}
    for (key, (_, mean_count)), (a, b) in zip(scores.items(), captured):
        out[int(key)] = (a, b, float(mean_count))
    return out
