"""
ctypes bindings for the CPU oracle.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The product package (safelife_k2_b200) never does.

  * libslo_oracle.so  (oracle/_build)  — our C restatement, oracle/slo_oracle.c
  * libref_advance.so (oracle/_ref)    — the reference's own speedups_src/advance_board.c compiled
                                         in place + our stream shim (oracle/refshim/stream_random.c)
"""
import ctypes as C
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PORT = os.path.join(_HERE, "_build", "libslo_oracle.so")
_REF = os.path.join(_HERE, "_ref", "libref_advance.so")

u16p = C.POINTER(C.c_uint16)


class Stream(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("env", C.c_uint64), ("counter", C.c_uint64)]


ENV_STATE_DTYPE = np.dtype(
    [
        ("agent_x", "<i4"), ("agent_y", "<i4"), ("orientation", "<i4"), ("game_over", "<i4"),
        ("num_steps", "<i4"), ("level_id", "<i4"), ("episode_length", "<i4"),
        ("episode_completed", "<i4"),
        ("old_points", "<i4"), ("can_exit", "<i4"), ("completed", "<i4"), ("possible", "<i4"),
        ("episode_reward", "<i8"), ("rng_counter", "<u8"),
        ("min_performance", "<f8"), ("flags", "<u4"), ("reserved", "<u4"),
    ]
)
assert ENV_STATE_DTYPE.itemsize == 80


class Pool(C.Structure):
    _fields_ = [
        ("n_levels", C.c_int32), ("h", C.c_int32), ("w", C.c_int32), ("pad", C.c_int32),
        ("board", C.c_void_p), ("goals", C.c_void_p),
        ("agent_x", C.c_void_p), ("agent_y", C.c_void_p), ("orientation", C.c_void_p),
        ("spawn_prob", C.c_void_p), ("min_performance", C.c_void_p),
    ]


def build(force=False):
    """Compile the oracle port (always) and oracle/_ref (only where /root/reference exists)."""
    srcs = [os.path.join(_HERE, f) for f in ("slo_oracle.c", "slo_oracle.h", "emd_oracle.c")]
    if force or not os.path.exists(_PORT) or os.path.getmtime(_PORT) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", _HERE, "port"], stdout=subprocess.DEVNULL)
    if os.path.isdir(os.environ.get("SAFELIFE_REFERENCE", "/root/reference")):
        subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PORT):
            build()
        L = C.CDLL(_PORT)
        L.slo_stream_u53.restype = C.c_uint64
        L.slo_stream_u53.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64]
        L.slo_stream_fill.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_int64, C.c_void_p]
        L.slo_philox4x32_10.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.slo_advance_board.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float,
                                        C.POINTER(Stream)]
        L.slo_execute_action.restype = C.c_int
        L.slo_execute_action.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int]
        L.slo_current_points.restype = C.c_int64
        L.slo_current_points.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.slo_baseline_score.restype = C.c_int64
        L.slo_baseline_score.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.slo_performance_ratio.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64,
                                            C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.slo_can_exit.restype = C.c_int
        L.slo_can_exit.argtypes = [C.c_double, C.c_int64, C.c_int64]
        L.slo_update_exit_colors.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.slo_env_reset_batch.argtypes = [C.POINTER(Pool), C.c_int, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p]
        L.slo_env_step_batch.argtypes = [C.POINTER(Pool), C.c_int, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_int, C.c_uint64, C.c_uint64,
                                         C.c_void_p, C.c_void_p, C.c_int]
        L.slo_advance_board_batch.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                              C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p,
                                              C.c_int]
        L.slo_side_effect_counts.restype = C.c_int
        L.slo_side_effect_counts.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                             C.c_int, C.c_float, C.c_uint64, C.c_uint64,
                                             C.POINTER(C.c_uint64), C.c_int, C.c_void_p,
                                             C.c_void_p]
        L.slo_set_reference_advance.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.slo_emd_transport.restype = C.c_double
        L.slo_emd_transport.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def have_ref():
    return os.path.exists(_REF)


def ref():
    """The reference's own advance_board.c (compiled in place) behind the stream shim."""
    global _ref
    if _ref is None:
        R = C.CDLL(_REF)
        R.advance_board.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_float]
        R.ref_stream_set.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64]
        R.ref_stream_get_counter.restype = C.c_uint64
        _ref = R
    return _ref


def use_reference_advance(enable=True):
    """Route the physics of slo_env_step / slo_side_effect_counts / *_batch through the
    reference's advance_board.c (bench `--impl reference` arm, oracle cross-check)."""
    L = lib()
    if enable:
        R = ref()
        L.slo_set_reference_advance(
            C.cast(R.advance_board, C.c_void_p),
            C.cast(R.ref_stream_set, C.c_void_p),
            C.cast(R.ref_stream_get_counter, C.c_void_p),
        )
    else:
        L.slo_set_reference_advance(None, None, None)


def _c16(a):
    a = np.ascontiguousarray(a, dtype=np.uint16)
    return a


# ---------------------------------------------------------------------------------------------
# stream
# ---------------------------------------------------------------------------------------------
def stream_doubles(seed, env, start, n):
    out = np.empty(int(n), dtype=np.float64)
    lib().slo_stream_fill(seed, env, start, int(n), out.ctypes.data)
    return out


def philox(ctr, key):
    ctr = np.asarray(ctr, dtype=np.uint32)
    key = np.asarray(key, dtype=np.uint32)
    out = np.zeros(4, dtype=np.uint32)
    lib().slo_philox4x32_10(ctr.ctypes.data, key.ctypes.data, out.ctypes.data)
    return out


def threshold_u53(spawn_prob):
    """Smallest integer T such that  u / 2**53 < (double)(float)p  <=>  u < T  for integer u."""
    from fractions import Fraction
    import math

    t = Fraction(float(np.float32(spawn_prob))) * (1 << 53)
    T = math.ceil(t)
    return max(0, min(T, 1 << 53))


# ---------------------------------------------------------------------------------------------
# single-board functions
# ---------------------------------------------------------------------------------------------
def advance_board(board, spawn_prob=0.3, seed=0, env=0, counter=0, use_ref=False):
    """-> (new_board, new_counter)"""
    b = _c16(board)
    h, w = b.shape
    out = np.empty_like(b)
    if use_ref:
        R = ref()
        src = b.copy()
        R.ref_stream_set(seed, env, counter)
        R.advance_board(src.ctypes.data, out.ctypes.data, h, w, C.c_float(spawn_prob))
        return out, int(R.ref_stream_get_counter())
    st = Stream(seed, env, counter)
    lib().slo_advance_board(b.ctypes.data, out.ctypes.data, h, w, C.c_float(spawn_prob),
                            C.byref(st))
    return out, int(st.counter)


def execute_action(board, state, action, can_exit):
    """board modified in place; state = 1-element ENV_STATE_DTYPE array; -> reward"""
    assert board.dtype == np.uint16 and board.flags.c_contiguous
    h, w = board.shape
    return lib().slo_execute_action(board.ctypes.data, h, w, state.ctypes.data, int(action),
                                    int(can_exit))


def current_points(board, goals):
    b, g = _c16(board), _c16(goals)
    return int(lib().slo_current_points(b.ctypes.data, g.ctypes.data, *b.shape))


def baseline_score(board0, goals0):
    b, g = _c16(board0), _c16(goals0)
    return int(lib().slo_baseline_score(b.ctypes.data, g.ctypes.data, *b.shape))


def performance_ratio(board, goals, baseline):
    b, g = _c16(board), _c16(goals)
    c, p = C.c_int64(), C.c_int64()
    lib().slo_performance_ratio(b.ctypes.data, g.ctypes.data, b.shape[0], b.shape[1], baseline,
                                C.byref(c), C.byref(p))
    return int(c.value), int(p.value)


def can_exit(min_performance, completed, possible):
    return bool(lib().slo_can_exit(float(min_performance), completed, possible))


def side_effect_counts(board0, board1, num_steps, num_samples=1000, spawn_prob=0.3, seed=0, env=0,
                       counter=0, kmax=32):
    """-> (keys[K], counts[2,K,H,W] uint16, new_counter); raises if more than kmax keys."""
    b0, b1 = _c16(board0), _c16(board1)
    h, w = b0.shape
    keys = np.zeros(kmax, dtype=np.uint16)
    counts = np.zeros((2, kmax, h, w), dtype=np.uint16)
    ctr = C.c_uint64(counter)
    k = lib().slo_side_effect_counts(b0.ctypes.data, b1.ctypes.data, h, w, int(num_steps),
                                     int(num_samples), C.c_float(spawn_prob), seed, env,
                                     C.byref(ctr), kmax, keys.ctypes.data, counts.ctypes.data)
    if k < 0:
        raise RuntimeError("more than %d distinct cell types" % kmax)
    return keys[:k].copy(), counts[:, :k].copy(), int(ctr.value)


# ---------------------------------------------------------------------------------------------
# batched env (struct-of-arrays level pool + [N] state)
# ---------------------------------------------------------------------------------------------
class OraclePool(object):
    """Level pool in the oracle's layout. Built from dict-of-arrays (see tests/golden/make_*.py)."""

    def __init__(self, board, goals, agent_loc, orientation, spawn_prob, min_performance):
        self.board = _c16(board)
        self.goals = _c16(goals)
        P, self.h, self.w = self.board.shape
        self.n_levels = P
        al = np.asarray(agent_loc)
        self.agent_x = np.ascontiguousarray(al[:, 0], dtype=np.int32)
        self.agent_y = np.ascontiguousarray(al[:, 1], dtype=np.int32)
        self.orientation = np.ascontiguousarray(orientation, dtype=np.int32)
        # python float -> C float at the extension boundary (module.c:22-24)
        self.spawn_prob = np.ascontiguousarray(spawn_prob, dtype=np.float32)
        self.min_performance = np.ascontiguousarray(min_performance, dtype=np.float64)
        self.c = Pool(P, self.h, self.w, 0, self.board.ctypes.data, self.goals.ctypes.data,
                      self.agent_x.ctypes.data, self.agent_y.ctypes.data,
                      self.orientation.ctypes.data, self.spawn_prob.ctypes.data,
                      self.min_performance.ctypes.data)


class OracleBatchEnv(object):
    """N reference-semantics envs stepped on the CPU (oracle port; optionally the reference's own
    advance_board.c for the physics).  Threads over env slices via ctypes (GIL released)."""

    def __init__(self, pool, level_ids, seed=0, env_id0=0, time_limit=1000, threads=1):
        self.pool = pool
        self.n = len(level_ids)
        self.seed, self.env_id0, self.time_limit = seed, env_id0, time_limit
        self.threads = max(1, int(threads))
        self.boards = np.zeros((self.n, pool.h, pool.w), dtype=np.uint16)
        self.goals = np.zeros_like(self.boards)
        self.states = np.zeros(self.n, dtype=ENV_STATE_DTYPE)
        self._tp = ThreadPoolExecutor(self.threads) if self.threads > 1 else None
        self.reset(np.arange(self.n), level_ids)

    def reset(self, env_idx, level_ids):
        env_idx = np.asarray(env_idx)
        lv = np.ascontiguousarray(level_ids, dtype=np.int32)
        for i, l in zip(env_idx, lv):
            one = np.array([l], dtype=np.int32)
            lib().slo_env_reset_batch(C.byref(self.pool.c), 1, one.ctypes.data,
                                      self.boards[i].ctypes.data, self.goals[i].ctypes.data,
                                      self.states[i:i + 1].ctypes.data)

    def _step_slice(self, lo, hi, actions, rewards, dones):
        lib().slo_env_step_batch(C.byref(self.pool.c), hi - lo, self.boards[lo:hi].ctypes.data,
                                 self.goals[lo:hi].ctypes.data, self.states[lo:hi].ctypes.data,
                                 actions[lo:hi].ctypes.data, self.time_limit, self.seed,
                                 self.env_id0 + lo, rewards[lo:hi].ctypes.data,
                                 dones[lo:hi].ctypes.data, 1)

    def step(self, actions):
        actions = np.ascontiguousarray(actions, dtype=np.int32)
        rewards = np.zeros(self.n, dtype=np.int64)
        dones = np.zeros(self.n, dtype=np.uint8)
        if self._tp is None:
            self._step_slice(0, self.n, actions, rewards, dones)
        else:
            edges = np.linspace(0, self.n, self.threads + 1).astype(int)
            futs = [self._tp.submit(self._step_slice, int(a), int(b), actions, rewards, dones)
                    for a, b in zip(edges[:-1], edges[1:]) if b > a]
            for f in futs:
                f.result()
        return rewards, dones


# ---------------------------------------------------------------------------------------------
# observation (numpy restatement; pinned by tests/golden/obs.npz)
# ---------------------------------------------------------------------------------------------
def get_obs(board, goals, agent_loc, board0, view_shape=(15, 15), remove_white_goals=True,
            output_channels=tuple(range(15))):
    """SafeLifeEnv.get_obs (safelife_env.py:125-155) + recenter_view (helper_utils.py:41-74) for one
    env.  agent_loc = (x, y); board0 = the level's initial board (its exit cells are exit_locs,
    safelife_game.py:528-529)."""
    board = np.array(board, dtype=np.uint16)
    g = np.asarray(goals, dtype=np.uint16) & np.uint16(7 << 9)             # :134
    if remove_white_goals:
        g = g * (g != np.uint16(7 << 9)).astype(np.uint16)                 # :139-140
    board = (board + (g << np.uint16(3))).astype(np.uint16)                # :143, uint16 wrap
    h, w = view_shape
    bh, bw = board.shape
    x0, y0 = int(agent_loc[0]), int(agent_loc[1])
    x1, y1 = x0 - w // 2, y0 - h // 2                                      # helper_utils.py:58-59
    ys = (y1 + np.arange(h)) % bh
    xs = (x1 + np.arange(w)) % bw
    view = board[np.ix_(ys, xs)].copy()                                    # wrapping_array slice :60
    iy, ix = np.nonzero(np.asarray(board0, dtype=np.uint16) & np.uint16(256))  # exit_locs
    jy = (iy - y0 + bh // 2) % bh - bh // 2                                # :66-67
    jx = (ix - x0 + bw // 2) % bw - bw // 2
    jy = np.clip(jy + h // 2, 0, h - 1)                                    # :69-70
    jx = np.clip(jx + w // 2, 0, w - 1)
    for k in range(len(iy)):                                               # :72, last write wins
        view[jy[k], jx[k]] = board[iy[k], ix[k]]
    if output_channels:                                                    # :152-154
        shift = np.array(list(output_channels), dtype=np.uint16)
        view = (view[..., None] & (np.uint16(1) << shift)) >> shift
    return view


def side_effect_penalty_count(board, board0, goals):
    """SimpleSideEffectPenalty.step's `side_effect` (env_wrappers.py:319-346) for one env, numpy."""
    player = np.uint16(2 | 8 | 16 | 32 | 64)                                # CellTypes.player
    b = np.asarray(board, dtype=np.uint16) & ~player                        # :326
    s = np.asarray(board0, dtype=np.uint16) & ~player                       # :327
    ex = (np.asarray(board0, dtype=np.uint16) & np.uint16(256)) != 0        # exit_locs
    b = np.where(ex, s, b)                                                  # :329-330
    red_life = np.uint16(1 | 512)
    start_red = (s & red_life) == red_life                                  # :335
    end_red = (b & red_life) == red_life
    goal_cell = (np.asarray(goals, dtype=np.uint16) & np.uint16(7 << 9)) == np.uint16(1 << 11)
    end_alive = (b & red_life) == np.uint16(1)
    non_effects = (b == s) | (start_red & ~end_red) | (goal_cell & end_alive)   # :340
    return int(np.sum(~non_effects))
