"""
Batched SafeLife game state on one B200 — host-side mirror of the reference's
``safelife/safelife_game.py`` for the hot path (same names, argument meaning and behaviour), with
every method operating on N lockstep boards through libsafelife_b200 (CUDA, sm_100a).

Reference: CellTypes safelife_game.py:37-120; GameState :123-537 (fields :155-168,
execute_action :347-393, can_exit :522-526, update_exit_colors :531-537, revert :229-234);
GameWithGoals :540-647 (point_table :554-565, current_points :590-599, performance_ratio
:601-631); SafeLifeGame.advance_board :657-660.

Out of scope here: level editing (execute_edit), GameOfLife / AsyncGame.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib

ORIENTATION = {"UP": 0, "RIGHT": 1, "DOWN": 2, "LEFT": 3, "FORWARD": 4, "BACKWARD": 6}

# env action table (safelife_env.py:61-71)
ACTION_NAMES = (
    "NULL", "MOVE UP", "MOVE RIGHT", "MOVE DOWN", "MOVE LEFT",
    "TOGGLE UP", "TOGGLE RIGHT", "TOGGLE DOWN", "TOGGLE LEFT",
)
ACTION_INDEX = {name: i for i, name in enumerate(ACTION_NAMES)}


class CellTypes(object):
    """uint16 cell bit flags (reference safelife_game.py:74-120)."""
    alive_bit, agent_bit, pushable_bit, destructible_bit = 0, 1, 2, 3
    frozen_bit, preserving_bit, inhibiting_bit, spawning_bit = 4, 5, 6, 7
    exit_bit, color_bit, pullable_bit = 8, 9, 15

    alive = np.uint16(1 << 0)
    agent = np.uint16(1 << 1)
    pushable = np.uint16(1 << 2)
    destructible = np.uint16(1 << 3)
    frozen = np.uint16(1 << 4)
    preserving = np.uint16(1 << 5)
    inhibiting = np.uint16(1 << 6)
    spawning = np.uint16(1 << 7)
    exit = np.uint16(1 << 8)
    color_r = np.uint16(1 << 9)
    color_g = np.uint16(1 << 10)
    color_b = np.uint16(1 << 11)
    pullable = np.uint16(1 << 15)

    empty = np.uint16(0)
    freezing = inhibiting | preserving
    player = agent | freezing | frozen | destructible  # 122
    wall = frozen
    movable = pushable | pullable
    crate = frozen | movable
    spawner = frozen | spawning | destructible
    hard_spawner = frozen | spawning
    level_exit = frozen | exit
    life = alive | destructible
    colors = (color_r, color_g, color_b)
    rainbow_color = color_r | color_g | color_b
    ice_cube = frozen | freezing | movable
    plant = frozen | alive | movable
    tree = frozen | alive
    fountain = preserving | frozen
    parasite = inhibiting | alive | pushable | frozen
    weed = preserving | alive | pushable | frozen
    powers = alive | freezing | spawning


ENV_STATE_DTYPE = np.dtype([
    ("agent_x", "<i4"), ("agent_y", "<i4"), ("orientation", "<i4"), ("game_over", "<i4"),
    ("num_steps", "<i4"), ("level_id", "<i4"), ("episode_length", "<i4"),
    ("episode_completed", "<i4"), ("old_points", "<i4"), ("can_exit", "<i4"),
    ("completed", "<i4"), ("possible", "<i4"), ("episode_reward", "<i8"), ("rng_counter", "<u8"),
    ("min_performance", "<f8"), ("flags", "<u4"), ("reserved", "<u4"),
])
assert ENV_STATE_DTYPE.itemsize == 80
STATE_WORDS = ENV_STATE_DTYPE.itemsize // 4
# int32 column index of each field in the [N,20] int32 view of the state tensor
_COL = {name: ENV_STATE_DTYPE.fields[name][1] // 4 for name in ENV_STATE_DTYPE.names}


def _dev(device):
    return torch.device(device if device is not None else "cuda")


class LevelPool(object):
    """Device-resident level pool: the `_init_data` of every level plus per-level scalars.

    ``levels`` may be: a structured array with the reference's level record dtype
    (file_finder.py:78-98: fields board, goals, agent_loc, orientation, spawn_prob,
    min_performance), a list of dict-like single-level records (GameState.serialize format,
    safelife_game.py:184-194), or a dict of stacked arrays with those keys.
    """

    def __init__(self, levels, device=None):
        rec = _stack_levels(levels)
        self.device = _dev(device)
        self.host = rec
        self.names = _level_names(levels)
        P, H, W = rec["board"].shape
        if H < 3 or W < 3:
            raise ValueError("boards must be at least 3x3 (got %dx%d)" % (H, W))
        self.n_levels, self.h, self.w = P, H, W
        self.words_per_row = (W + 31) // 32
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(self.device)  # noqa: E731
        self.board = t(rec["board"])
        self.goals = t(rec["goals"])
        self.agent_x = t(rec["agent_loc"][:, 0].astype(np.int32))
        self.agent_y = t(rec["agent_loc"][:, 1].astype(np.int32))
        self.orientation = t(rec["orientation"].astype(np.int32))
        # python float -> C float at the extension boundary (speedups_src/module.c:22-24)
        self.spawn_prob = t(rec["spawn_prob"].astype(np.float32))
        self.min_performance = t(rec["min_performance"].astype(np.float64))
        self.baseline = torch.zeros(P, dtype=torch.int32, device=self.device)
        self.exit_mask = torch.zeros((P, H, self.words_per_row), dtype=torch.int32, device=self.device)
        # levels in the resident (compute) layout + the OR of all cell bits seen (see the header)
        self.resident_bytes = int(_lib.load().slb_resident_bytes(H, W))
        self.board_res = torch.zeros(P * self.resident_bytes, dtype=torch.uint8, device=self.device)
        self.goals_res = torch.zeros_like(self.board_res)
        self.plane_bits = torch.zeros(2, dtype=torch.int32, device=self.device)
        self.level_consts = torch.zeros((P, 32), dtype=torch.uint8, device=self.device)  # slb_level_consts
        self.c = _lib.Pool(P, H, W, self.words_per_row, self.board.data_ptr(), self.goals.data_ptr(),
                           self.agent_x.data_ptr(), self.agent_y.data_ptr(),
                           self.orientation.data_ptr(), self.spawn_prob.data_ptr(),
                           self.min_performance.data_ptr(), self.baseline.data_ptr(),
                           self.exit_mask.data_ptr(), self.board_res.data_ptr(),
                           self.goals_res.data_ptr(), self.plane_bits.data_ptr(),
                           (C.c_uint32 * 2)(0, 0), self.level_consts.data_ptr(), 0, 0)
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().slb_pool_prepare(self.c, _lib.current_stream()))

    def update(self, slots, levels):
        """Replace the levels in `slots` (indices into the pool) with new ones of the same board shape —
        the feed from a host-side level generator (the reference's procedural generation, proc_gen.py via
        speedups.gen_pattern, stays on the host): envs pick the new levels up at their next reset /
        auto-reset into those slots.  An env that is IN an episode on one of the slots keeps its board but
        would see the new level's constants (baseline, exit cells, min_performance): update slots that no
        env is playing, or reset those envs right after.  Re-prepares the pool (slb_pool_prepare)."""
        rec = _stack_levels(levels)
        slots = np.asarray(slots, dtype=np.int64).reshape(-1)
        if rec["board"].shape != (slots.size, self.h, self.w):
            raise ValueError("need one %dx%d level per slot" % (self.h, self.w))
        if slots.size and (slots.min() < 0 or slots.max() >= self.n_levels):
            raise ValueError("slot out of range")
        idx = torch.from_numpy(slots).to(self.device)
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(self.device)  # noqa: E731
        if not getattr(self, "_host_owned", False):  # never write through to the caller's arrays
            self.host = {k: np.array(v) for k, v in self.host.items()}
            self._host_owned = True
        for k in rec:
            self.host[k][slots] = rec[k]
        self.board.view(torch.int16)[idx] = t(rec["board"]).view(torch.int16)
        self.goals.view(torch.int16)[idx] = t(rec["goals"]).view(torch.int16)
        self.agent_x[idx] = t(rec["agent_loc"][:, 0].astype(np.int32))
        self.agent_y[idx] = t(rec["agent_loc"][:, 1].astype(np.int32))
        self.orientation[idx] = t(rec["orientation"].astype(np.int32))
        self.spawn_prob[idx] = t(rec["spawn_prob"].astype(np.float32))
        self.min_performance[idx] = t(rec["min_performance"].astype(np.float64))
        if self.names is not None:
            new_names = _level_names(levels)
            for j, sl in enumerate(slots):
                self.names[sl] = new_names[j] if new_names else "level-%d" % sl
        keep = int(self.c.external_writes)  # out-of-band board writes stay known across the re-prepare
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().slb_pool_prepare(self.c, _lib.current_stream()))
        self.c.external_writes = keep


def _stack_levels(levels):
    keys = ("board", "goals", "agent_loc", "orientation", "spawn_prob", "min_performance")
    if isinstance(levels, dict):
        d = {k: np.asarray(levels[k]) for k in keys}
    elif isinstance(levels, np.ndarray) and levels.dtype.fields:
        d = {k: levels[k] for k in keys}
    else:
        recs = list(levels)
        d = {}
        for k in keys:
            default = {"orientation": 1, "spawn_prob": 0.3, "min_performance": -1, "agent_loc": (0, 0)}
            vals = []
            for r in recs:
                has = (k in r.dtype.fields) if hasattr(r, "dtype") and r.dtype.fields else (k in r)
                vals.append(np.asarray(r[k]) if has else np.asarray(default[k]))
            d[k] = np.stack(vals)
    out = {
        "board": np.ascontiguousarray(d["board"], dtype=np.uint16),
        "goals": np.ascontiguousarray(d["goals"], dtype=np.uint16),
        "agent_loc": np.asarray(d["agent_loc"], dtype=np.int64).reshape(-1, 2),
        "orientation": np.asarray(d["orientation"], dtype=np.int64).reshape(-1),
        "spawn_prob": np.asarray(d["spawn_prob"], dtype=np.float64).reshape(-1),
        "min_performance": np.asarray(d["min_performance"], dtype=np.float64).reshape(-1),
    }
    if out["board"].ndim != 3 or out["board"].shape != out["goals"].shape:
        raise ValueError("levels must stack to board/goals of shape [P,H,W]")
    return out


def _level_names(levels):
    """Level titles when the records carry them (the `name` field of reference archives)."""
    try:
        if isinstance(levels, np.ndarray) and levels.dtype.fields and "name" in levels.dtype.fields:
            return [n.decode() if isinstance(n, bytes) else str(n) for n in levels["name"]]
        if isinstance(levels, dict) and "name" in levels:
            return [str(n) for n in levels["name"]]
    except Exception:
        pass
    return None


def load_levels(path):
    """Read level records from a reference-format .npz: either an archive with a structured
    `levels` array (benchmarks/v1.0/*.npz) or a single saved game (GameState.save)."""
    with np.load(path) as data:
        if "levels" in data:
            return data["levels"].copy()
        return [{k: data[k] for k in data.keys()}]


class SafeLifeGame(object):
    """N lockstep SafeLife games on the device.  Attribute and method names follow the reference's
    SafeLifeGame; values are tensors with a leading env dimension.

    board, goals : uint16 [N,H,W] cuda tensors — materialised on access from the RESIDENT state
                   (board_res / goals_res, the bit-sliced layout the kernels compute in; see
                   include/safelife_b200.h); assigning to them imports the cells
    state        : int32 [N,20] cuda tensor (80-byte slb_env_state records)
    """

    points_on_level_exit = +1  # safelife_game.py:163

    def __init__(self, pool, num_envs, seed=0, env_id0=0, level_ids=None):
        self.pool = pool
        self.n = int(num_envs)
        self.seed, self.env_id0 = int(seed), int(env_id0)
        dev = pool.device
        self.board_res = torch.zeros(self.n * pool.resident_bytes, dtype=torch.uint8, device=dev)
        self.goals_res = torch.zeros_like(self.board_res)
        self.state = torch.zeros((self.n, STATE_WORDS), dtype=torch.int32, device=dev)
        self._lib = _lib.load()
        if level_ids is None:
            level_ids = np.arange(self.n) % pool.n_levels
        self.revert(level_ids)

    # -- plumbing ---------------------------------------------------------------------------
    def _s(self):
        return _lib.current_stream()

    def _field(self, name):
        if name == "episode_reward" or name == "rng_counter":
            return self.state.view(torch.int64)[:, _COL[name] // 2]
        if name == "min_performance":
            return self.state.view(torch.float64)[:, _COL[name] // 2]
        return self.state[:, _COL[name]]

    def state_numpy(self):
        return self.state.cpu().numpy().view(ENV_STATE_DTYPE).reshape(self.n)

    def _export(self, res):
        out = torch.empty((self.n, self.pool.h, self.pool.w), dtype=torch.uint16, device=self.pool.device)
        with torch.cuda.device(self.pool.device):
            _lib.check(self._lib.slb_resident_export(res.data_ptr(), out.data_ptr(), self.n,
                                                     self.pool.h, self.pool.w, self._s()))
        return out

    def cells_of(self, i, which="board"):
        """uint16 [H,W] cells of env i alone (board or goals) without exporting the whole batch."""
        res = self.board_res if which == "board" else self.goals_res
        rb = self.pool.resident_bytes
        out = torch.empty((1, self.pool.h, self.pool.w), dtype=torch.uint16, device=self.pool.device)
        with torch.cuda.device(self.pool.device):
            _lib.check(self._lib.slb_resident_export(res.data_ptr() + int(i) * rb, out.data_ptr(), 1,
                                                     self.pool.h, self.pool.w, self._s()))
        return out[0]

    def _import(self, cells, res, which):
        c = torch.as_tensor(cells)
        if c.dtype == torch.int16:
            c = c.view(torch.uint16)
        c = c.to(self.pool.device).contiguous()
        if c.dtype != torch.uint16 or tuple(c.shape) != (self.n, self.pool.h, self.pool.w):
            raise ValueError("need a uint16 [N,H,W] array")
        with torch.cuda.device(self.pool.device):
            _lib.check(self._lib.slb_resident_import(c.data_ptr(), res.data_ptr(), self.n, self.pool.h,
                                                     self.pool.w, self.pool.c, which, self._s()))

    @property
    def board(self):
        return self._export(self.board_res)

    @board.setter
    def board(self, cells):
        self._import(cells, self.board_res, 0)

    @property
    def goals(self):
        return self._export(self.goals_res)

    @goals.setter
    def goals(self, cells):
        self._import(cells, self.goals_res, 1)

    # -- reference attributes ---------------------------------------------------------------
    @property
    def width(self):
        return self.pool.w

    @property
    def height(self):
        return self.pool.h

    @property
    def agent_loc(self):
        return torch.stack([self._field("agent_x"), self._field("agent_y")], dim=1)

    @property
    def orientation(self):
        return self._field("orientation")

    @property
    def game_over(self):
        return self._field("game_over") != 0

    @property
    def num_steps(self):
        return self._field("num_steps")

    @property
    def level_id(self):
        return self._field("level_id")

    @property
    def rng_counter(self):
        return self._field("rng_counter")

    @property
    def spawn_prob(self):
        return self.pool.spawn_prob[self.level_id.long()]

    @property
    def min_performance(self):
        """GameState.min_performance (safelife_game.py:166) per env: the level's value after a reset
        (deserialize, :205) until something assigns to it."""
        return self._field("min_performance")

    @min_performance.setter
    def min_performance(self, value):
        self.set_min_performance(value)

    def set_min_performance(self, value, mask=None):
        """``game.min_performance = value`` for every env (or the envs in `mask`): can_exit() sees it from the
        next execute_action on and the exits are repainted by the next step, like the reference's
        update_exit_colors (slb_env_set_min_performance)."""
        m = None
        if mask is not None:
            m = torch.as_tensor(mask).to(self.pool.device).to(torch.uint8).contiguous()
        with torch.cuda.device(self.pool.device):
            _lib.check(self._lib.slb_env_set_min_performance(self.n, self.state.data_ptr(), _lib.ptr(m),
                                                             float(value), self._s()))

    # -- reference methods ------------------------------------------------------------------
    def revert(self, level_ids=None, mask=None):
        """GameState.revert / deserialize (safelife_game.py:196-212, 229-234) followed by what
        SafeLifeEnv.reset does to the game (update_exit_colors, safelife_env.py:190).  With
        `level_ids` None every env restarts its current level."""
        if level_ids is None:
            lv = self._field("level_id").contiguous()
        else:
            lv = torch.as_tensor(np.asarray(level_ids), dtype=torch.int32).to(self.pool.device)
        if lv.numel() != self.n:
            raise ValueError("level_ids must have one entry per env")
        m = None
        if mask is not None:
            m = torch.as_tensor(mask).to(self.pool.device).to(torch.uint8).contiguous()
        with torch.cuda.device(self.pool.device):
            _lib.check(self._lib.slb_env_reset(self.pool.c, self.n, lv.data_ptr(), _lib.ptr(m),
                                               self.board_res.data_ptr(), self.goals_res.data_ptr(),
                                               self.state.data_ptr(), self._s()))
        return True

    def execute_action(self, action):
        """Execute one action per env; `action` is a name (same for all), an int, or an
        int array/tensor [N] of env action indices 0..8.  Returns the int32 [N] reward tensor."""
        acts = self._actions(action)
        reward = torch.empty(self.n, dtype=torch.int32, device=self.pool.device)
        with torch.cuda.device(self.pool.device):
            _lib.check(self._lib.slb_execute_action(self.pool.c, self.n, self.board_res.data_ptr(),
                                                    self.state.data_ptr(), acts.data_ptr(),
                                                    reward.data_ptr(), self._s()))
        return reward

    def _actions(self, action):
        if isinstance(action, str):
            if action not in ACTION_INDEX:
                raise ValueError("only the nine env actions are supported on the device: %r" % (action,))
            action = ACTION_INDEX[action]
        if isinstance(action, (int, np.integer)):
            return torch.full((self.n,), int(action), dtype=torch.uint8, device=self.pool.device)
        a = torch.as_tensor(action)
        if a.numel() != self.n:
            raise ValueError("need one action per env")
        return a.to(self.pool.device).to(torch.uint8).contiguous()

    def advance_board(self):
        """SafeLifeGame.advance_board (safelife_game.py:657-660): num_steps += 1, board then goals
        (that is the order the per-env stream is consumed in)."""
        with torch.cuda.device(self.pool.device):
            _lib.check(self._lib.slb_game_advance(self.pool.c, self.n, self.board_res.data_ptr(),
                                                  self.goals_res.data_ptr(), self.state.data_ptr(),
                                                  self.seed, self.env_id0, self._s()))

    def _points_perf(self):
        dev = self.pool.device
        out = [torch.empty(self.n, dtype=torch.int32, device=dev) for _ in range(3)]
        ce = torch.empty(self.n, dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _lib.check(self._lib.slb_points_and_performance(
                self.pool.c, self.n, self.board_res.data_ptr(), self.goals_res.data_ptr(),
                self.state.data_ptr(), out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(),
                ce.data_ptr(), self._s()))
        return out[0], out[1], out[2], ce

    def current_points(self):
        return self._points_perf()[0]

    def performance_ratio(self):
        """-> (completed [N], possible [N]) int32 (safelife_game.py:601-631, unit rewards)."""
        _, c, p, _ = self._points_perf()
        return c, p

    def can_exit(self):
        return self._points_perf()[3] != 0

    def update_exit_colors(self):
        with torch.cuda.device(self.pool.device):
            _lib.check(self._lib.slb_update_exit_colors(self.pool.c, self.n, self.board_res.data_ptr(),
                                                        self.goals_res.data_ptr(), self.state.data_ptr(),
                                                        self._s()))

    # -- checkpoint / resume (GameState.serialize / save / load, safelife_game.py:184-260) -----
    def serialize(self, i):
        """GameWithGoals.serialize (safelife_game.py:184-194, 567-574) for env i: a dict in the reference's
        saved-game format (np.savez_compressed(**d) gives a file its GameState.load reads)."""
        st = self.state[int(i)].cpu().numpy().view(ENV_STATE_DTYPE)[0]
        return {
            "spawn_prob": float(self.pool.host["spawn_prob"][int(st["level_id"])]),
            "orientation": int(st["orientation"]),
            "agent_loc": (int(st["agent_x"]), int(st["agent_y"])),
            "board": self.cells_of(i, "board").cpu().numpy(),
            "goals": self.cells_of(i, "goals").cpu().numpy(),
            "class": "safelife.safelife_game.SafeLifeGame",
            "min_performance": float(st["min_performance"]),
        }

    def save(self, file_name, i=0):
        """GameState.save (safelife_game.py:214-227) for env i."""
        import os
        file_name = os.path.abspath(os.path.expanduser(file_name))
        if not file_name.endswith(".npz"):
            file_name += ".npz"
        np.savez_compressed(file_name, **self.serialize(i))
        return file_name

    def state_dict(self):
        """Everything a batch needs to resume bit-exactly: the resident boards / goals as they are, the env
        state records (agent, counters, cached score, spawn-stream counters, min_performance), seed and global
        env index.  CPU tensors; the level pool is not part of it (load into a game on an equal pool)."""
        return {
            "format": "safelife_k2_b200.SafeLifeGame/1",
            "num_envs": self.n, "shape": (self.pool.h, self.pool.w), "n_levels": self.pool.n_levels,
            "plane_groups": tuple(int(x) for x in self.pool.c.plane_groups),
            "seed": self.seed, "env_id0": self.env_id0,
            "board_res": self.board_res.cpu(), "goals_res": self.goals_res.cpu(), "state": self.state.cpu(),
        }

    def load_state_dict(self, d):
        if d.get("format") != "safelife_k2_b200.SafeLifeGame/1":
            raise ValueError("not a SafeLifeGame state_dict")
        if (d["num_envs"] != self.n or tuple(d["shape"]) != (self.pool.h, self.pool.w)
                or d["n_levels"] != self.pool.n_levels):
            raise ValueError("state_dict of a different batch: %d envs of %s on %d levels" % (
                d["num_envs"], tuple(d["shape"]), d["n_levels"]))
        if tuple(d["plane_groups"]) != tuple(int(x) for x in self.pool.c.plane_groups):
            raise ValueError("state_dict was taken on a level pool with other cell types (plane groups differ)")
        self.seed, self.env_id0 = int(d["seed"]), int(d["env_id0"])
        self.board_res.copy_(d["board_res"])
        self.goals_res.copy_(d["goals_res"])
        self.state.copy_(d["state"])

    @property
    def is_stochastic(self):
        return ((self.board.view(torch.int16) & int(CellTypes.spawning)) != 0).flatten(1).any(1)
