"""
Batched SafeLifeEnv — host-side mirror of the reference's ``safelife/safelife_env.py`` (old-gym
API: reset() -> obs, step(a) -> (obs, reward, done, info), seed(s) -> [s]) for N lockstep
environments on one B200.  One fused CUDA launch per step (slb_env_step) does
execute_action -> advance_board (board, goals) -> current_points -> reward -> update_exit_colors
-> done for every env (safelife_env.py:157-186); observations come from slb_get_obs
(safelife_env.py:125-155).

Keyword parameters keep the reference names: time_limit, remove_white_goals, view_shape,
output_channels, global_counter (safelife_env.py:73-95).
"""
import os

import numpy as np
import torch

from . import _lib
from .safelife_game import ACTION_NAMES, LevelPool, SafeLifeGame


# the current stream's handle without building a torch.cuda.Stream object (0.7 us per step)
_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
if _raw_stream is None:  # pragma: no cover
    def _raw_stream(index):
        return torch.cuda.current_stream(index).cuda_stream


def _hptr(x):
    return x if isinstance(x, int) or x is None else _lib.ptr(x)


class HostStepper(object):
    """See SafeLifeEnv.host_stepper.  Keeps its env alive (the env owns the native stepper)."""
    __slots__ = ("_env", "_sp", "_step", "_wait", "_dev")

    def __init__(self, env, sp):
        self._env, self._dev = env, env._dev_index
        try:  # the CPython binding of the two calls (csrc/slb_pyfast.c): no ctypes marshalling per step
            if os.environ.get("SLB_NO_PYFAST"):
                raise ImportError
            from . import _slbfast
            self._sp, self._step, self._wait = sp.value, _slbfast.step, _slbfast.wait
        except ImportError:  # not built: the same two C-ABI calls through ctypes
            self._sp, self._step, self._wait = sp, env._lib.slb_stepper_step, env._lib.slb_stepper_wait

    def step(self, actions_ptr, reward_ptr, done_ptr):
        rc = self._step(self._sp, actions_ptr, reward_ptr, done_ptr)
        if rc:
            _lib.check(rc)

    def wait(self):
        rc = self._wait(self._sp)
        if rc:
            _lib.check(rc)

    def bind_outputs(self, reward_ptr, done_ptr):
        """The outputs of later step(a, reward_ptr, done_ptr) calls are one page-locked block [int32 reward x n |
        uint8 done x n] (e.g. a torch pin_memory() uint8 tensor of 5 n bytes, n a multiple of 16) that stays alive:
        the step then writes them there directly (slb_stepper_bind_outputs).  None, None unbinds."""
        _lib.check(self._env._lib.slb_stepper_bind_outputs(self._sp, reward_ptr, done_ptr))


class _Info(dict):
    """The step's info dict (safelife_env.py:176-186): 'board', 'goals' (uint16 [N,H,W] tensors exported
    from the resident state), 'agent_loc', 'times_up', 'episode'.  Every value is read from the env state
    when it is first asked for, not on every step: a policy loop that never looks at them pays nothing
    for them.  Read them before the next step() — the state they come from moves on."""
    _KEYS = ("board", "goals", "agent_loc", "times_up", "episode")

    def __init__(self, game, time_limit):
        super().__init__()
        self._game, self._time_limit = game, time_limit

    def __missing__(self, key):
        g = self._game
        if key in ("board", "goals"):
            val = getattr(g, key)
        elif key == "agent_loc":
            val = g.agent_loc
        elif key == "times_up":
            val = g._field("episode_length") > self._time_limit
        elif key == "episode":
            val = {"length": g._field("episode_length"), "reward": g._field("episode_reward")}
        else:
            raise KeyError(key)
        self[key] = val
        return val

    def get(self, key, default=None):
        return self[key] if key in self._KEYS else super().get(key, default)

    def keys(self):
        return list(super().keys() | set(self._KEYS))

    def __iter__(self):
        return iter(self.keys())

    def __len__(self):
        return len(self.keys())

    def items(self):
        return [(k, self[k]) for k in self.keys()]

    def values(self):
        return [self[k] for k in self.keys()]

    def __contains__(self, key):
        return key in self._KEYS or super().__contains__(key)


class GlobalCounter(object):
    """SafeLifeEnv.global_counter (safelife_env.py:81-85: episodes_started, episodes_completed, num_steps)
    without a host synchronisation per step: every env batch that uses the counter owns a small device
    buffer that its step kernel adds to (slb_dev_stepper_create's d_global_counter), and the attributes are
    summed up when they are READ (one device->host copy per batch then).  Assigning to an attribute works
    (`counter.num_steps = 0`).  Like the reference's class-level SimpleNamespace, one GlobalCounter is the
    class-level default and is shared by every SafeLifeEnv that is not given its own."""

    def __init__(self, episodes_started=0, episodes_completed=0, num_steps=0):
        object.__setattr__(self, "_base", {"episodes_started": int(episodes_started),
                                            "episodes_completed": int(episodes_completed),
                                            "num_steps": int(num_steps)})
        object.__setattr__(self, "_sources", [])  # weak references to the envs' device buffers

    def _register(self, buf):
        import weakref
        self._sources.append(weakref.ref(buf))

    def _device_sums(self):
        steps = done = 0
        alive = []
        for r in self._sources:
            buf = r()
            if buf is None:
                continue
            alive.append(r)
            v = buf.cpu().view(-1, 2).sum(0)
            steps += int(v[0])
            done += int(v[1])
        self._sources[:] = alive
        return steps, done

    def _retire(self, buf):
        """fold a device buffer that is going away into the host totals"""
        v = buf.cpu().view(-1, 2).sum(0)
        self._base["num_steps"] += int(v[0])
        self._base["episodes_completed"] += int(v[1])
        self._sources[:] = [r for r in self._sources if r() is not None and r() is not buf]

    def __getattr__(self, name):
        base = object.__getattribute__(self, "_base")
        if name not in base:
            raise AttributeError(name)
        if name == "episodes_started":
            return base[name]
        steps, done = self._device_sums()
        return base[name] + (steps if name == "num_steps" else done)

    def __setattr__(self, name, value):
        if name not in self._base:
            raise AttributeError(name)
        if name == "episodes_started":
            self._base[name] = int(value)
            return
        steps, done = self._device_sums()
        self._base[name] = int(value) - (steps if name == "num_steps" else done)

    def __repr__(self):
        return "GlobalCounter(episodes_started=%d, episodes_completed=%d, num_steps=%d)" % (
            self.episodes_started, self.episodes_completed, self.num_steps)


class SafeLifeEnv(object):
    metadata = {"render.modes": []}
    action_names = ACTION_NAMES
    game = None

    time_limit = 1000
    remove_white_goals = True
    view_shape = (15, 15)
    output_channels = tuple(range(15))
    global_counter = GlobalCounter()  # any other object with the three attributes is updated eagerly (two syncs per step)

    # batched-only knobs
    num_envs = 1
    env_id0 = 0          # global index of this shard's first env (placement-independent streams)
    auto_reset = False   # reset finished envs inside the step kernel (ContinuingEnv-like)
    compute_obs = True   # False: step()/reset() return obs=None (policy reads game.board itself)
    level_ids = None     # first level of each env (default: (global env index) % pool size)
    shard = None         # (lo, hi, total) when made by SafeLifeEnv.sharded

    def __init__(self, level_iterator, **kwargs):
        """level_iterator: a LevelPool, or anything LevelPool accepts (structured level array,
        list of level records).  Env i plays levels i, i+N, i+2N, ... (mod pool size)."""
        for key, val in kwargs.items():
            if not key.startswith("_") and hasattr(self, key) and not callable(getattr(self, key)):
                setattr(self, key, val)
            else:
                raise ValueError("Unrecognized parameter: '%s'" % (key,))
        self.pool = level_iterator if isinstance(level_iterator, LevelPool) else LevelPool(level_iterator)
        self.num_envs = int(self.num_envs)
        self._seed = 0
        self._lib = _lib.load()
        self._steppers = {}
        self._dev_sp, self._dev_key, self._dev_cur, self._obs_shape = None, None, None, None
        self._obs_scratch = False
        self._fused_obs = True
        self._dev_index = self.pool.device.index if self.pool.device.index is not None else torch.cuda.current_device()
        dev = self.pool.device
        self._reward = torch.zeros(self.num_envs, dtype=torch.int32, device=dev)
        self._gc_dev = None
        if isinstance(self.global_counter, GlobalCounter):
            self._gc_dev = torch.zeros(128, dtype=torch.int64, device=dev)  # 64 x {num_steps, episodes_completed}
            self.global_counter._register(self._gc_dev)
        self._done = torch.zeros(self.num_envs, dtype=torch.uint8, device=dev)
        if self.level_ids is None:
            self._next_level = (np.arange(self.num_envs) + self.env_id0) % self.pool.n_levels
        else:
            self._next_level = np.asarray(self.level_ids, dtype=np.int64).reshape(self.num_envs) % self.pool.n_levels
        self.seed()

    @classmethod
    def sharded(cls, level_iterator, total_envs, rank=None, world_size=None, device=None, **kwargs):
        """This rank's contiguous share of a job of `total_envs` envs (envs shard by global index,
        one process per GPU, no data-path collective).  rank / world_size default to the torchrun environment
        (distributed.init_from_env, which also creates the process group for the statistics all-reduce); the
        level pool is put on this rank's GPU.  An env's levels and spawn stream depend on its GLOBAL index only,
        so the job's results do not depend on the number of GPUs."""
        from . import distributed
        if rank is None or world_size is None:
            rank, world_size, local = distributed.init_from_env()
            if device is None:
                device = torch.device("cuda", local)
        lo, hi = distributed.shard_range(int(total_envs), int(rank), int(world_size))
        pool = level_iterator if isinstance(level_iterator, LevelPool) else LevelPool(level_iterator, device=device)
        base = int(kwargs.pop("env_id0", 0))
        env = cls(pool, num_envs=hi - lo, env_id0=base + lo, **kwargs)
        env.shard = (lo, hi, int(total_envs))
        return env

    def state_dict(self):
        """Checkpoint of the env batch: the game's state_dict plus the level cursor, the seed and the
        global counter's values."""
        gc = self.global_counter
        return {"game": self.game.state_dict() if self.game is not None else None,
                "next_level": np.array(self._next_level), "seed": self._seed,
                "global_counter": None if gc is None else {k: int(getattr(gc, k)) for k in
                                                           ("episodes_started", "episodes_completed", "num_steps")}}

    def load_state_dict(self, d):
        if d["game"] is not None:
            if self.game is None:
                self.reset()
            self.game.load_state_dict(d["game"])
        self._next_level = np.array(d["next_level"])
        self._seed = int(d["seed"])
        if d.get("global_counter") is not None and self.global_counter is not None:
            for k, v in d["global_counter"].items():
                setattr(self.global_counter, k, v)
        self._dev_key = self._dev_cur = None  # rebind the native steppers (the seed may have changed)
        self._drop_steppers()

    # ------------------------------------------------------------------------------------------
    def seed(self, seed=None):
        """safelife_env.py:120-123: sets the seed of the spawn streams (the reference reseeds the
        global numpy stream through speedups.seed).  Draw counters restart at the next reset()."""
        if seed is None:
            seed = int(np.random.SeedSequence().generate_state(1)[0] & 0x7FFFFFFF)
        self._seed = int(seed)
        if self.game is not None:
            self.game.seed = self._seed
            self.game.rng_counter.zero_()
        return [seed]

    def _stride(self):
        return self.num_envs % self.pool.n_levels if self.num_envs % self.pool.n_levels else self.pool.n_levels

    def reset(self, mask=None):
        """safelife_env.py:188-198 for every env (or the envs selected by `mask`)."""
        if self.game is None:
            self.game = SafeLifeGame(self.pool, self.num_envs, seed=self._seed, env_id0=self.env_id0,
                                     level_ids=self._next_level)
            started = self.num_envs
            self._next_level = (self._next_level + self.num_envs) % self.pool.n_levels
        else:
            m = np.ones(self.num_envs, bool) if mask is None else np.asarray(
                mask.cpu() if hasattr(mask, "cpu") else mask).astype(bool)
            self.game.revert(self._next_level, mask=m)
            self._next_level = np.where(m, (self._next_level + self.num_envs) % self.pool.n_levels,
                                        self._next_level)
            started = int(m.sum())
        if self.global_counter is not None:
            self.global_counter.episodes_started += started
        return self.get_obs() if self.compute_obs else None

    def get_obs(self, out=None):
        """uint16 [N,vh,vw,C] (output_channels) or [N,vh,vw] (output_channels None).  `out`: an
        existing cuda tensor of that shape to fill instead of allocating a new one every call."""
        g = self.game
        vh, vw = self.view_shape
        dev = self.pool.device
        if torch.cuda.current_device() != self._dev_index:
            torch.cuda.set_device(self._dev_index)
        st = _lib.current_stream()
        if not self.output_channels:
            words = out if out is not None else torch.empty((g.n, vh, vw), dtype=torch.uint16, device=dev)
            _lib.check(self._lib.slb_get_obs(self.pool.c, g.n, g.board_res.data_ptr(), g.goals_res.data_ptr(),
                                             g.state.data_ptr(), vh, vw, int(self.remove_white_goals),
                                             words.data_ptr(), st))
            return words
        ch = self._channels()
        if out is None:
            out = torch.empty((g.n, vh, vw, ch.size), dtype=torch.uint16, device=dev)
        scratch = torch.empty((g.n, vh, vw), dtype=torch.uint16, device=dev) if self._obs_scratch else None
        rc = self._lib.slb_get_obs_channels(self.pool.c, g.n, g.board_res.data_ptr(), g.goals_res.data_ptr(),
                                            g.state.data_ptr(), vh, vw, int(self.remove_white_goals),
                                            ch.ctypes.data, ch.size, _lib.ptr(scratch), out.data_ptr(), st)
        if rc == _lib.SLB_ERR_UNSUPPORTED and scratch is None:
            self._obs_scratch = True  # this board / view shape takes the two-kernel path
            return self.get_obs(out)
        _lib.check(rc)
        return out

    def _channels(self):
        key = tuple(self.output_channels)
        if getattr(self, "_ch_key", None) != key:
            if len(key) > 16 or any(c < 0 or c > 15 for c in key):
                raise ValueError("output_channels must be at most 16 bit indices in 0..15")
            self._ch = np.array(key, dtype=np.uint8)  # host side: slb_obs_channels takes the list by value
            self._ch_key = key
        return self._ch

    def step(self, action):
        """action: int array/tensor [N] (0..8).  Returns (obs, reward int32 [N], done bool [N], info); the
        reference's reward is a numpy.int64 scalar, here an int32 tensor (|reward| <= 5 * H * W).  reward and
        done are fresh tensors every step (a rollout buffer may keep them).
        The call path is kept short for a policy on the GPU: a contiguous uint8 cuda tensor of actions is
        used as it is, `done` is a bool view of the step's uint8 output, `info` is filled on access, and the
        default global_counter (a GlobalCounter) is counted by the step kernel itself."""
        g = self.game
        assert g is not None, "Game state is not initialized."
        if (type(action) is torch.Tensor and action.dtype == torch.uint8 and action.is_cuda
                and action.numel() == g.n and action.is_contiguous()):
            acts = action
        else:
            acts = g._actions(action)
        eager = self.global_counter is not None and self._gc_dev is None
        already = g._field("episode_completed").clone() if eager else None
        if torch.cuda.current_device() != self._dev_index:
            torch.cuda.set_device(self._dev_index)
        st = _raw_stream(self._dev_index)
        # every argument that does not change from step to step is bound once in a native stepper
        oc = self.output_channels
        # (compared by identity / equality against the values the stepper was bound with: no tuple(oc) per step)
        cur = (g, self.time_limit, g.seed, self.auto_reset, self.view_shape, oc, self.remove_white_goals,
               self.compute_obs, self._fused_obs, self.global_counter is not None)
        if cur != self._dev_cur:
            key = (id(g), self.time_limit, g.seed, self.auto_reset, self.view_shape, tuple(oc) if oc else None,
                   self.remove_white_goals, self.compute_obs, self._fused_obs, self.global_counter is not None)
            if key != self._dev_key:
                self._bind_dev_stepper(key)
            self._dev_cur = cur[:5] + (list(oc) if isinstance(oc, list) else oc,) + cur[6:]
        dev = self.pool.device
        done = torch.empty(g.n, dtype=torch.bool, device=dev)  # the step writes 0 / 1; fresh every step
        reward = torch.empty(g.n, dtype=torch.int32, device=dev)
        self._done, self._reward = done.view(torch.uint8), reward
        obs = torch.empty(self._obs_shape, dtype=torch.uint16, device=dev) if self._obs_shape is not None else None
        rc = self._dev_step(self._dev_sp, acts.data_ptr(), reward.data_ptr(), done.data_ptr(),
                            obs.data_ptr() if obs is not None else None, st)
        if rc == _lib.SLB_ERR_UNSUPPORTED and obs is not None:
            # this board / view shape has no fused step + observation: step, then get_obs
            self._fused_obs = False
            k0 = self._dev_key
            self._bind_dev_stepper(k0[:8] + (False,) + k0[9:])
            self._dev_cur = None
            obs = None
            rc = self._dev_step(self._dev_sp, acts.data_ptr(), reward.data_ptr(), done.data_ptr(), None, st)
        if rc:
            _lib.check(rc)
        if eager:
            fresh = already == 0
            self.global_counter.episodes_completed += int((done & fresh).sum().item())
            self.global_counter.num_steps += int(fresh.sum().item())
        info = _Info(g, self.time_limit)
        if obs is None and self.compute_obs:
            obs = self.get_obs()
        return obs, reward, done, info

    def _bind_dev_stepper(self, key):
        import ctypes
        g = self.game
        if self._dev_sp is not None:
            self._lib.slb_dev_stepper_destroy(self._dev_sp)
            self._dev_sp = None
        fused = self.compute_obs and self._fused_obs
        vh, vw = self.view_shape if fused else (0, 0)
        ch = self._channels() if (fused and self.output_channels) else None
        self._obs_shape = None
        if fused:
            self._obs_shape = (g.n, vh, vw, ch.size) if ch is not None else (g.n, vh, vw)
        out = ctypes.c_void_p()
        _lib.check(self._lib.slb_dev_stepper_create(
            self.pool.c, g.n, g.board_res.data_ptr(), g.goals_res.data_ptr(), g.state.data_ptr(),
            int(self.time_limit), g.seed, g.env_id0, self._stride() if self.auto_reset else 0,
            self._reward.data_ptr(), vh, vw, int(self.remove_white_goals),
            ch.ctypes.data if ch is not None else None, ch.size if ch is not None else 0,
            self._gc_dev.data_ptr() if (self._gc_dev is not None and self.global_counter is not None) else None,
            ctypes.byref(out)))
        self._dev_sp, self._dev_key, self._dev_step = out, key, self._lib.slb_dev_stepper_step
        try:  # the CPython binding of the call (csrc/slb_pyfast.c): no ctypes marshalling per step
            if os.environ.get("SLB_NO_PYFAST"):
                raise ImportError
            from . import _slbfast
            self._dev_sp, self._dev_step = out.value, _slbfast.dev_step
        except (ImportError, AttributeError):
            pass

    def step_host(self, actions_host, reward_host, done_host, sync=True, stream=None):
        """Host-buffer step (slb_env_step_host): uint8 actions in host memory in, int32 rewards / uint8
        dones out; board state stays resident on the device.  Pass raw pointers (ints) or tensors /
        arrays.  sync=False only queues the copies and the step on `stream` (a torch stream; default the
        current one) — call wait(stream) before reading the outputs; independent env batches stepped on
        different streams overlap."""
        st = (stream if stream is not None else torch.cuda.current_stream()).cuda_stream
        g = self.game
        key = (st, id(g), self.time_limit, g.seed, self.auto_reset)
        sp = self._steppers.get(key)
        if sp is None:
            sp = self._make_stepper(key, st)
        if torch.cuda.current_device() != self._dev_index:
            torch.cuda.set_device(self._dev_index)
        rc = self._lib.slb_stepper_step(sp, _hptr(actions_host), _hptr(reward_host), _hptr(done_host))
        if rc == 0 and sync:
            rc = self._lib.slb_stepper_wait(sp)
        if rc:
            _lib.check(rc)

    def host_stepper(self, stream=None):
        """The host-buffer step bound to this env batch and `stream` (default: the current one) as a
        small object for hot actor loops: `hs.step(a_ptr, r_ptr, d_ptr)` queues one step (raw host
        pointers as ints: uint8 actions in, int32 rewards / uint8 dones out) and `hs.wait()` returns once
        that step's outputs are in the caller's buffers — one FFI call each, nothing else.  Make a new
        one after seed() or after changing time_limit / auto_reset."""
        st = (stream if stream is not None else torch.cuda.current_stream()).cuda_stream
        g = self.game
        key = (st, id(g), self.time_limit, g.seed, self.auto_reset)
        sp = self._steppers.get(key)
        if sp is None:
            sp = self._make_stepper(key, st)
        return HostStepper(self, sp)

    def _make_stepper(self, key, st):
        """Bind the per-env-batch arguments once (slb_stepper_create): the per-step FFI call then
        carries only the three host pointers."""
        import ctypes
        g = self.game
        if len(self._steppers) >= 64:
            self._drop_steppers()
        out = ctypes.c_void_p()
        _lib.check(self._lib.slb_stepper_create(
            self.pool.c, g.n, g.board_res.data_ptr(), g.goals_res.data_ptr(), g.state.data_ptr(),
            int(self.time_limit), g.seed, g.env_id0, self._stride() if self.auto_reset else 0, st,
            ctypes.byref(out)))
        self._steppers[key] = out
        return out

    def _drop_steppers(self):
        for sp in self._steppers.values():
            self._lib.slb_stepper_destroy(sp)
        self._steppers.clear()
        if self._dev_sp is not None:
            self._lib.slb_dev_stepper_destroy(self._dev_sp)
            self._dev_sp, self._dev_key, self._dev_cur = None, None, None

    def __del__(self):
        try:
            self._drop_steppers()
        except Exception:
            pass

    def wait(self, stream=None):
        st = (stream if stream is not None else torch.cuda.current_stream()).cuda_stream
        _lib.check(self._lib.slb_stream_wait(st))

    def rollout(self, actions, reward=None, done=None):
        """T fused steps with pre-staged actions: uint8 [T,N] cuda tensor.  Returns (reward int32
        [T,N], done uint8 [T,N]).  Board state stays on chip between steps on fast-path shapes."""
        g = self.game
        T = actions.shape[0]
        dev = self.pool.device
        if reward is None:
            reward = torch.empty((T, g.n), dtype=torch.int32, device=dev)
        if done is None:
            done = torch.empty((T, g.n), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _lib.check(self._lib.slb_env_rollout(
                self.pool.c, g.n, g.board_res.data_ptr(), g.goals_res.data_ptr(), g.state.data_ptr(),
                actions.data_ptr(), int(T), int(self.time_limit), g.seed, g.env_id0,
                reward.data_ptr(), done.data_ptr(), self._stride() if self.auto_reset else 0,
                _lib.current_stream()))
        return reward, done

    def render(self, mode="ansi"):
        raise NotImplementedError("rendering is out of scope (SURVEY.md §2.1)")

    def close(self):
        pass
