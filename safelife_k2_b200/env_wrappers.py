"""
Training-time wrappers — host-side mirror of the reference's ``safelife/env_wrappers.py`` for the
batched env: ``MovementBonusWrapper`` (:39-95), ``ContinuingEnv`` (:289-303),
``SimpleSideEffectPenalty`` (:306-346).  Same class / attribute names and arithmetic, with a leading
env dimension; rewards become float64 tensors once a wrapper adds a float term, like the reference's
python floats.  The side-effect count runs on the GPU (slb_side_effect_penalty).

``SafeLifeRecorder`` (:97-136) and ``RecordingSafeLifeWrapper`` (:139-286) keep the parts that are
data: the trajectory ``.npz`` (orientation / board / goals per frame) and the YAML episode log.  The
video frames and the TensorBoard writer are left out: they need a renderer / TensorFlow, unavailable on
this box.
"""
import os

import numpy as np
import torch

from . import _lib


class BaseWrapper(object):
    """env_wrappers.py:16-36: keyword attributes, attribute forwarding, ``scheduled``."""

    def __init__(self, env, **kwargs):
        for key, val in kwargs.items():
            if not key.startswith("_") and hasattr(self, key) and not callable(getattr(self, key)):
                setattr(self, key, val)
            else:
                raise ValueError("Unrecognized parameter: '%s'" % (key,))
        self.env = env

    def __getattr__(self, name):
        if name.startswith("_") or name == "env":
            raise AttributeError(name)
        return getattr(self.env, name)

    def scheduled(self, val):
        if not callable(val):
            return val  # (the counter is only read — one device sync with a GlobalCounter — for a schedule)
        counter = self.global_counter
        return val(0 if counter is None else counter.num_steps)

    def step(self, action):
        return self.env.step(action)

    def reset(self, mask=None):
        return self.env.reset(mask)

    def seed(self, seed=None):
        return self.env.seed(seed)


def _mask_tensor(mask, n, device):
    if mask is None:
        return torch.ones(n, dtype=torch.bool, device=device)
    return torch.as_tensor(mask).to(device).to(torch.bool)


class MovementBonusWrapper(BaseWrapper):
    """env_wrappers.py:39-95.  The deque of prior positions becomes a [period, N, 2] history with a
    per-env fill count."""
    movement_bonus = 0.1
    movement_bonus_power = 0.01
    movement_bonus_period = 4

    def step(self, action):
        obs, reward, done, info = self.env.step(action)
        n = self.movement_bonus_period
        p0 = self.game.agent_loc.to(torch.int64)              # [N,2]
        ln = self._len                                        # entries held, 1..n
        # oldest entry that is looked at: prior[-n] when full, else prior[0]
        idx = (n - ln).clamp(min=0)
        p1 = self._hist[idx, torch.arange(p0.shape[0], device=p0.device)]
        dist = (p0 - p1).abs().sum(1) + (n - ln).clamp(min=0)  # treat the start as continuous motion
        speed = dist.to(torch.float64) / n
        reward = reward.to(torch.float64) + self.movement_bonus * speed ** self.movement_bonus_power
        self._push(p0)
        return obs, reward, done, info

    def _push(self, p):
        self._hist = torch.cat([self._hist[1:], p[None]], 0)
        self._len = (self._len + 1).clamp(max=self.movement_bonus_period)

    def reset(self, mask=None):
        obs = self.env.reset(mask)
        n = self.movement_bonus_period
        loc = self.game.agent_loc.to(torch.int64)
        m = _mask_tensor(mask, loc.shape[0], loc.device)
        if getattr(self, "_hist", None) is None or self._hist.shape[1] != loc.shape[0]:
            self._hist = loc[None].repeat(n, 1, 1)
            self._len = torch.ones(loc.shape[0], dtype=torch.int64, device=loc.device)
        else:
            self._hist[-1] = torch.where(m[:, None], loc, self._hist[-1])
            self._len = torch.where(m, torch.ones_like(self._len), self._len)
        return obs

    _hist = None


class ContinuingEnv(BaseWrapper):
    """env_wrappers.py:289-303: an episode only ends when the time is up; an env that finished its
    level otherwise is reset on the spot and keeps going."""

    def step(self, action):
        obs, reward, done, info = self.env.step(action)
        restart = done & ~info["times_up"]
        if bool(restart.any()):
            done = done & ~restart
            obs = self.env.reset(restart)
        return obs, reward, done, info


class SimpleSideEffectPenalty(BaseWrapper):
    """env_wrappers.py:306-346: penalise departures from the starting state."""
    penalty_coef = 0.0
    min_performance = 0.01

    def reset(self, mask=None):
        obs = self.env.reset(mask)
        g = self.game
        m = _mask_tensor(mask, g.n, g.pool.device)
        if getattr(self, "last_side_effect", None) is None or self.last_side_effect.shape[0] != g.n:
            self.last_side_effect = torch.zeros(g.n, dtype=torch.int32, device=g.pool.device)
        else:
            self.last_side_effect = torch.where(m, torch.zeros_like(self.last_side_effect), self.last_side_effect)
        # `self.game.min_performance = ...` (:316) for the envs that were just reset: per-env game state,
        # the envs that are in the middle of an episode keep the value scheduled at their own reset
        g.set_min_performance(float(self.scheduled(self.min_performance)), None if mask is None else m)
        return obs

    def side_effect_count(self):
        g = self.game
        out = torch.empty(g.n, dtype=torch.int32, device=g.pool.device)
        with torch.cuda.device(g.pool.device):
            _lib.check(_lib.load().slb_side_effect_penalty(g.pool.c, g.n, g.board_res.data_ptr(),
                                                           g.goals_res.data_ptr(), g.state.data_ptr(),
                                                           out.data_ptr(), _lib.current_stream()))
        return out

    def step(self, action):
        observation, reward, done, info = self.env.step(action)
        side_effect = self.side_effect_count()
        delta_effect = side_effect - self.last_side_effect
        reward = reward.to(torch.float64) - delta_effect.to(torch.float64) * float(self.scheduled(self.penalty_coef))
        self.last_side_effect = side_effect
        return observation, reward, done, info

    last_side_effect = None


class SafeLifeRecorder(object):
    """env_wrappers.py:97-136 without the video: the trajectory of ONE env of the batch.  `close()`
    writes ``base_path + '.npz'`` with the reference's arrays: orientation [T], board [T,H,W], goals
    [T,H,W] (one frame per capture while the env's game is not over)."""

    def __init__(self, env, env_index=0, enabled=True, base_path=None):
        self.env, self.env_index, self.enabled, self.base_path = env, int(env_index), enabled, base_path
        self.trajectory = {"orientation": [], "board": [], "goals": []}

    def capture_frame(self):
        game = self.env.game
        if not self.enabled or game is None:
            return
        i = self.env_index
        st = game.state[i].cpu().numpy()
        if st[3]:  # game_over
            return
        self.trajectory["orientation"].append(int(st[2]))
        self.trajectory["board"].append(game.cells_of(i, "board").cpu().numpy())
        self.trajectory["goals"].append(game.cells_of(i, "goals").cpu().numpy())

    def close(self):
        if self.enabled and self.base_path is not None:
            np.savez_compressed(self.base_path + ".npz", **self.trajectory)
        self.enabled = False


class RecordingSafeLifeWrapper(BaseWrapper):
    """env_wrappers.py:139-286 for the batched env: end-of-episode YAML records (same text as the
    reference, one per env episode, written to `log_file`) and trajectory recordings of env
    `record_env` every `video_recording_freq` episodes of that env.  Side effects are scored for all
    envs that finished in a step with one launch."""
    log_file = None
    video_name = None
    video_recorder = None
    video_recording_freq = 100
    record_side_effects = True
    other_episode_data = {}
    record_env = 0
    level_names = None
    num_samples = 1000
    emd = True  # False: skip the host-side earth mover distance (its value is written as 0.00)

    def __init__(self, env, **kwargs):
        super().__init__(env, **kwargs)
        self._logged = None
        self.episode_num = None

    def _names(self):
        names = self.level_names or getattr(self.pool, "names", None)
        return names or ["level-%d" % k for k in range(self.pool.n_levels)]

    def log_episodes(self, idx):
        """Records of the envs `idx` (which just finished) -> list of message strings."""
        from .benchmarking import CellTypes, format_episode
        from .side_effects import side_effect_score
        g = self.game
        st = g.state_numpy()
        pool = self.pool
        green_life = int(CellTypes.life | CellTypes.color_g)
        init = pool.host["board"]
        side = None
        if self.record_side_effects:
            side = side_effect_score(g, num_samples=self.num_samples, envs=idx, emd=self.emd)
        names = self._names()
        msgs = []
        for k, i in enumerate(idx):
            lv = int(st["level_id"][i])
            initial_green = int(((init[lv] | int(CellTypes.destructible)) == green_life).sum())
            se = None
            if side is not None:
                se = {c: [0.0 if v[0] is None else v[0], v[1]] for c, v in side[k].items()}
            msg = format_episode(names[lv], int(self.episode_num[i]), int(st["episode_length"][i]),
                                 int(st["episode_reward"][i]), int(st["completed"][i]), int(st["possible"][i]),
                                 max(0, float(pool.host["min_performance"][lv])), initial_green, None)
            for key, val in self.other_episode_data.items():
                msg += "  {}: {:0.4g}\n".format(key, self.scheduled(val))
            if se is not None:
                msg += "  side effects:\n"
                from .benchmarking import cell_name
                msg += "\n".join("    {}: [{:0.2f}, {:0.2f}]".format(cell_name(c), v[0], v[1]) for c, v in se.items())
            msgs.append(msg)
            if self.log_file is not None:
                self.log_file.write(msg)
        if self.log_file is not None and msgs:
            self.log_file.flush()
        return msgs

    def step(self, action):
        obs, reward, done, info = self.env.step(action)
        if self.video_recorder is not None:
            self.video_recorder.capture_frame()
        d = done.cpu().numpy().astype(bool)
        fresh = d & ~self._logged
        if fresh.any():
            self._logged |= fresh
            self.log_episodes(np.nonzero(fresh)[0])
        return obs, reward, done, info

    def reset(self, mask=None):
        n = self.num_envs
        m = np.ones(n, bool) if mask is None else np.asarray(mask.cpu() if hasattr(mask, "cpu") else mask).astype(bool)
        if self._logged is None:
            self._logged = np.zeros(n, bool)
            self.episode_num = np.full(n, -1, np.int64)
        self._logged[m] = False
        counter = self.global_counter
        started = 0 if counter is None else counter.episodes_started
        obs = self.env.reset(mask)
        if counter is not None:
            # the reference numbers an episode by episodes_started after its reset (one env at a time)
            self.episode_num[m] = started + 1 + np.arange(int(m.sum()))
        if m[self.record_env]:
            self.reset_video_recorder()
        return obs

    def close(self):
        if self.video_recorder is not None:
            self.video_recorder.close()
            self.video_recorder = None

    def reset_video_recorder(self):
        self.close()
        num = int(self.episode_num[self.record_env])
        if self.video_name and num % self.video_recording_freq == 0:
            counter = self.global_counter
            lv = int(self.game.state[self.record_env, 5].item())
            name = self.video_name.format(level_title=self._names()[lv], episode_num=num,
                                          step_num=0 if counter is None else counter.num_steps)
            path = p0 = os.path.abspath(name)
            os.makedirs(os.path.split(path)[0], exist_ok=True)
            k = 1
            while os.path.exists(path + ".npz"):
                k += 1
                path = p0 + " ({})".format(k)
            self.video_recorder = SafeLifeRecorder(self.env, self.record_env, base_path=path)
            self.video_recorder.capture_frame()
