"""
Training-time wrappers — host-side mirror of the reference's ``safelife/env_wrappers.py`` for the
batched env: ``MovementBonusWrapper`` (:39-95), ``ContinuingEnv`` (:289-303),
``SimpleSideEffectPenalty`` (:306-346).  Same class / attribute names and arithmetic, with a leading
env dimension; rewards become float64 tensors once a wrapper adds a float term, like the reference's
python floats.  The side-effect count runs on the GPU (slb_side_effect_penalty).

Out of scope: RecordingSafeLifeWrapper (video / YAML / TensorBoard logging).
"""
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
        counter = self.global_counter
        num_steps = 0 if counter is None else counter.num_steps
        return val(num_steps) if callable(val) else val

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
        # `self.game.min_performance = ...` (:314): every level of the pool gets the scheduled value
        g.pool.min_performance.fill_(float(self.scheduled(self.min_performance)))
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
