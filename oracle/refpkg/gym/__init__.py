"""Minimal stand-in for the `gym` package (absent from this image) so that the reference's
safelife_env.py / env_wrappers.py import unmodified.  Base classes and a seed pass-through only —
no arithmetic lives here.  Oracle use only."""
from . import spaces  # noqa: F401
from . import utils  # noqa: F401


class Env(object):
    metadata = {}
    action_space = None
    observation_space = None

    def step(self, action):
        raise NotImplementedError

    def reset(self):
        raise NotImplementedError

    def seed(self, seed=None):
        return [seed]

    def close(self):
        pass


class Wrapper(Env):
    def __init__(self, env):
        self.env = env
        self.action_space = getattr(env, "action_space", None)
        self.observation_space = getattr(env, "observation_space", None)

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self.env, name)

    def step(self, action):
        return self.env.step(action)

    def reset(self, **kwargs):
        return self.env.reset(**kwargs)

    def seed(self, seed=None):
        return self.env.seed(seed)


_registry = {}


def register(id, entry_point=None, kwargs=None, **kw):
    _registry[id] = (entry_point, kwargs or {})
