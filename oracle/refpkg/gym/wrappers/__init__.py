"""gym.wrappers stub (oracle use only): only the import path the reference's env_wrappers.py needs."""
from . import monitoring  # noqa: F401
