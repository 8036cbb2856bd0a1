"""Minimal stand-in for gym 0.26.2 (the API generation the reference targets: ``reset`` returns
``(obs, info)``, ``step`` returns ``(obs, reward, terminated, truncated, info)``): ``Env``, ``Wrapper``,
``spaces.Box/Dict``, the two wrappers the reference uses (reference src/rlsp/agents/simple_ddpg.py:51-98)
and ``utils.seeding``.  Stand-in for the un-installable dependency; no part of the CUDA hot path."""
from . import spaces, utils, wrappers  # noqa: F401
from .core import Env, Wrapper  # noqa: F401
from .envs import registration  # noqa: F401

__version__ = "0.26.2-compat"


def make(*a, **k):
    raise NotImplementedError("gym.make is not part of the compat shim (the reference constructs GymEnv directly)")
