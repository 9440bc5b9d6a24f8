"""Minimal stand-in for OpenAI `gym` (not installed here) — TEST INFRASTRUCTURE, just enough for the
reference's gym_air_hockey to import and run: Env, spaces.Discrete/Box, envs.registration.register, make."""
import importlib

from . import spaces  # noqa: F401
from .envs import registration  # noqa: F401


class Env(object):
    pass


def make(env_id):
    module, cls = registration.registry[env_id].split(':')
    return getattr(importlib.import_module(module), cls)()
