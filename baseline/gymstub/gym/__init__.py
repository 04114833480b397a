"""Minimal stand-in for the slice of ``gym`` 0.25 that the reference's torch env imports
(/root/reference/torch_impl/env/env.py:3-5, wrappers.py:2-8): ``gym.Env``, ``gym.ObservationWrapper``
(``new_step_api=True``) and ``gym.spaces.{Discrete, Box}``.  gym is not installed in this image and cannot be
(no network); the stub lets ``bench.py --impl reference`` time the reference's UNMODIFIED env + WindowedGridView
(copied to ``baseline/_ref`` by ``__graft_entry__.build()``).  It contains no env logic.
"""
from . import spaces  # noqa: F401


class Env:
    metadata = {}
    action_space = None
    observation_space = None

    def reset(self, **kwargs):
        raise NotImplementedError

    def step(self, action):
        raise NotImplementedError


class Wrapper(Env):
    def __init__(self, env, new_step_api=False):
        self.env = env
        self.new_step_api = new_step_api

    def __getattr__(self, name):          # gym forwards unknown attributes to the wrapped env
        if name.startswith('_'):
            raise AttributeError(name)
        return getattr(self.env, name)


class ObservationWrapper(Wrapper):
    """gym 0.25 semantics: reset -> observation(obs) (the env's info is dropped unless return_info=True);
    step -> (observation(obs), reward, terminated, truncated, info) with the new step API."""

    def reset(self, **kwargs):
        if kwargs.get('return_info', False):
            obs, info = self.env.reset(**kwargs)
            return self.observation(obs), info
        return self.observation(self.env.reset(**kwargs))

    def step(self, action):
        obs, reward, terminated, truncated, info = self.env.step(action)
        return self.observation(obs), reward, terminated, truncated, info

    def observation(self, observation):
        raise NotImplementedError
