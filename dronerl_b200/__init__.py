"""dronerl_b200 -- B200-native batched DroneRL environment (reset / step / get_obs as sm_100a CUDA
kernels behind a C ABI), presented through the reference's own env API.

Reference: nyx-ai/droneRL, jax_impl/env/env.py (DeliveryDrones, DroneEnvParams, DroneEnvState) and
common/constants.py (Action, Object).
"""
from .params import Action, Object, DroneEnvParams
from .state import DroneEnvState
from .env import DeliveryDrones
from . import random
from . import gym_api

__all__ = ['Action', 'Object', 'DroneEnvParams', 'DroneEnvState', 'DeliveryDrones', 'random', 'gym_api']
