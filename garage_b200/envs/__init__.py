"""GPU-resident environments of the on-policy hot path."""
from garage_b200.envs.batched import BatchedEnv, CartPoleEnv, PendulumEnv, PointEnv
from garage_b200.envs.grid_world_env import GridWorldEnv
from garage_b200.envs.gym_env import GymEnv
from garage_b200.envs.normalized_env import NormalizedEnv, normalize

__all__ = ['BatchedEnv', 'PointEnv', 'PendulumEnv', 'CartPoleEnv', 'NormalizedEnv', 'normalize',
           'GymEnv', 'GridWorldEnv']
