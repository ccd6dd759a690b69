"""GPU-resident environments of the on-policy hot path."""
from garage_b200.envs.batched import BatchedEnv, CartPoleEnv, PendulumEnv, PointEnv

__all__ = ['BatchedEnv', 'PointEnv', 'PendulumEnv', 'CartPoleEnv']
