from garage_b200.torch.algos.ppo import PPO
from garage_b200.torch.algos.trpo import TRPO
from garage_b200.torch.algos.vpg import VPG

__all__ = ['VPG', 'PPO', 'TRPO']
