from garage_b200.torch.policies.gaussian_mlp_policy import GaussianMLPPolicy, Policy, StochasticPolicy

__all__ = ['Policy', 'StochasticPolicy', 'GaussianMLPPolicy']
