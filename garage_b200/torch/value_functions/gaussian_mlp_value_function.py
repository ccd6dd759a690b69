"""GaussianMLPValueFunction drop-in (reference torch/value_functions/
{value_function,gaussian_mlp_value_function}.py)."""
import torch
from torch import nn

from garage_b200 import ops
from garage_b200.torch.modules import GaussianMLPModule


class ValueFunction(nn.Module):
    """Reference torch/value_functions/value_function.py."""

    def __init__(self, env_spec, name):
        super().__init__()
        self._mdp_spec = env_spec
        self.name = name


class GaussianMLPValueFunction(ValueFunction):
    """Gaussian MLP value function with a learned scalar std and NO std clamp
    (`min_std=None`, reference `gaussian_mlp_value_function.py:46-79`)."""

    def __init__(self, env_spec, hidden_sizes=(32, 32), hidden_nonlinearity=torch.tanh,
                 hidden_w_init=nn.init.xavier_uniform_, hidden_b_init=nn.init.zeros_,
                 output_nonlinearity=None, output_w_init=nn.init.xavier_uniform_,
                 output_b_init=nn.init.zeros_, learn_std=True, init_std=1.0,
                 layer_normalization=False, name='GaussianMLPValueFunction'):
        super().__init__(env_spec, name)
        self.module = GaussianMLPModule(
            input_dim=env_spec.observation_space.flat_dim, output_dim=1, hidden_sizes=hidden_sizes,
            hidden_nonlinearity=hidden_nonlinearity, hidden_w_init=hidden_w_init,
            hidden_b_init=hidden_b_init, output_nonlinearity=output_nonlinearity,
            output_w_init=output_w_init, output_b_init=output_b_init, learn_std=learn_std,
            init_std=init_std, min_std=None, max_std=None, std_parameterization='exp',
            layer_normalization=layer_normalization)

    @property
    def flat(self):
        return self.module.flat

    @property
    def mlp_shape(self):
        return self.module.shape

    def compute_loss(self, obs, returns):
        """-mean log N(returns; V(obs), std) (reference `:81-97`) as a 0-dim tensor
        (no autograd graph: gradients come from libb2g's fused backward)."""
        obs = torch.as_tensor(obs, dtype=torch.float32).to(self.module.device).reshape(-1, self.module._input_dim).contiguous()
        ret = torch.as_tensor(returns, dtype=torch.float32).to(self.module.device).reshape(-1).contiguous()
        n = obs.shape[0]
        nb = ops.mlp_eval(self.module.shape, self.flat, obs, ret=ret, loss_kind=ops.LOSS_VF)
        stats = torch.empty(ops.MLP_NSTATS, dtype=torch.float64, device=self.module.device)
        ops.mlp_reduce(self.module.shape, nb, 1.0, stats_out=stats)
        return (stats[0] / n).to(torch.float32)

    def forward(self, obs):
        """Baselines: obs `(..., O)` -> `(...)` (reference `:100-112`)."""
        obs = torch.as_tensor(obs, dtype=torch.float32).to(self.module.device)
        lead = obs.shape[:-1]
        return self.module.mean(obs.reshape(-1, obs.shape[-1]).contiguous()).reshape(*lead)
