"""GaussianMLPPolicy drop-in (reference torch/policies/{policy,stochastic_policy,
gaussian_mlp_policy}.py): same constructor, `forward`, `get_action(s)`,
`get_param_values/set_param_values`, `reset`, `name`, and state_dict keys."""
import numpy as np
import torch
from torch import nn
from torch.distributions import Normal
from torch.distributions.independent import Independent

from garage_b200 import ops
from garage_b200.torch.modules import GaussianMLPModule


class Policy(nn.Module):
    """Reference torch/policies/policy.py:9-79."""

    def __init__(self, env_spec, name):
        super().__init__()
        self._env_spec = env_spec
        self._name = name

    def get_param_values(self):
        """state_dict (reference `policy.py:49-58`), detached copies."""
        return {k: v.detach().clone() for k, v in self.state_dict().items()}

    def set_param_values(self, state_dict):
        """load_state_dict (reference `policy.py:60-69`); CPU tensors are copied H2D."""
        self.load_state_dict(state_dict)

    def reset(self, do_resets=None):
        """Stateless policy: nothing to reset."""

    @property
    def name(self):
        return self._name

    @property
    def env_spec(self):
        return self._env_spec

    @property
    def observation_space(self):
        return self._env_spec.observation_space

    @property
    def action_space(self):
        return self._env_spec.action_space


class StochasticPolicy(Policy):
    """Reference torch/policies/stochastic_policy.py:11-104 (numpy in, numpy out)."""

    _sample_seed = 0
    _sample_counter = 0

    def _obs_to_device(self, observations):
        if isinstance(observations, list):
            observations = (torch.stack(observations) if isinstance(observations[0], torch.Tensor)
                            else np.stack([np.asarray(o) for o in observations]))
        if isinstance(observations, torch.Tensor):
            t = observations
        else:
            t = torch.from_numpy(np.asarray(observations, dtype=np.float32))
        t = t.to(device=self._module.device, dtype=torch.float32)
        return t.reshape(t.shape[0], -1).contiguous()

    def get_action(self, observation):
        if isinstance(observation, torch.Tensor):
            obs = observation.reshape(1, -1)
        else:
            obs = np.asarray(observation, dtype=np.float32).reshape(1, -1)
        action, infos = self.get_actions(obs)
        return action[0], {k: v[0] for k, v in infos.items()}

    def get_actions(self, observations):
        with torch.no_grad():
            obs = self._obs_to_device(observations)
            mean = self._module.mean(obs)
            log_std = self._module.log_std_value()
            # counter RNG instead of torch's global CPU generator (stochastic_policy.py:86)
            act, _ = ops.policy_sample(mean, log_std, self._sample_seed, 0, self._sample_counter,
                                       want_logp=False)
            self._sample_counter += 1
            return act.cpu().numpy(), dict(mean=mean.cpu().numpy(),
                                           log_std=np.full(mean.shape, np.float32(log_std)))


class GaussianMLPPolicy(StochasticPolicy):
    """MLP whose outputs are fed into a Normal distribution (reference
    torch/policies/gaussian_mlp_policy.py:13-102).  Parameters live in one flat CUDA
    vector (`self.flat`) that the rollout / update kernels use directly."""

    def __init__(self, env_spec, hidden_sizes=(32, 32), hidden_nonlinearity=torch.tanh,
                 hidden_w_init=nn.init.xavier_uniform_, hidden_b_init=nn.init.zeros_,
                 output_nonlinearity=None, output_w_init=nn.init.xavier_uniform_,
                 output_b_init=nn.init.zeros_, learn_std=True, init_std=1.0, min_std=1e-6,
                 max_std=None, std_parameterization='exp', layer_normalization=False,
                 name='GaussianMLPPolicy'):
        super().__init__(env_spec, name)
        self._obs_dim = env_spec.observation_space.flat_dim
        self._action_dim = env_spec.action_space.flat_dim
        self._module = GaussianMLPModule(
            input_dim=self._obs_dim, output_dim=self._action_dim, hidden_sizes=hidden_sizes,
            hidden_nonlinearity=hidden_nonlinearity, hidden_w_init=hidden_w_init,
            hidden_b_init=hidden_b_init, output_nonlinearity=output_nonlinearity,
            output_w_init=output_w_init, output_b_init=output_b_init, learn_std=learn_std,
            init_std=init_std, min_std=min_std, max_std=max_std,
            std_parameterization=std_parameterization, layer_normalization=layer_normalization)

    @property
    def flat(self):
        return self._module.flat

    @property
    def mlp_shape(self):
        return self._module.shape

    def forward(self, observations):
        """(dist, info) like the reference (`:89-102`); computed by the libb2g forward
        kernel, so the result carries no autograd graph -- training goes through the
        fused loss/backward kernels, not through this call."""
        obs = torch.as_tensor(observations, dtype=torch.float32).to(self._module.device)
        lead = obs.shape[:-1]
        mean = self._module.mean(obs.reshape(-1, obs.shape[-1]).contiguous()).reshape(*lead, self._action_dim)
        log_std = torch.full_like(mean, self._module.log_std_value())
        dist = Independent(Normal(mean, log_std.exp()), 1)
        return dist, dict(mean=mean, log_std=log_std)
