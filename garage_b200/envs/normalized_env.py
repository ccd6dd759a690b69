"""`normalize(env, ...)`: action rescaling + running observation / reward normalisation
(reference `envs/normalized_env.py:8-171`, `normalize = NormalizedEnv` :171), fused into the
device step: the wrapper is a DESCRIPTION (`norm_config()`) that the rollout kernel, the
reset kernel and the single-step kernel apply per env copy; the running estimates of every
copy live on the device (`ops.EnvState.norm_stats`, float64 like the reference's numpy).
"""
import numpy as np
import torch

from garage_b200 import ops
from garage_b200._dtypes import StepType
from garage_b200._environment import EnvStep
from garage_b200.envs.batched import BatchedEnv
from garage_b200.spaces import Box


class NormalizedEnv(BatchedEnv):
    """Drop-in for `garage.envs.normalize` around a `BatchedEnv`.

    Args (reference :29-39): env, scale_reward=1., normalize_obs=False, normalize_reward=False,
        expected_action_scale=1., flatten_obs=True, obs_alpha=0.001, reward_alpha=0.001.
    """

    def __init__(self, env, scale_reward=1., normalize_obs=False, normalize_reward=False,
                 expected_action_scale=1., flatten_obs=True, obs_alpha=0.001, reward_alpha=0.001):
        if not isinstance(env, BatchedEnv):
            raise TypeError('normalize() wraps a garage_b200 BatchedEnv; got {}'.format(type(env)))
        if isinstance(env, NormalizedEnv):
            raise NotImplementedError('nested normalize() wrappers are not built')
        if not flatten_obs and len(env.observation_space.shape) != 1:
            raise NotImplementedError('flatten_obs=False needs flat observations')
        self._env = env
        self._scale_reward = scale_reward
        self._normalize_obs = normalize_obs
        self._normalize_reward = normalize_reward
        self._expected_action_scale = expected_action_scale
        self._flatten_obs = flatten_obs
        self._obs_alpha = obs_alpha
        self._reward_alpha = reward_alpha
        self._es = None
        self._seed = 0

    # ---- the wrapped env's description, passed through (garage Wrapper.__getattr__)
    env_id = property(lambda self: self._env.env_id)
    has_success = property(lambda self: self._env.has_success)
    never_terminates = property(lambda self: self._env.never_terminates)
    max_len = property(lambda self: self._env.max_len)
    spec = property(lambda self: self._env.spec)
    action_space = property(lambda self: self._env.action_space)
    observation_space = property(lambda self: self._env.observation_space)
    render_modes = property(lambda self: self._env.render_modes)
    unwrapped = property(lambda self: self._env)

    def params(self):
        return self._env.params()

    def params_array(self):
        return self._env.params_array()

    def episode_info(self):
        return self._env.episode_info()

    def env_info_consts(self):
        return self._env.env_info_consts()

    def __getattr__(self, name):
        if name.startswith('_'):
            raise AttributeError(name)
        return getattr(self._env, name)

    def norm_config(self):
        """`ops.NormConfig` for the kernels (flags as the reference evaluates them, :91-101)."""
        flags = 0
        sp = self._env.action_space
        lo, hi = np.zeros(2, np.float32), np.zeros(2, np.float32)
        if isinstance(sp, Box):
            lb, ub = sp.low, sp.high
            # the reference tests `ub != -inf` (sic); kept
            if np.all(lb != -np.inf) and np.all(ub != -np.inf):
                flags |= ops.NORM_ACTION
                lo[:lb.size], hi[:ub.size] = lb.ravel(), ub.ravel()
        if self._normalize_obs:
            flags |= ops.NORM_OBS
        if self._normalize_reward:
            flags |= ops.NORM_REWARD
        return ops.NormConfig(flags, lo, hi, self._expected_action_scale, self._scale_reward,
                              self._obs_alpha, self._reward_alpha)

    # ---- single-env Environment API (one env copy on the device)
    def seed(self, seed):
        self._seed = int(seed)

    def reset(self):
        es = self._state()
        ops.env_reset(es, self.params_array(), self._seed, norm=self.norm_config())
        return self._obs_out(es.obs[0]), dict(self.episode_info())

    def step(self, action):
        if self._es is None or int(self._es.step_cnt.item()) < 0:
            raise RuntimeError('reset() must be called before step()!')
        es = self._es
        act = torch.as_tensor(np.asarray(action, dtype=np.float32).reshape(1, -1), device=es.device)
        obs, rew, st, succ = ops.env_step(es, self.params_array(), self.max_len, act,
                                          want_success=self.has_success, norm=self.norm_config())
        info = dict(self.env_info_consts())
        if self.has_success:
            info['success'] = bool(succ.item())
        return EnvStep(env_spec=self.spec, action=action, reward=float(rew.item()),
                       observation=self._obs_out(obs[0]), env_info=info,
                       step_type=StepType(int(st.item())))

    def _obs_out(self, obs_dev):
        obs = obs_dev.cpu().numpy()
        # normalised observations are float64 in the reference (float32 obs minus float64 mean)
        return obs.astype(np.float64) if self._normalize_obs else obs

    def statistics(self):
        """(obs_mean, obs_var, reward_mean, reward_var) of the single-env copy (host numpy)."""
        if self._es is None or self._es.norm_stats is None:
            O = self.observation_space.flat_dim
            return np.zeros(O), np.ones(O), 0., 1.
        s = self._es.norm_stats[:, 0].cpu().numpy()
        O = (s.size - 2) // 2
        return s[:O], s[O:2 * O], float(s[2 * O]), float(s[2 * O + 1])

    def __getstate__(self):
        d = self.__dict__.copy()
        d['_es'] = None
        return d

    def __setstate__(self, d):
        self.__dict__.update(d)


normalize = NormalizedEnv
