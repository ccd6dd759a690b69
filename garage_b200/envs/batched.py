"""GPU-resident environments.

Each class is BOTH a garage `Environment` (single-env `reset()` / `step()`, same
signatures and errors as the reference, computed by libb2g's step kernel on a
1-env device state) AND a description (`env_id`, `params`) that
`garage_b200.sampler.GPUSampler` uses to run thousands of copies inside the
persistent rollout kernel.
"""
import math

import numpy as np
import torch

from garage_b200 import ops
from garage_b200._dtypes import StepType
from garage_b200._environment import Environment, EnvSpec, EnvStep, finite_len
from garage_b200.spaces import Box


class BatchedEnv(Environment):
    env_id = None

    def __init__(self, max_episode_length):
        self._max_episode_length = max_episode_length
        self._es = None
        self._seed = 0
        self._spec = EnvSpec(action_space=self.action_space, observation_space=self.observation_space,
                             max_episode_length=max_episode_length)

    # -- description used by the GPU sampler
    def params(self):
        raise NotImplementedError

    def params_array(self):
        return ops.env_params_array(self.params())

    @property
    def max_len(self):
        return finite_len(self._max_episode_length)

    def episode_info(self):
        return {}

    def env_info_consts(self):
        return {}

    has_success = False
    never_terminates = False  # True: episodes only end at max_episode_length (static shapes)

    # -- Environment API
    @property
    def spec(self):
        return self._spec

    def seed(self, seed):
        self._seed = int(seed)

    def _state(self):
        if self._es is None:
            self._es = ops.EnvState(self.env_id, 1, torch.device('cuda', torch.cuda.current_device()))
        return self._es

    def reset(self):
        es = self._state()
        ops.env_reset(es, self.params_array(), self._seed)
        return self._cast_obs(es.obs[0].cpu().numpy()), dict(self.episode_info())

    def step(self, action):
        es = self._state()
        act = torch.as_tensor(np.asarray(action, dtype=np.float32).reshape(1, -1), device=es.device)
        obs, rew, st, succ = ops.env_step(es, self.params_array(), self.max_len, act,
                                          want_success=self.has_success)
        info = dict(self.env_info_consts())
        if self.has_success:
            info['success'] = bool(succ.item())
        return EnvStep(env_spec=self.spec, action=action, reward=float(rew.item()),
                       observation=self._cast_obs(obs[0].cpu().numpy()), env_info=info,
                       step_type=StepType(int(st.item())))

    def _cast_obs(self, obs):
        return obs

    def __getstate__(self):
        d = self.__dict__.copy()
        d['_es'] = None  # device handles are rebuilt lazily after unpickling
        return d


class PointEnv(BatchedEnv):
    """2-D point-to-goal task; reference `envs/point_env.py:13-211`."""
    env_id = ops.ENV_POINT
    has_success = True

    def __init__(self, goal=np.array((1., 1.), dtype=np.float32), arena_size=5., done_bonus=0.,
                 never_done=False, max_episode_length=math.inf):
        goal = np.array(goal, dtype=np.float32)
        assert ((goal >= -arena_size) & (goal <= arena_size)).all()
        self._goal = goal
        self._arena_size = arena_size
        self._done_bonus = done_bonus
        self._never_done = never_done
        self._task = {'goal': self._goal}
        self._observation_space = Box(low=-np.inf, high=np.inf, shape=(3, ), dtype=np.float32)
        self._action_space = Box(low=-0.1, high=0.1, shape=(2, ), dtype=np.float32)
        super().__init__(max_episode_length)

    @property
    def action_space(self):
        return self._action_space

    @property
    def observation_space(self):
        return self._observation_space

    @property
    def render_modes(self):
        return ['ascii']

    @property
    def never_terminates(self):
        return bool(self._never_done)

    def params(self):
        return [float(self._goal[0]), float(self._goal[1]), float(self._arena_size),
                float(self._done_bonus), 1.0 if self._never_done else 0.0]

    def episode_info(self):
        return dict(goal=self._goal)

    def env_info_consts(self):
        return dict(task=self._task)

    def step(self, action):
        if self._es is None or int(self._es.step_cnt.item()) < 0:
            raise RuntimeError('reset() must be called before step()!')
        return super().step(action)

    def render(self, mode):
        es = self._state()
        return 'Point: {}, Goal: {}'.format(es.state[:, 0].cpu().numpy().astype(np.float32), self._goal)

    def sample_tasks(self, num_tasks):
        goals = np.random.uniform(-2, 2, size=(num_tasks, 2))
        return [{'goal': g} for g in goals]

    def set_task(self, task):
        self._task = task
        self._goal = np.array(task['goal'], dtype=np.float32)


class PendulumEnv(BatchedEnv):
    """`GymEnv('Pendulum-v0')` of the reference (envs/gym_env.py:115-270 around
    gym==0.17.2 classic_control/pendulum.py), TimeLimit 200."""
    env_id = ops.ENV_PENDULUM
    never_terminates = True

    def __init__(self, max_episode_length=200, max_speed=8., max_torque=2., dt=.05, g=10.0, m=1., l=1.):
        self._p = [max_speed, max_torque, dt, g, m, l]
        high = np.array([1., 1., max_speed], dtype=np.float32)
        self._observation_space = Box(low=-high, high=high, dtype=np.float32)
        self._action_space = Box(low=-max_torque, high=max_torque, shape=(1, ), dtype=np.float32)
        super().__init__(max_episode_length)

    @property
    def action_space(self):
        return self._action_space

    @property
    def observation_space(self):
        return self._observation_space

    def params(self):
        return list(self._p)


class CartPoleEnv(BatchedEnv):
    """CartPole-v1 dynamics (gym==0.17.2 classic_control/cartpole.py) with a
    1-d continuous action for Gaussian policies: action > 0 pushes right
    (`act_mode=0`, gym's two discrete forces) or force = F * clip(a, -1, 1)
    (`act_mode=1`).  TimeLimit 500."""
    env_id = ops.ENV_CARTPOLE

    def __init__(self, max_episode_length=500, act_mode=0):
        self._p = [9.8, 1.0, 0.1, 0.5, 10.0, 0.02, 2.4, 12 * 2 * math.pi / 360, float(act_mode)]
        high = np.array([4.8, np.finfo(np.float32).max, 24 * 2 * math.pi / 360, np.finfo(np.float32).max],
                        dtype=np.float32)
        self._observation_space = Box(low=-high, high=high, dtype=np.float32)
        self._action_space = Box(low=-1., high=1., shape=(1, ), dtype=np.float32)
        super().__init__(max_episode_length)

    @property
    def action_space(self):
        return self._action_space

    @property
    def observation_space(self):
        return self._observation_space

    def params(self):
        return list(self._p)
