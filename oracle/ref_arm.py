"""TEST INFRASTRUCTURE ONLY -- the reference arm of bench.py: the UNMODIFIED garage v2021.03.0
package (imported from /root/reference/src or its staged copy oracle/_ref through
oracle/ref_shims.py) running its own stock code path on the host cores:

    LocalSampler(worker_class=VecWorker)          sampler/local_sampler.py:134-166,
                                                  sampler/vec_worker.py:107-219
    garage.torch.algos.PPO.train path             torch/algos/vpg.py:136-293, ppo.py:96-132
    GaussianMLPPolicy / GaussianMLPValueFunction  torch/policies, torch/value_functions
    OptimizerWrapper + BatchDataset               torch/optimizers/optimizer_wrapper.py

The ONLY piece that is not the reference's own code is the environment: the reference reaches
Pendulum through `GymEnv('Pendulum-v0')` and gym==0.17.2 is not on either box (setup.py:8,38),
so `PendulumRef` / `CartPoleRef` implement garage's `Environment` interface
(_environment.py:128-249) around the oracle's restatement of gym 0.17.2's step functions
(np_oracle.pendulum_step / cartpole_step) -- the same dynamics the CUDA kernels are checked
against.  Nothing under garage_b200/ imports this module.
"""
import math
import time

import numpy as np

from oracle import np_oracle as O
from oracle import ref_shims


def _env_classes():
    """Environment subclasses are created after ref_shims.install() (they derive from the
    reference's own base class)."""
    import akro
    from garage import Environment, EnvSpec, EnvStep, StepType

    class _ClassicControl(Environment):
        obs_dim, act_dim = 0, 0

        def __init__(self, max_episode_length, seed):
            self._max_episode_length = max_episode_length
            self._rng = np.random.RandomState(seed)
            self._step_cnt = None
            self._obs_space = akro.Box(low=-np.inf, high=np.inf, shape=(self.obs_dim, ), dtype=np.float32)
            self._act_space = self._make_action_space()
            self._spec = EnvSpec(action_space=self._act_space, observation_space=self._obs_space,
                                 max_episode_length=max_episode_length)

        observation_space = property(lambda self: self._obs_space)
        action_space = property(lambda self: self._act_space)
        spec = property(lambda self: self._spec)
        render_modes = property(lambda self: [])

        def render(self, mode):
            return None

        def visualize(self):
            pass

        def close(self):
            pass

        def _finish(self, action, reward, obs, done):
            self._step_cnt += 1
            st = StepType.get_step_type(step_cnt=self._step_cnt,
                                        max_episode_length=self._max_episode_length, done=done)
            if st in (StepType.TERMINAL, StepType.TIMEOUT):
                self._step_cnt = None
            return EnvStep(env_spec=self._spec, action=action, reward=float(reward), observation=obs,
                           env_info={}, step_type=st)

    class PendulumRef(_ClassicControl):
        """gym 0.17.2 Pendulum-v0 (classic_control/pendulum.py) behind garage's Environment."""
        obs_dim, act_dim = 3, 1

        def _make_action_space(self):
            return akro.Box(low=-2.0, high=2.0, shape=(1, ), dtype=np.float32)

        def reset(self):
            high = np.array([np.pi, 1])
            self._state = self._rng.uniform(low=-high, high=high)[None]
            self._step_cnt = 0
            return O.pendulum_obs(self._state)[0], {}

        def step(self, action):
            if self._step_cnt is None:
                raise RuntimeError('reset() must be called before step()!')
            self._state, obs, rew, _ = O.pendulum_step(self._state, np.asarray(action, np.float32)[None])
            return self._finish(action, rew[0], obs[0], False)

    class CartPoleRef(_ClassicControl):
        """gym 0.17.2 CartPole-v1 dynamics with a continuous push (sign of the action)."""
        obs_dim, act_dim = 4, 1

        def _make_action_space(self):
            return akro.Box(low=-1.0, high=1.0, shape=(1, ), dtype=np.float32)

        def reset(self):
            self._state = self._rng.uniform(low=-0.05, high=0.05, size=(4, ))[None]
            self._step_cnt = 0
            return self._state[0].astype(np.float32), {}

        def step(self, action):
            if self._step_cnt is None:
                raise RuntimeError('reset() must be called before step()!')
            self._state, obs, rew, done = O.cartpole_step(self._state, np.asarray(action, np.float32)[None])
            return self._finish(action, rew[0], obs[0], bool(done[0]))

    return dict(pendulum=PendulumRef, cartpole=CartPoleRef)


class ReferenceIteration:
    """One garage PPO (or VPG) iteration = `sampler.obtain_samples` + `algo._train_once`,
    exactly the two calls of `VPG.train` (torch/algos/vpg.py:208-228)."""

    def __init__(self, env_name='pendulum', n_envs=32, horizon=200, hidden=64, epochs=10,
                 minibatch_size=None, kind='ppo', gae_lambda=0.95, seed=1):
        if not ref_shims.install():
            raise RuntimeError('reference package not present (neither /root/reference nor oracle/_ref)')
        import torch
        from garage.sampler import LocalSampler, VecWorker
        from garage.torch.algos import PPO, VPG
        from garage.torch.optimizers import OptimizerWrapper
        from garage.torch.policies import GaussianMLPPolicy
        from garage.torch.value_functions import GaussianMLPValueFunction
        torch.manual_seed(seed)
        np.random.seed(seed)
        cls = _env_classes()[env_name]
        envs = [cls(horizon, seed + i) for i in range(n_envs)]
        spec = envs[0].spec
        self.policy = GaussianMLPPolicy(spec, hidden_sizes=(hidden, hidden))
        self.vf = GaussianMLPValueFunction(spec, hidden_sizes=(hidden, hidden))
        # one worker that steps all n_envs copies in lock-step: the VecWorker configuration of the
        # north star.  LocalSampler hands worker i the i-th env update; VecWorker takes a list.
        self.sampler = LocalSampler(agents=self.policy, envs=envs[0], max_episode_length=horizon,
                                    n_workers=1, worker_class=VecWorker, worker_args=dict(n_envs=n_envs))
        w = self.sampler._workers[0]
        w.update_env(envs)
        mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), m,
                                        max_optimization_epochs=epochs, minibatch_size=minibatch_size)
        algo_cls = PPO if kind == 'ppo' else VPG
        kw = dict(lr_clip_range=0.2) if kind == 'ppo' else {}
        self.algo = algo_cls(spec, self.policy, self.vf, self.sampler, policy_optimizer=mk(self.policy),
                             vf_optimizer=mk(self.vf), discount=0.99, gae_lambda=gae_lambda,
                             center_adv=True, **kw)
        self.batch = n_envs * horizon
        self.itr = 0

    def run(self):
        """Returns (env steps, rollout seconds, update seconds)."""
        t0 = time.perf_counter()
        eps = self.sampler.obtain_samples(self.itr, self.batch, self.policy.get_param_values())
        t1 = time.perf_counter()
        self.algo._train_once(self.itr, eps)
        t2 = time.perf_counter()
        self.itr += 1
        return int(np.sum(eps.lengths)), t1 - t0, t2 - t1


def time_iteration(iters=1, **kw):
    """Same return value as oracle.port.time_iteration, measured on the real reference."""
    it = ReferenceIteration(**kw)
    steps, tr, tt = 0, 0.0, 0.0
    for _ in range(iters):
        s, a, b = it.run()
        steps += s
        tr += a
        tt += b
    return dict(env_steps=steps, rollout_s=tr, train_s=tt, total_s=tr + tt)


def time_gae_reference(T=200, N=4096, discount=0.99, gae_lambda=0.95, reps=3):
    """The reference's own GAE and return code on [N, T] float32/float64 arrays:
    garage.torch.compute_advantages (torch/_functions.py:25-85: conv2d with a length-T filter) and
    garage.np.discount_cumsum (np/_functions.py:111-128: scipy lfilter), called per episode row as
    vpg.py:149-153 does.  Returns seconds per call of each and the 17 B/element rate."""
    if not ref_shims.install():
        raise RuntimeError('reference package not present')
    import torch
    from garage.np import discount_cumsum
    from garage.torch import compute_advantages
    rng = np.random.default_rng(0)
    rew = torch.from_numpy(rng.standard_normal((N, T)).astype(np.float32))
    val = torch.from_numpy(rng.standard_normal((N, T)).astype(np.float32))
    rew64 = rew.numpy().astype(np.float64)
    compute_advantages(discount, gae_lambda, T, val, rew)
    t0 = time.perf_counter()
    for _ in range(reps):
        compute_advantages(discount, gae_lambda, T, val, rew)
    t_adv = (time.perf_counter() - t0) / reps
    t0 = time.perf_counter()
    for _ in range(reps):
        np.stack([discount_cumsum(r, discount) for r in rew64])
    t_ret = (time.perf_counter() - t0) / reps
    total = t_adv + t_ret
    return dict(T=T, N=N, compute_advantages_s=t_adv, discount_cumsum_s=t_ret,
                gbs=17.0 * T * N / total / 1e9, threads=torch.get_num_threads())


if __name__ == '__main__':
    print(time_iteration(iters=1, n_envs=8, horizon=50, epochs=2, minibatch_size=200))
    print(time_gae_reference(T=200, N=1024))
