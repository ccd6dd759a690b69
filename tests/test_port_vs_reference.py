"""Build-container only (needs /root/reference): pin oracle/port.py and oracle/np_oracle.py
against the LIVE reference -- the VecWorker/LocalSampler bookkeeping under replayed actions
(lengths, step types, episode order bit-exact) and a fresh PPO/VPG _train_once."""
import numpy as np
import pytest
import torch

from oracle import np_oracle as O
from oracle import port

pytestmark = pytest.mark.reference


class ReplayPolicy:
    """Feeds pre-recorded actions [T, n_envs, A] to a worker (Policy duck type)."""

    def __init__(self, actions):
        self.actions, self.t = actions, 0

    def get_actions(self, observations):
        a = self.actions[self.t]
        self.t += 1
        return a, {}

    def reset(self, do_resets=None):
        pass


def test_vec_worker_bookkeeping_under_replay():
    from oracle import ref_shims
    ref_shims.install()
    from garage.envs import PointEnv
    from garage.sampler import LocalSampler, VecWorker

    class NoInfoPointEnv(PointEnv):  # VecWorker breaks on non-empty episode_info (quirk Q1)
        def reset(self):
            obs, _ = super().reset()
            return obs, {}

    n_envs, max_len, T = 6, 12, 80
    rng = np.random.default_rng(3)
    goal = np.array((0.3, 0.25), dtype=np.float32)
    # envs drift towards the goal at different speeds -> different termination times
    speed = rng.uniform(0.2, 1.0, size=(1, n_envs, 1))
    actions = (speed * 0.1 * np.sign(goal)[None, None] + rng.standard_normal((T, n_envs, 2)) * 0.03).astype(np.float32)
    env = NoInfoPointEnv(goal=goal, max_episode_length=max_len)
    pol = ReplayPolicy(actions)
    s = LocalSampler(agents=pol, envs=env, max_episode_length=max_len, n_workers=1,
                     worker_class=VecWorker, worker_args=dict(n_envs=n_envs))
    ref = s.obtain_samples(0, 150, pol)
    envs = [port.PointNp(goal=tuple(goal), max_episode_length=max_len) for _ in range(n_envs)]
    got = port.vec_rollout(None, envs, max_len, 150, action_fn=lambda t, obs: actions[t])
    assert np.array_equal(got['lengths'], ref.lengths)                       # bit-exact
    assert np.array_equal(got['step_types'], np.asarray(ref.step_types, np.uint8))
    assert len(set(got['lengths'])) > 1                                      # ragged batch
    # Reference quirk Q12: VecWorker appends `self._prev_obs[i]`, a VIEW of a row that it later
    # overwrites in place (vec_worker.py:188,196), so with vector observations every stored
    # observation of an episode aliases the episode's LAST observation.  Pin the quirk, and check
    # the port's (correct, DefaultWorker-like) observations through next_observations instead.
    stops = np.cumsum(ref.lengths)
    for a, b, last in zip(stops - ref.lengths, stops, ref.last_observations):
        assert np.allclose(ref.observations[a:b], last[None])
        assert np.allclose(got['observations'][a], [0, 0, np.linalg.norm(goal)])   # reset obs
    assert np.array_equal(got['actions'], ref.actions)
    assert np.allclose(got['rewards'], ref.rewards, rtol=1e-6)
    assert np.allclose(got['last_observations'], ref.last_observations, rtol=1e-6, atol=1e-7)
    # the vectorised numpy replay used by the GPU tests gives the same batch
    rr = O.replay_rollout('point', actions, max_len, 0, env_kw=dict(goal=tuple(goal)))
    lengths, packed, order = O.pack_episodes(rr['step_type'], rr['ep_len'], 150,
                                             [rr['obs'], actions, rr['rew'], rr['step_type']])
    assert np.array_equal(lengths, ref.lengths) and order == got['order']
    assert np.allclose(packed[0], got['observations'], rtol=1e-6, atol=1e-7)
    assert np.allclose(packed[2], ref.rewards, rtol=1e-6)
    assert np.array_equal(packed[3], np.asarray(ref.step_types, np.uint8))


@pytest.mark.parametrize('kind', ['ppo', 'vpg'])
def test_port_iteration_equals_reference(kind):
    """Same seeds -> the port's sampler + _train_once reproduce the reference's parameters."""
    from oracle import ref_shims
    ref_shims.install()
    from garage.envs import PointEnv
    from garage.sampler import LocalSampler
    from garage.torch.algos import PPO, VPG
    from garage.torch.optimizers import OptimizerWrapper
    from garage.torch.policies import GaussianMLPPolicy
    from garage.torch.value_functions import GaussianMLPValueFunction
    from oracle.gen_golden import flat_of
    max_len = 20
    torch.manual_seed(5)
    env = PointEnv(max_episode_length=max_len)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=(32, 32))
    vf = GaussianMLPValueFunction(env.spec, hidden_sizes=(32, 32))
    sampler = LocalSampler(agents=policy, envs=env, max_episode_length=max_len, n_workers=1)
    mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=1e-3)), m, max_optimization_epochs=2,
                                    minibatch_size=32)
    cls = PPO if kind == 'ppo' else VPG
    algo = cls(env.spec, policy, vf, sampler, policy_optimizer=mk(policy), vf_optimizer=mk(vf),
               gae_lambda=0.9)
    pol = port.GaussianMLP(3, 2, (32, 32))
    vfp = port.GaussianMLP(3, 1, (32, 32), min_std=None)
    pol.load_flat(flat_of(policy.state_dict(), '_module.'))
    vfp.load_flat(flat_of(vf.state_dict(), 'module.'))
    eps = sampler.obtain_samples(0, 100, policy.get_param_values())
    np.random.seed(3)
    algo._train_once(0, eps)
    palgo = port.PPOPort(pol, vfp, max_len, kind=kind, lr=1e-3, epochs=2, minibatch_size=32, gae_lambda=0.9)
    batch = dict(observations=eps.observations, actions=eps.actions, rewards=eps.rewards, lengths=eps.lengths)
    np.random.seed(3)
    palgo.train_once(batch)
    assert np.allclose(pol.flat(), flat_of(policy.state_dict(), '_module.'), rtol=1e-5, atol=1e-6)
    assert np.allclose(vfp.flat(), flat_of(vf.state_dict(), 'module.'), rtol=1e-5, atol=1e-6)
