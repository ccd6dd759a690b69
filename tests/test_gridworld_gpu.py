"""GridWorldEnv + ScriptedPolicy on the device against the reference's own VecWorker /
LocalSampler (the setting of tests/garage/sampler/test_vec_worker.py).  The
fixture tests/golden/gridworld.npz was written by oracle/gen_golden.py::gen_gridworld running the
unmodified reference; everything here is integer work, so the bar is bit-exact."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'gridworld.npz')


def _case(d, k):
    return {name[len('c%d_' % k):]: d[name] for name in d.files if name.startswith('c%d_' % k)}


@pytest.mark.parametrize('k', range(6))
def test_sampler_matches_reference_vecworker(k):
    from garage_b200 import StepType
    from garage_b200.envs import GridWorldEnv
    from garage_b200.np.policies import ScriptedPolicy
    from garage_b200.sampler import LocalSampler, VecWorker
    c = _case(np.load(GOLD), k)
    env = GridWorldEnv(desc=str(c['desc']), max_episode_length=int(c['max_len']))
    policy = ScriptedPolicy(scripted_actions=[int(a) for a in c['table']])
    sampler = LocalSampler(agents=policy, envs=env, max_episode_length=int(c['max_len']), n_workers=1,
                           worker_class=VecWorker, worker_args=dict(n_envs=int(c['n_envs'])))
    eps = sampler.obtain_samples(0, int(c['num_samples']), policy)
    np.testing.assert_array_equal(eps.lengths, c['lengths'])
    np.testing.assert_array_equal(eps.observations, c['observations'])
    np.testing.assert_array_equal(eps.actions, c['actions'])
    np.testing.assert_array_equal(eps.last_observations, c['last_observations'])
    np.testing.assert_array_equal(eps.rewards, c['rewards'])
    np.testing.assert_array_equal(np.array([int(s) for s in eps.step_types], dtype=np.uint8), c['step_types'])
    assert eps.observations.dtype == np.int64 and eps.observations.shape == c['observations'].shape
    assert eps.agent_infos == {} and eps.step_types.dtype == StepType
    assert sampler.total_env_steps == int(c['lengths'].sum())
    host = eps.to_host()  # passes EpisodeBatch validation with Discrete spaces
    assert len(host.split()) == len(c['lengths'])
    # a second call restarts every env from reset(): same batch
    again = sampler.obtain_samples(1, int(c['num_samples']), policy)
    np.testing.assert_array_equal(again.observations, c['observations'])


def test_env_step_api_matches_rollout_and_reference_semantics():
    """Environment.reset/step one env at a time follows the same trajectory as the fixture's
    first episode; errors as the reference (step before reset, step after the episode ended)."""
    from garage_b200.envs import GridWorldEnv
    c = _case(np.load(GOLD), 0)
    env = GridWorldEnv(desc='4x4', max_episode_length=16)
    with pytest.raises(RuntimeError):
        env.step(0)
    obs, info = env.reset()
    assert obs == 0 and info == {}
    n0 = int(c['lengths'][0])
    for t in range(n0):
        assert obs == c['observations'][t]
        es = env.step(int(c['table'][obs]))
        assert es.reward == c['rewards'][t] and int(es.step_type) == c['step_types'][t]
        obs = es.observation
    assert es.terminal and obs == c['last_observations'][0] == 15 and es.reward == 1.0
    with pytest.raises(RuntimeError):
        env.step(0)  # grid_world_env.py:171-172: reset() needed after the episode ended
    # walls and borders: 4x4_safe, moving into a wall or off the grid stays in place
    safe = GridWorldEnv(desc='4x4_safe', max_episode_length=None)
    safe.reset()
    assert safe.step(0).observation == 0 and safe.step(3).observation == 0  # left / up at the corner
    assert safe.step(1).observation == 4 and safe.step(2).observation == 4  # (1,1) is a wall
    assert safe.spec.observation_space.n == 16 and safe.spec.action_space.n == 4


def test_rollout_many_envs_bookkeeping():
    """16,384 copies of the 8x8 map, random table, no time limit on the env but on the sampler:
    every env walks the same path, so all columns are equal and the packed batch is the
    reference-ordered repetition of one episode."""
    from garage_b200.envs import GridWorldEnv
    from garage_b200.np.policies import ScriptedPolicy
    from garage_b200.sampler import GPUSampler
    c = _case(np.load(GOLD), 5)
    N = 16384
    env = GridWorldEnv(desc='8x8', max_episode_length=200)
    s = GPUSampler(ScriptedPolicy([int(a) for a in c['table']]), env, max_episode_length=200, n_envs=N)
    eps = s.obtain_samples(0, N * 5 * 3, None)
    L = int(c['lengths'][0])
    assert eps.n_episodes == 3 * N and (eps.lengths == L).all()
    obs = eps.dev.observations.view(3 * N, L)
    assert torch.equal(obs, obs[:1].expand_as(obs))
    np.testing.assert_array_equal(obs[0].cpu().numpy().astype(np.int64), c['observations'][:L])
    np.testing.assert_array_equal(eps.dev.ep_env.view(3, N).cpu().numpy(), np.tile(np.arange(N), (3, 1)))
