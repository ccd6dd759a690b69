"""`log_performance`, `log_multitask_performance`, `obtain_evaluation_episodes`,
`obtain_exact_episodes` on the GPU: the reference's own known-answer test
(tests/garage/test_functions.py:50-98), the episode-stats kernel against the numpy oracle on
ragged batches, at configs[1] size against torch reductions, and the structure of evaluation
batches."""
import math

import numpy as np
import pytest
import torch

from oracle import np_oracle

pytestmark = pytest.mark.gpu


def _spec():
    from garage_b200 import EnvSpec
    from garage_b200.spaces import Box
    return EnvSpec(Box(np.zeros(3), np.ones(3)), Box(-np.ones(2), np.zeros(2)))


def _kat_batch():
    from garage_b200 import EpisodeBatch, StepType
    lengths = np.array([10, 5, 1, 1])
    n = int(lengths.sum())
    st = ([StepType.FIRST] + [StepType.MID] * 8 + [StepType.TERMINAL] + [StepType.FIRST] + [StepType.MID] * 3 +
          [StepType.TERMINAL] + [StepType.FIRST] + [StepType.FIRST])
    return EpisodeBatch(
        _spec(), observations=np.ones((n, 3), dtype=np.float32),
        last_observations=np.ones((4, 3), dtype=np.float32), actions=np.zeros((n, 2), dtype=np.float32),
        rewards=np.array([0.34026529, 0.58263177, 0.84307509, 0.97651095, 0.81723901, 0.22631398,
                          0.03421301, 0.97515046, 0.64311832, 0.65068933, 0.17657714, 0.04783857,
                          0.73904013, 0.41364329, 0.52235551, 0.24203526, 0.43328910]),
        step_types=np.array(st, dtype=StepType),
        env_infos={'success': np.array([0, 0, 0, 0, 0, 0, 1, 1, 0, 0, 0, 0, 0, 0, 1, 0, 1], dtype=bool)},
        agent_infos={}, episode_infos={}, lengths=lengths)


def test_log_performance_reference_known_answer():
    """Input and expected values of the reference's test_log_performance."""
    from garage_b200 import log_performance
    from garage_b200.tabular import tabular
    tabular.clear()
    und = log_performance(7, _kat_batch(), 0.8, prefix='test_log_performance')
    res = tabular.as_dict
    assert res['test_log_performance/Iteration'] == 7
    assert res['test_log_performance/NumEpisodes'] == 4
    assert math.isclose(res['test_log_performance/SuccessRate'], 0.75)
    assert math.isclose(res['test_log_performance/TerminationRate'], 0.5)
    assert math.isclose(res['test_log_performance/AverageDiscountedReturn'], 1.1131040640673113)
    assert math.isclose(res['test_log_performance/AverageReturn'], 2.1659965525)
    assert math.isclose(res['test_log_performance/StdReturn'], 2.354067152038576)
    assert math.isclose(res['test_log_performance/MaxReturn'], und.max()) and len(und) == 4
    np.testing.assert_allclose(und[2:], [0.24203526, 0.43328910], rtol=1e-15)


@pytest.mark.parametrize('dtype', [torch.float32, torch.float64])
def test_episode_stats_vs_oracle_ragged(dtype):
    from garage_b200 import ops
    rng = np.random.default_rng(3)
    lengths = np.concatenate([[1, 2, 31, 32, 33, 64, 65, 1000, 999], rng.integers(1, 300, size=500)])
    B = int(lengths.sum())
    rew = rng.normal(size=B)
    if dtype == torch.float32:
        rew = rew.astype(np.float32)
    st = np.ones(B, dtype=np.uint8)
    ends = np.cumsum(lengths) - 1
    st[ends - lengths + 1] = 0
    term = rng.random(len(lengths)) < 0.4
    st[ends] = np.where(term, 2, 3)
    succ = (rng.random(B) < 0.002).astype(np.uint8)
    dev = torch.device('cuda')
    per, out = ops.episode_stats(torch.as_tensor(rew).to(dev), 0.99, step_types=torch.as_tensor(st).to(dev),
                                 success=torch.as_tensor(succ).to(dev), lengths=torch.as_tensor(lengths).to(dev))
    per, out = per.cpu().numpy(), out.cpu().numpy()
    starts = ends - lengths + 1
    und = np.array([rew[a:b + 1].astype(np.float64).sum() for a, b in zip(starts, ends)])
    disc = np.array([np_oracle.discount_cumsum(rew[a:b + 1].astype(np.float64), 0.99)[0]
                     for a, b in zip(starts, ends)])
    sx = np.array([succ[a:b + 1].any() for a, b in zip(starts, ends)], dtype=np.float64)
    np.testing.assert_allclose(per[0], und, rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(per[1], disc, rtol=1e-12, atol=1e-12)
    np.testing.assert_array_equal(per[2], term.astype(np.float64))
    np.testing.assert_array_equal(per[3], sx)
    np.testing.assert_allclose(out, [disc.mean(), und.mean(), und.std(), und.max(), und.min(), term.mean(),
                                     sx.mean()], rtol=1e-12)
    # no success field -> NaN success rate; bit-reproducible
    per2, out2 = ops.episode_stats(torch.as_tensor(rew).to(dev), 0.99, step_types=torch.as_tensor(st).to(dev),
                                   lengths=torch.as_tensor(lengths).to(dev))
    assert np.isnan(out2[6].item()) and torch.equal(per2[:3].cpu(), torch.as_tensor(per[:3]))


def test_episode_stats_full_size_fixed_len():
    """configs[1]: 4,096 episodes x 200 steps, against torch float64 reductions."""
    from garage_b200 import ops
    dev = torch.device('cuda')
    g = torch.Generator(device=dev).manual_seed(5)
    n_eps, L = 4096, 200
    rew = torch.randn(n_eps * L, device=dev, generator=g)
    st = torch.ones(n_eps * L, dtype=torch.uint8, device=dev)
    st.view(n_eps, L)[:, -1] = 3
    per, out = ops.episode_stats(rew, 0.99, step_types=st, fixed_len=L)
    r = rew.view(n_eps, L).double()
    w = 0.99 ** torch.arange(L, device=dev, dtype=torch.float64)
    torch.testing.assert_close(per[0], r.sum(1), rtol=1e-12, atol=1e-12)
    torch.testing.assert_close(per[1], (r * w).sum(1), rtol=1e-11, atol=1e-11)
    assert per[2].sum().item() == 0 and out[5].item() == 0
    torch.testing.assert_close(out[2], r.sum(1).std(unbiased=False), rtol=1e-12, atol=0)
    lengths = torch.full((n_eps, ), L, dtype=torch.int64, device=dev)
    per_r, out_r = ops.episode_stats(rew, 0.99, step_types=st, lengths=lengths)
    assert torch.equal(per_r, per) and torch.equal(out_r[:6], out[:6])


def _policy(env, hidden=(64, 64)):
    from garage_b200.torch.policies import GaussianMLPPolicy
    torch.manual_seed(2)
    return GaussianMLPPolicy(env.spec, hidden_sizes=hidden)


def test_obtain_evaluation_episodes_deterministic():
    from garage_b200 import StepType, log_performance, obtain_evaluation_episodes
    from garage_b200.envs import CartPoleEnv, PendulumEnv
    from garage_b200.tabular import tabular
    env = CartPoleEnv(max_episode_length=60)
    pol = _policy(env)
    eps = obtain_evaluation_episodes(pol, env, max_episode_length=60, num_eps=100)
    assert len(eps.lengths) == 100 and eps.lengths.max() <= 60 and eps.lengths.min() >= 1
    assert eps.lengths.min() < 60  # an untrained policy drops the pole early somewhere
    np.testing.assert_array_equal(eps.dev.ep_env.cpu().numpy(), np.arange(100))  # one per env, env order
    # deterministic=True acts with the mean, bit for bit
    np.testing.assert_array_equal(eps.actions, eps.agent_infos['mean'])
    # (the stand-alone forward kernel sums in another order than the rollout kernel: 1e-6)
    np.testing.assert_allclose(eps.actions, pol._module.mean(eps.dev.observations).cpu().numpy(), atol=2e-6)
    ends = np.cumsum(eps.lengths) - 1
    st = np.array([int(s) for s in eps.step_types])
    assert (st[ends - eps.lengths + 1][eps.lengths > 1] == int(StepType.FIRST)).all()
    assert np.isin(st[ends], [int(StepType.TERMINAL), int(StepType.TIMEOUT)]).all()
    assert (st[ends][eps.lengths < 60] == int(StepType.TERMINAL)).all()
    tabular.clear()
    und = log_performance(0, eps, 0.99)
    np.testing.assert_allclose(und, eps.lengths.astype(np.float64))  # CartPole: reward 1 per step
    assert tabular.as_dict['Evaluation/TerminationRate'] == (eps.lengths < 60).mean()
    # stochastic evaluation differs from the mean; fixed-length env gives num_eps x L
    env2 = PendulumEnv(max_episode_length=40)
    pol2 = _policy(env2)
    e2 = obtain_evaluation_episodes(pol2, env2, max_episode_length=40, num_eps=7, deterministic=False)
    assert list(e2.lengths) == [40] * 7 and not np.array_equal(e2.actions, e2.agent_infos['mean'])
    host = e2.to_host()
    assert host.observations.shape == (280, 3)


def test_obtain_exact_episodes_env_major():
    from garage_b200.envs import CartPoleEnv
    from garage_b200.sampler import GPUSampler
    env = CartPoleEnv(max_episode_length=50)
    pol = _policy(env)
    s = GPUSampler(pol, env, max_episode_length=50, n_envs=33, seed=11)
    eps = s.obtain_exact_episodes(3, pol)
    assert eps.n_episodes == 99 and s.total_env_steps == eps.n_steps == int(eps.lengths.sum())
    np.testing.assert_array_equal(eps.dev.ep_env.cpu().numpy(), np.repeat(np.arange(33), 3))
    # the three episodes of an env follow each other in time
    end = eps.dev.ep_end.cpu().numpy().reshape(33, 3)
    ln = eps.lengths.reshape(33, 3)
    np.testing.assert_array_equal(end[:, 0] + 1, ln[:, 0])
    np.testing.assert_array_equal(end[:, 1] - end[:, 0], ln[:, 1])
    np.testing.assert_array_equal(end[:, 2] - end[:, 1], ln[:, 2])
    # same seed, same call -> the first episode of every env equals a k=1 collection
    s2 = GPUSampler(pol, env, max_episode_length=50, n_envs=33, seed=11)
    one = s2.obtain_exact_episodes(1, pol)
    np.testing.assert_array_equal(one.lengths, ln[:, 0])
    off3 = np.cumsum(eps.lengths) - eps.lengths
    first_rows = np.concatenate([np.arange(o, o + l) for o, l in zip(off3[::3], ln[:, 0])])
    np.testing.assert_array_equal(one.observations, eps.observations[first_rows])
    np.testing.assert_array_equal(one.rewards, eps.rewards[first_rows])
    with pytest.raises(ValueError):
        s.obtain_exact_episodes(0, pol)


def test_log_multitask_performance():
    from garage_b200 import EpisodeBatch, log_multitask_performance
    from garage_b200.tabular import tabular
    b = _kat_batch()
    infos = dict(b.env_infos)
    infos['task_id'] = np.repeat([0, 1, 0, 2], b.lengths)
    mt = EpisodeBatch(b.env_spec, episode_infos={}, observations=b.observations,
                      last_observations=b.last_observations, actions=b.actions, rewards=b.rewards,
                      env_infos=infos, agent_infos={}, step_types=b.step_types, lengths=b.lengths)
    tabular.clear()
    und = log_multitask_performance(3, mt, 0.8, name_map={0: 'a', 1: 'b', 2: 'c', 3: 'd'})
    d = tabular.as_dict
    assert d['a/NumEpisodes'] == 2 and d['b/NumEpisodes'] == 1 and d['c/NumEpisodes'] == 1
    assert d['d/NumEpisodes'] == 0 and np.isnan(d['d/AverageReturn'])
    assert d['Average/NumEpisodes'] == 4 and len(und) == 4
    assert math.isclose(d['Average/AverageDiscountedReturn'], 1.1131040640673113)
    assert math.isclose(d['b/AverageReturn'], b.rewards[10:15].sum())
    assert d['a/TerminationRate'] == 0.5 and d['b/TerminationRate'] == 1.0
