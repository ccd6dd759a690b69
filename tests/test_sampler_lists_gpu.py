"""Per-worker `agents` / `envs` lists of `LocalSampler` (reference sampler/local_sampler.py:50-86,
156-166 and sampler/worker_factory.py:68-95): every worker gets its own env (multi-task sampling),
`obtain_samples` visits the workers round-robin, takes what one `rollout()` of a worker completes
and stops as soon as the total reaches `num_samples`.

The referee is an independent replay: each worker's episodes come from a stand-alone single-env
sampler with the same seed and env-id offset (already pinned against the oracle in
tests/test_rollout_gpu.py), and the visiting order is recomputed here from their completion rows.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(goals, L, per):
    from garage_b200.envs import PointEnv
    from garage_b200.torch.policies import GaussianMLPPolicy
    torch.manual_seed(3)
    envs = [PointEnv(goal=g, max_episode_length=L) for g in goals]
    policy = GaussianMLPPolicy(envs[0].spec, hidden_sizes=(32, 32))
    return envs, policy


def _expected(policy, envs, L, per, seed, num_samples):
    """Replay of local_sampler.py:156-166 over stand-alone samplers."""
    from garage_b200.sampler import GPUSampler
    groups = []
    for i, env in enumerate(envs):
        s = GPUSampler(policy, env, max_episode_length=L, n_envs=per, seed=seed, env_id_offset=i * per)
        b = s.obtain_samples(0, num_samples, policy.get_param_values())
        ends = b.dev.ep_end.cpu().numpy()
        host = b.to_host()
        starts = np.concatenate([[0], np.cumsum(host.lengths)])
        rows = sorted(set(ends.tolist()))
        groups.append((host, starts, [np.nonzero(ends == r)[0] for r in rows]))
    order, total, r = [], 0, 0
    while True:
        for gi, (host, starts, per_row) in enumerate(groups):
            before_visit = total
            for e in per_row[r]:
                order.append((gi, int(e)))
                total += int(host.lengths[e])
            if total >= num_samples:
                return groups, order, total, before_visit
        r += 1


@pytest.mark.parametrize('goals,num_samples', [
    ([(1.0, 1.0), (-2.0, 0.5), (0.3, -0.4)], 500),      # all episodes run to the time limit
    ([(3.0, 3.0), (0.05, 0.05)], 450),                   # the second worker's episodes end at once: ragged rounds
])
def test_per_worker_env_lists_round_robin(goals, num_samples):
    from garage_b200.sampler import LocalSampler, VecWorker
    L, per, seed = 25, 4, 11
    envs, policy = _setup(goals, L, per)
    s = LocalSampler(agents=policy, envs=envs, max_episode_length=L, n_workers=len(envs),
                     worker_class=VecWorker, worker_args=dict(n_envs=per), seed=seed)
    assert s.n_envs == per * len(envs)
    batch = s.obtain_samples(0, num_samples, policy.get_param_values())
    groups, order, total, before_visit = _expected(policy, envs, L, per, seed, num_samples)
    assert int(batch.lengths.sum()) == total >= num_samples > before_visit  # stops at the first visit that suffices
    assert s.total_env_steps == total
    assert len(batch.lengths) == len(order)
    pos = 0
    for k, (g, e) in enumerate(order):
        host, starts, _ = groups[g]
        n = int(host.lengths[e])
        sl, bl = slice(starts[e], starts[e] + n), slice(pos, pos + n)
        assert batch.lengths[k] == n
        assert np.array_equal(batch.observations[bl], host.observations[sl])
        assert np.array_equal(batch.actions[bl], host.actions[sl])
        assert np.array_equal(batch.rewards[bl], host.rewards[sl])
        assert np.array_equal(batch.step_types[bl], host.step_types[sl])
        assert np.array_equal(batch.last_observations[k], host.last_observations[e])
        assert np.allclose(batch.episode_infos_by_episode['goal'][k], goals[g])          # the worker's own task
        assert np.array_equal(batch.agent_infos['mean'][bl], host.agent_infos['mean'][sl])
        pos += n
    assert order[0][0] == 0  # the first visit is worker 0


def test_lists_of_the_same_objects_collapse_and_bad_lengths_raise():
    from garage_b200.sampler import LocalSampler, VecWorker
    envs, policy = _setup([(1.0, 1.0)], 10, 4)
    s = LocalSampler(agents=[policy, policy], envs=[envs[0], envs[0]], max_episode_length=10, n_workers=2,
                     worker_class=VecWorker, worker_args=dict(n_envs=4), seed=1)
    assert s._groups is None and s.n_envs == 8
    b = s.obtain_samples(0, 80, policy.get_param_values())
    assert int(np.sum(b.lengths)) >= 80
    with pytest.raises(ValueError, match="Length of list doesn't match number of workers"):
        LocalSampler(agents=policy, envs=[envs[0]] * 3, max_episode_length=10, n_workers=2)


def test_ppo_trains_on_a_per_worker_batch():
    """The host EpisodeBatch of the grouped path goes through `_train_once` like a reference batch."""
    from garage_b200.sampler import LocalSampler, VecWorker
    from garage_b200.torch.algos import PPO
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    L, per = 20, 8
    envs, policy = _setup([(1.0, 1.0), (-1.0, 2.0)], L, per)
    vf = GaussianMLPValueFunction(envs[0].spec, hidden_sizes=(32, 32))
    s = LocalSampler(agents=policy, envs=envs, max_episode_length=L, n_workers=2, worker_class=VecWorker,
                     worker_args=dict(n_envs=per), seed=2)
    mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=1e-3)), m, max_optimization_epochs=2,
                                    minibatch_size=64)
    algo = PPO(envs[0].spec, policy, vf, s, policy_optimizer=mk(policy), vf_optimizer=mk(vf))
    before = policy.flat.clone()
    batch = s.obtain_samples(0, 2 * per * L, policy.get_param_values())
    ret = algo._train_once(0, batch)
    assert np.isfinite(ret)
    assert not torch.equal(before, policy.flat)
    ex = s.obtain_exact_episodes(2, policy.get_param_values())
    assert len(ex.lengths) == 2 * 2 * per
