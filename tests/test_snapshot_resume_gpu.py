"""Snapshot -> restore -> resume on the GPU path (reference
tests/garage/experiment/test_resume.py, test_snapshotter_integration.py, examples
torch/ppo_pendulum.py + torch/resume_training.py): a launcher written like the reference's
runs through `wrap_experiment`, `Trainer(ctxt)` writes cloudpickle snapshots, and a second
trainer restored from the first epoch's snapshot ends on the SAME parameters as the
uninterrupted run (rollout RNG counters, shuffle seeds, Adam moments and step all travel)."""
import csv
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _launcher(tmp_path, hidden, env_id, n_envs, mb):
    from garage_b200 import Trainer, wrap_experiment
    from garage_b200.envs import GymEnv
    from garage_b200.experiment.deterministic import set_seed
    from garage_b200.sampler import LocalSampler, VecWorker
    from garage_b200.torch.algos import PPO
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction

    @wrap_experiment(log_dir=str(tmp_path / 'run'), snapshot_mode='all', use_existing_dir=True)
    def ppo_pendulum(ctxt=None, seed=1, n_epochs=3):
        set_seed(seed)
        env = GymEnv(env_id, max_episode_length=50)
        trainer = Trainer(ctxt)
        policy = GaussianMLPPolicy(env.spec, hidden_sizes=hidden, hidden_nonlinearity=torch.tanh,
                                   output_nonlinearity=None)
        value_function = GaussianMLPValueFunction(env_spec=env.spec, hidden_sizes=hidden,
                                                  hidden_nonlinearity=torch.tanh,
                                                  output_nonlinearity=None)
        sampler = LocalSampler(agents=policy, envs=env, max_episode_length=env.spec.max_episode_length,
                               n_workers=1, worker_class=VecWorker, worker_args=dict(n_envs=n_envs))
        algo = PPO(env_spec=env.spec, policy=policy, value_function=value_function, sampler=sampler,
                   policy_optimizer=OptimizerWrapper((torch.optim.Adam, dict(lr=1e-3)), policy,
                                                     max_optimization_epochs=2, minibatch_size=mb),
                   vf_optimizer=OptimizerWrapper((torch.optim.Adam, dict(lr=1e-3)), value_function,
                                                 max_optimization_epochs=2, minibatch_size=mb),
                   discount=0.99, center_adv=False)
        trainer.setup(algo, env)
        trainer.train(n_epochs=n_epochs, batch_size=n_envs * 50)
        return trainer, algo

    return ppo_pendulum


@pytest.mark.parametrize('hidden, env_id', [((64, 64), 'Pendulum-v0'), ((32, 32), 'CartPole-v1')])
def test_resume_reproduces_uninterrupted_run(tmp_path, hidden, env_id):
    from garage_b200 import SnapshotConfig, Trainer
    n_envs, mb = 128, 1600
    trainer, algo = _launcher(tmp_path, hidden, env_id, n_envs, mb)(seed=3, n_epochs=3)
    run = tmp_path / 'run'
    assert {'itr_0.pkl', 'itr_1.pkl', 'itr_2.pkl', 'progress.csv', 'variant.json'} <= set(os.listdir(run))
    with open(run / 'progress.csv', newline='') as fh:
        rows = list(csv.DictReader(fh))
    assert len(rows) == 3
    for key in ('Evaluation/AverageReturn', 'GaussianMLPPolicy/KL', 'GaussianMLPValueFunction/LossAfter',
                'TotalEnvSteps'):
        assert key in rows[0], key
    steps = [int(float(r['TotalEnvSteps'])) for r in rows]
    if env_id == 'Pendulum-v0':  # fixed-length episodes; CartPole batches end on whole episodes
        assert steps == [6400, 12800, 19200]
    assert steps[0] >= 6400 and steps[1] - steps[0] >= 6400 and steps[2] == trainer.total_env_steps
    full_p = algo.policy.flat.cpu().numpy().copy()
    full_v = algo._value_function.flat.cpu().numpy().copy()
    full_log = trainer.log

    # a fresh process would do exactly this (examples/torch/resume_training.py)
    t2 = Trainer(SnapshotConfig(str(tmp_path / 'resumed'), 'last', 1))
    args = t2.restore(str(run), from_epoch=0)
    assert (args.start_epoch, args.n_epochs, args.batch_size) == (1, 3, 6400)
    assert t2.total_env_steps == steps[0]
    a2 = t2._algo
    assert a2 is not algo and a2.policy.flat.is_cuda
    assert not np.array_equal(a2.policy.flat.cpu().numpy(), full_p)
    t2.resume()
    assert t2.total_env_steps == steps[2] and len(t2.log) == 2
    np.testing.assert_array_equal(a2.policy.flat.cpu().numpy(), full_p)
    np.testing.assert_array_equal(a2._value_function.flat.cpu().numpy(), full_v)
    for k in ('Evaluation/AverageReturn', 'GaussianMLPPolicy/LossAfter'):
        assert t2.log[-1][k] == full_log[-1][k]
    assert os.path.exists(tmp_path / 'resumed' / 'params.pkl')

    # 'last' picks the final snapshot; resuming it with more epochs continues from there
    t3 = Trainer()
    assert t3.restore(str(run)).start_epoch == 3
    t3.resume(n_epochs=4)
    assert t3.total_env_steps >= steps[2] + 6400 and len(t3.log) == 1


def test_state_dict_interchange_after_restore(tmp_path):
    """The snapshot's algo exposes the same state_dict keys as before pickling, and loading them
    into a fresh policy gives the same actions' means (checkpoint interchange)."""
    from garage_b200.envs import PendulumEnv
    from garage_b200.experiment import Snapshotter
    from garage_b200.torch.policies import GaussianMLPPolicy
    env = PendulumEnv(max_episode_length=20)
    torch.manual_seed(0)
    pol = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
    s = Snapshotter(str(tmp_path), 'last', 1)
    s.save_itr_params(0, dict(policy=pol, env=env))
    back = s.load(str(tmp_path))
    p2, e2 = back['policy'], back['env']
    assert list(p2.state_dict()) == list(pol.state_dict())
    for k, v in pol.state_dict().items():
        assert torch.equal(v.cpu(), p2.state_dict()[k].cpu()), k
    obs = torch.randn(16, 3, device='cuda')
    assert torch.equal(pol._module.mean(obs), p2._module.mean(obs))
    assert e2.spec.max_episode_length == 20
