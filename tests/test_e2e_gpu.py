"""End-to-end on the GPU through the garage-compatible API: sampler -> EpisodeBatch ->
PPO/VPG._train_once -> Trainer."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

# state_dict() of the reference's GaussianMLPPolicy (printed by running the reference here)
REF_POLICY_KEYS = [
    '_module._init_std', '_module.min_std_param',
    '_module._mean_module._layers.0.linear.weight', '_module._mean_module._layers.0.linear.bias',
    '_module._mean_module._layers.1.linear.weight', '_module._mean_module._layers.1.linear.bias',
    '_module._mean_module._output_layers.0.linear.weight',
    '_module._mean_module._output_layers.0.linear.bias',
]


def build(env, hidden=(64, 64), n_envs=64, algo='ppo', epochs=2, mb=None, **algo_kw):
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.algos import PPO, VPG
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    torch.manual_seed(1)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=hidden)
    vf = GaussianMLPValueFunction(env.spec, hidden_sizes=hidden)
    sampler = GPUSampler(policy, env, max_episode_length=env.spec.max_episode_length, n_envs=n_envs,
                         seed=1234)
    po = OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), policy, max_optimization_epochs=epochs,
                          minibatch_size=mb)
    vo = OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), vf, max_optimization_epochs=epochs,
                          minibatch_size=mb)
    cls = PPO if algo == 'ppo' else VPG
    a = cls(env.spec, policy, vf, sampler, policy_optimizer=po, vf_optimizer=vo, discount=0.99,
            gae_lambda=0.95, **algo_kw)
    return a, policy, vf, sampler


def test_state_dict_keys_match_reference():
    from garage_b200.envs import PointEnv
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    env = PointEnv(max_episode_length=50)
    pol = GaussianMLPPolicy(env.spec)
    sd = pol.state_dict()
    assert list(sd.keys()) == REF_POLICY_KEYS
    assert sd['_module._mean_module._layers.0.linear.weight'].shape == (32, 3)
    assert sd['_module._mean_module._output_layers.0.linear.weight'].shape == (2, 32)
    assert sd['_module._init_std'].shape == (1, )
    vf = GaussianMLPValueFunction(env.spec)
    assert list(vf.state_dict().keys()) == [k.replace('_module.', 'module.', 1) for k in REF_POLICY_KEYS if 'min_std' not in k]
    # parameters are views of the flat vector the kernels use: a CPU state_dict loads through
    cpu = {k: torch.randn_like(v).cpu() for k, v in sd.items() if 'min_std' not in k}
    cpu['_module.min_std_param'] = sd['_module.min_std_param'].cpu()
    pol.set_param_values(cpu)
    assert torch.equal(pol.flat[0:1].cpu(), cpu['_module._init_std'])
    assert torch.equal(pol.flat[1:1 + 96].cpu(), cpu['_module._mean_module._layers.0.linear.weight'].reshape(-1))
    # xavier/zeros initialisation as in the reference: zero biases, log_std = log(1) = 0
    pol2 = GaussianMLPPolicy(env.spec)
    assert float(pol2.flat[0]) == 0.0
    assert torch.count_nonzero(pol2.state_dict()['_module._mean_module._layers.0.linear.bias']) == 0


def test_get_actions_api():
    from garage_b200.envs import PointEnv
    from garage_b200.torch.policies import GaussianMLPPolicy
    env = PointEnv(max_episode_length=50)
    pol = GaussianMLPPolicy(env.spec)
    obs, _ = env.reset()
    a, info = pol.get_action(obs)
    assert a.shape == (2, ) and info['mean'].shape == (2, ) and info['log_std'].shape == (2, )
    acts, infos = pol.get_actions(np.stack([obs] * 5))
    assert acts.shape == (5, 2) and infos['mean'].shape == (5, 2)
    assert np.allclose(infos['mean'][0], infos['mean'][4]) and np.allclose(infos['log_std'], 0.0)
    dist, info = pol(torch.from_numpy(np.stack([obs] * 3)))
    assert dist.mean.shape == (3, 2) and np.allclose(dist.mean.cpu().numpy(), infos['mean'][:3], atol=1e-6)
    es = env.step(a)
    assert es.observation.shape == (3, ) and isinstance(es.reward, float)


def test_sampler_semantics_vecworker():
    """161 samples, max length 16, 8 env copies: the VecWorker loop completes 8 episodes
    per 16 steps -> 16 whole episodes (256 steps), in env order inside each completion row."""
    from garage_b200 import EpisodeBatch
    from garage_b200.envs import PointEnv
    from garage_b200.sampler import LocalSampler
    from garage_b200.torch.policies import GaussianMLPPolicy
    # goal out of reach within 16 steps: every episode ends by TIMEOUT (dynamic-shape path)
    env = PointEnv(goal=np.array((3., 3.), dtype=np.float32), max_episode_length=16)
    pol = GaussianMLPPolicy(env.spec)
    s = LocalSampler(agents=pol, envs=env, max_episode_length=16, n_envs=8)
    eps = s.obtain_samples(0, 161, pol.get_param_values())
    assert list(eps.lengths) == [16] * 16 and eps.lengths.dtype == np.int64
    assert s.total_env_steps == 256
    assert eps.observations.shape == (256, 3) and eps.observations.dtype == np.float32
    assert eps.rewards.dtype == np.float64 and eps.actions.shape == (256, 2)
    st = eps.step_types
    assert st[0] == 0 and st[15] == 3 and (st[1:15] == 1).all()
    assert set(eps.agent_infos) == {'mean', 'log_std'} and 'success' in eps.env_infos
    assert eps.episode_infos_by_episode['goal'].shape == (16, 2)
    assert eps.padded_observations.shape == (16, 16, 3) and eps.valids.shape == (16, 16)
    host = eps.to_host()
    assert isinstance(host, EpisodeBatch) and len(host.split()) == 16
    # first observation of every episode is the reset observation (0, 0, |goal|)
    first = eps.observations[::16]
    assert np.allclose(first, [[0, 0, math.sqrt(18)]] * 16)
    # next_observations ends with last_observations
    assert np.array_equal(eps.next_observations[15], eps.last_observations[0])
    with pytest.raises(TypeError):
        LocalSampler(agents=pol, envs=env)


@pytest.mark.parametrize('algo,mb', [('ppo', 2000), ('ppo', None), ('vpg', None)])
def test_train_once_pendulum(algo, mb):
    from garage_b200.envs import PendulumEnv
    env = PendulumEnv(max_episode_length=50)
    a, policy, vf, sampler = build(env, n_envs=128, algo=algo, mb=mb)
    before = policy.flat.clone()
    eps = sampler.obtain_samples(0, 128 * 50, policy.get_param_values())
    assert eps.n_steps == 6400 and eps.n_episodes == 128
    r = a._train_once(0, eps)
    s = a.last_stats
    assert np.isfinite(r) and r < 0
    assert abs(s['GaussianMLPPolicy/KLBefore']) < 1e-6
    assert s['GaussianMLPPolicy/KL'] >= 0 and s['GaussianMLPPolicy/KL'] < 0.1
    assert abs(s['GaussianMLPPolicy/Entropy'] - (0.5 + 0.5 * math.log(2 * math.pi))) < 1e-2
    assert s['GaussianMLPValueFunction/LossAfter'] < s['GaussianMLPValueFunction/LossBefore']
    assert s['Evaluation/NumEpisodes'] == 128 and s['Evaluation/TerminationRate'] == 0.0
    assert not torch.equal(before, policy.flat)
    # log_performance cross-check on the host copy of the batch
    und = eps.rewards.reshape(128, 50).sum(1)
    assert abs(s['Evaluation/AverageReturn'] - und.mean()) < 1e-3 * abs(und.mean())
    assert abs(s['Evaluation/MaxReturn'] - und.max()) < 1e-3 * abs(und.max())


def test_trainer_loop_cartpole_vpg_and_point_ppo():
    from garage_b200 import Trainer
    from garage_b200.envs import CartPoleEnv, PointEnv
    env = CartPoleEnv(max_episode_length=100)
    a, policy, vf, sampler = build(env, n_envs=256, algo='vpg', epochs=1)
    tr = Trainer()
    tr.setup(a, env)
    ret = tr.train(n_epochs=3, batch_size=5000)
    assert np.isfinite(ret) and 5 < ret <= 100
    assert tr.total_env_steps >= 15000 and len(tr.log) == 3
    assert 0 < tr.log[-1]['Evaluation/TerminationRate'] <= 1.0
    env = PointEnv(max_episode_length=30)
    a, policy, vf, sampler = build(env, hidden=(32, 32), n_envs=64, algo='ppo', epochs=2, mb=512)
    tr = Trainer()
    tr.setup(a, env)
    ret = tr.train(n_epochs=2, batch_size=1000)
    assert np.isfinite(ret) and 'Evaluation/SuccessRate' in tr.log[-1]


def test_ppo_learns_point_env():
    """Reference-style 'does it learn' check (tests/garage/torch/algos/test_ppo.py is a
    MuJoCo learning test): mean return on PointEnv improves markedly."""
    from garage_b200 import Trainer
    from garage_b200.envs import PointEnv
    env = PointEnv(max_episode_length=40)
    a, policy, vf, sampler = build(env, hidden=(32, 32), n_envs=512, algo='ppo', epochs=4, mb=4096)
    a._policy_optimizer.lr = 3e-3
    a._vf_optimizer.lr = 3e-3
    tr = Trainer()
    tr.setup(a, env)
    tr.train(n_epochs=40, batch_size=512 * 40)
    first, last = tr.log[0]['Evaluation/AverageReturn'], tr.log[-1]['Evaluation/AverageReturn']
    assert last > first + 10, (first, last)


def test_baselines_pass_also_gives_vf_loss_before():
    """Fixed-length batches: the baselines pass doubles as the value function's LossBefore pass
    (returns are computed first).  Advantages and returns are bit-identical to the two-pass form
    and the statistics equal a separate evaluation on the packed batch."""
    from garage_b200 import ops
    from garage_b200.envs import CartPoleEnv, PendulumEnv
    env = PendulumEnv(max_episode_length=50)
    a, policy, vf, sampler = build(env, n_envs=256, algo='ppo', epochs=2, mb=1600)
    eps = sampler.obtain_samples(0, 256 * 50, policy)
    a._log_stats = True
    merged = a._prepare_device(eps)
    a._log_stats = False
    plain = a._prepare_device(eps)
    assert merged.vf_before is not None and plain.vf_before is None
    assert torch.equal(merged.adv, plain.adv) and torch.equal(merged.ret, plain.ret)
    out = torch.zeros(8, dtype=torch.float64, device='cuda')
    nb = ops.mlp_eval(vf.mlp_shape, vf.flat, plain.obs, ret=plain.ret, loss_kind=ops.LOSS_VF)
    ops.mlp_reduce(vf.mlp_shape, nb, 1.0, stats_out=out)
    torch.testing.assert_close(merged.vf_before, out, rtol=1e-6, atol=0)
    v, r = vf.module.mean(plain.obs).reshape(-1).double(), plain.ret.double()
    want = (0.5 * (r - v) ** 2 + 0.5 * math.log(2 * math.pi)).sum()
    torch.testing.assert_close(merged.vf_before[0], want, rtol=1e-6, atol=0)
    # ragged batches (early termination) keep the separate pass
    env2 = CartPoleEnv(max_episode_length=50)
    a2, policy2, _, sampler2 = build(env2, n_envs=128, algo='ppo', epochs=1, mb=None)
    f2 = a2._prepare_device(sampler2.obtain_samples(0, 128 * 50, policy2))
    assert f2.vf_before is None


def test_ppo_loss_before_without_forward_pass():
    """Before the update the likelihood ratio is 1 on every row (the reference's old policy is a
    synchronised copy), so PPO's LossBefore is -mean(adv) and KLBefore 0; the forward-pass form
    agrees to rounding.  VPG and entropy-regularised PPO keep the forward pass."""
    from garage_b200.envs import PendulumEnv
    env = PendulumEnv(max_episode_length=40)
    a, policy, vf, sampler = build(env, n_envs=128, algo='ppo', epochs=1, mb=None)
    f = a._prepare_device(sampler.obtain_samples(0, 128 * 40, policy))
    fast = torch.zeros(16, dtype=torch.float64, device='cuda')
    full = torch.zeros(16, dtype=torch.float64, device='cuda')
    a._evaluate(f, fast, before=True)
    a._evaluate(f, full, before=False)
    B = float(f.B)
    assert abs(fast[0].item() / B - full[0].item() / B) < 1e-6
    assert fast[1].item() == 0.0 and abs(full[1].item() / B) < 1e-9   # KL(old || new)
    torch.testing.assert_close(fast[8:16], full[8:16])                # value-function block untouched
    torch.testing.assert_close(fast[0], -f.adv.sum(dtype=torch.float64))
    v, _, _, s2 = build(env, n_envs=128, algo='vpg', epochs=1, mb=None)
    fv = v._prepare_device(s2.obtain_samples(0, 128 * 40, v.policy))
    o1 = torch.zeros(16, dtype=torch.float64, device='cuda')
    o2 = torch.zeros(16, dtype=torch.float64, device='cuda')
    v._evaluate(fv, o1, before=True)
    v._evaluate(fv, o2, before=False)
    assert torch.equal(o1, o2) and o1[2].item() != 0.0  # same forward pass either way (entropy sum present)


def test_graphed_iteration_with_host_io_equals_eager_api():
    """`GraphedIteration(algo, B, host_params=...)`: one CUDA graph holds the H2D copy of the policy
    parameters, the whole iteration and the D2H copies (EpisodeBatch, logging scalars, updated
    parameters).  Same bits as the eager public-API sequence obtain_samples(host state_dict) ->
    _train_once -> state_dict, iteration after iteration."""
    from garage_b200.envs import PendulumEnv
    from garage_b200.torch.algos.graphed import GraphedIteration
    B = 64 * 40

    def make():
        a, policy, vf, sampler = build(PendulumEnv(max_episode_length=40), n_envs=64, epochs=2, mb=512, shuffle='device')
        host = {k: v.detach().cpu().pin_memory() for k, v in policy.state_dict().items()}
        return a, policy, sampler, host
    a1, p1, s1, h1 = make()
    eager_rets, eager_obs = [], None
    for it in range(5):
        eps = s1.obtain_samples(it, B, h1)
        eps.prefetch_host()
        eager_rets.append(a1._train_once(it, eps))
        eager_obs = np.array(eps.observations)
        for k, v in p1.state_dict().items():
            h1[k].copy_(v)
    a2, p2, s2, h2 = make()
    runner = GraphedIteration(a2, B, warmup=2, host_params=h2)
    graph_rets = [runner.run(it) for it in range(5)]
    assert runner._graph is not None and runner.h2d_bytes > 0 and runner.d2h_bytes > B * 4 * 3
    assert graph_rets == eager_rets
    for k in h1:
        assert torch.equal(h1[k], h2[k]), k
    assert np.array_equal(np.asarray(runner.eps.observations), eager_obs)
    assert torch.equal(p1.flat, p2.flat)
