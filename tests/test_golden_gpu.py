"""GPU parity against fixtures produced by RUNNING the unmodified reference
(oracle/gen_golden.py -> tests/golden/*.npz): GAE/returns, PointEnv transitions under
replayed actions, PPO/VPG/value loss + gradients + Adam, and a whole PPO._train_once (C1)."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import np_oracle as O

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
TOL = 1e-5  # north_star: 1e-5 relative in fp32; integers bit-exact


def load(name):
    return np.load(os.path.join(G, name), allow_pickle=False)


def rel(got, want):
    return float(np.abs(np.asarray(got, np.float64) - want).max() / max(np.abs(want).max(), 1e-30))


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize('case', 'abcd')
def test_gae_returns_vs_reference(case):
    from garage_b200 import ops
    from garage_b200.np import discount_cumsum
    from garage_b200.torch import compute_advantages
    d = load('gae.npz')
    g, l = d[case + '_cfg']
    b, r = d[case + '_baselines'], d[case + '_rewards']
    adv = compute_advantages(g, l, b.shape[1], torch.from_numpy(b), torch.from_numpy(r)).numpy()
    assert np.allclose(adv, d[case + '_adv'], rtol=TOL, atol=TOL)
    ret = discount_cumsum(r.astype(np.float64), g)
    assert np.allclose(ret, d[case + '_ret'], rtol=1e-12, atol=1e-12)
    st = np.full(b.T.shape, O.MID, np.uint8)
    st[-1] = O.TIMEOUT
    a2, r2 = ops.gae_time_major(dev(r.T), dev(b.T), dev(st), discount=g, gae_lambda=l)
    assert np.allclose(a2.cpu().numpy().T, d[case + '_adv'], rtol=TOL, atol=TOL)
    assert np.allclose(r2.cpu().numpy().T, d[case + '_ret'], rtol=TOL, atol=TOL)


@pytest.mark.parametrize('case', 'abc')
def test_point_env_replay_vs_reference(case):
    from garage_b200 import ops
    d = load('point_env.npz')
    gx, gy, arena, bonus, never, max_len = d[case + '_params']
    es = ops.EnvState(0, 1, torch.device('cuda:0'))
    prm = ops.env_params_array([gx, gy, arena, bonus, never])
    ops.env_reset(es, prm, 0)
    assert np.array_equal(es.obs.cpu().numpy()[0], d[case + '_obs'][0])
    for t, a in enumerate(d[case + '_actions']):
        obs, rew, st, succ = ops.env_step(es, prm, int(max_len), dev(a[None]), want_success=True)
        assert np.allclose(obs.cpu().numpy()[0], d[case + '_obs'][t + 1], rtol=TOL, atol=1e-7)
        assert abs(rew.item() - d[case + '_rewards'][t]) <= TOL * max(1, abs(rew.item()))
        assert int(st.item()) == int(d[case + '_step_types'][t])          # bit-exact
        assert bool(succ.item()) == bool(d[case + '_success'][t])          # bit-exact


@pytest.mark.parametrize('case', ['p64', 'p32', 'c64'])
def test_loss_grad_adam_vs_reference(case):
    from garage_b200 import ops
    d = load('losses.npz')
    O_, A, H, n = (int(v) for v in d[case + '_dims'])
    device = torch.device('cuda:0')
    obs, act, adv = dev(d[case + '_obs']), dev(d[case + '_act']), dev(d[case + '_adv'])
    ret, olp = dev(d[case + '_ret']), dev(d[case + '_old_logp'])
    psh, vsh = ops.MlpShape(O_, H, A, device), ops.MlpShape(O_, H, 1, device)
    pf, vfl = dev(d[case + '_policy_flat']), dev(d[case + '_vf_flat'])
    lo = math.log(1e-6)

    def check(shape, nb, want_loss, want_grad, A_):
        grad = torch.empty(shape.P, device=device)
        stats = torch.empty(8, dtype=torch.float64, device=device)
        ops.mlp_reduce(shape, nb, 1.0 / n, grad_out=grad, stats_out=stats)
        assert abs(stats[0].item() / n - want_loss) <= TOL * max(1, abs(want_loss))
        got = O.unflatten_params(grad.cpu().numpy().astype(np.float64), O_, H, A_)
        for k, v in O.unflatten_params(want_grad.astype(np.float64), O_, H, A_).items():
            assert rel(got[k], v) < TOL, k

    for kind, code in (('ppo', ops.LOSS_PPO), ('vpg', ops.LOSS_VPG)):
        nb = ops.mlp_minibatch_grad(psh, pf, obs, act=act, adv=adv, old_logp=olp, loss_kind=code,
                                    min_log_std=lo)
        check(psh, nb, float(d['{}_{}_loss'.format(case, kind)]), d['{}_{}_grad'.format(case, kind)], A)
    nb = ops.mlp_minibatch_grad(vsh, vfl, obs, ret=ret, loss_kind=ops.LOSS_VF)
    check(vsh, nb, float(d[case + '_vf_loss']), d[case + '_vf_grad'], 1)
    # KL(old || new) and entropy statistics
    nb = ops.mlp_eval(psh, pf, obs, act=act, adv=adv, old_logp=olp, old_mean=dev(d[case + '_old_mean']),
                      old_log_std=torch.tensor([float(d[case + '_old_log_std'])], device=device),
                      loss_kind=ops.LOSS_PPO, min_log_std=lo)
    stats = torch.empty(8, dtype=torch.float64, device=device)
    ops.mlp_reduce(psh, nb, 1.0, stats_out=stats)
    kl, ent = stats[1].item() / n, stats[2].item() / n
    assert abs(kl - d[case + '_kl']) <= 1e-4 * max(1e-3, abs(kl))
    assert abs(ent - d[case + '_entropy']) < 1e-5
    # Adam: the reference's 3 policy + 3 value-function steps on the same index lists
    m, v, step = torch.zeros_like(pf), torch.zeros_like(pf), torch.zeros(1, dtype=torch.int64, device=device)
    mq, vq, stq = torch.zeros_like(vfl), torch.zeros_like(vfl), torch.zeros(1, dtype=torch.int64, device=device)
    p0, q0 = pf.clone(), vfl.clone()
    for mb in d[case + '_adam_idx']:
        idx = dev(mb.astype(np.int32))
        nb = ops.mlp_minibatch_grad(psh, pf, obs, idx=idx, act=act, adv=adv, old_logp=olp,
                                    loss_kind=ops.LOSS_PPO, min_log_std=lo, step_counter=step)
        ops.mlp_reduce(psh, nb, 1.0 / len(mb), params=pf, adam_m=m, adam_v=v, step_counter=step, lr=2.5e-4)
        nb = ops.mlp_minibatch_grad(vsh, vfl, obs, idx=idx, ret=ret, loss_kind=ops.LOSS_VF, step_counter=stq)
        ops.mlp_reduce(vsh, nb, 1.0 / len(mb), params=vfl, adam_m=mq, adam_v=vq, step_counter=stq, lr=2.5e-4)
    assert np.allclose(pf.cpu().numpy(), d[case + '_policy_flat_after'], rtol=TOL, atol=1e-6)
    assert np.allclose(vfl.cpu().numpy(), d[case + '_vf_flat_after'], rtol=TOL, atol=1e-6)
    assert rel((pf - p0).cpu().numpy(), (d[case + '_policy_flat_after'] - d[case + '_policy_flat']).astype(np.float64)) < 5e-3


@pytest.mark.parametrize('case', ['fixed', 'early'])
def test_train_once_c1_vs_reference(case):
    """BASELINE configs[0] end to end: the reference's own EpisodeBatch (LocalSampler on
    PointEnv) -> garage_b200 PPO._train_once with the reference's minibatch shuffles
    (shuffle='numpy', np.random.seed) and its padding quirk (pad_value_mode='reference')
    -> the reference's parameters and dowel values."""
    from garage_b200 import EpisodeBatch
    from garage_b200._dtypes import step_types_from_codes
    from garage_b200.envs import PointEnv
    from garage_b200.torch.algos import PPO
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    d = load('train_once.npz')
    max_len, gx, gy = d[case + '_cfg']
    env = PointEnv(goal=np.array([gx, gy], np.float32), max_episode_length=int(max_len))
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=(32, 32))
    vf = GaussianMLPValueFunction(env.spec, hidden_sizes=(32, 32))
    policy.flat.copy_(dev(d[case + '_policy_flat']))
    vf.flat.copy_(dev(d[case + '_vf_flat']))
    mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), m, max_optimization_epochs=3,
                                    minibatch_size=64)
    algo = PPO(env.spec, policy, vf, None, policy_optimizer=mk(policy), vf_optimizer=mk(vf),
               gae_lambda=0.95, shuffle='numpy', pad_value_mode='reference')
    n = len(d[case + '_rew'])
    eps = EpisodeBatch(env_spec=env.spec, episode_infos={}, observations=d[case + '_obs'],
                       last_observations=d[case + '_last_obs'], actions=d[case + '_act'],
                       rewards=d[case + '_rew'], env_infos=dict(success=d[case + '_success']),
                       agent_infos={}, step_types=step_types_from_codes(d[case + '_step_types']),
                       lengths=d[case + '_lengths'])
    np.random.seed(11)
    ret = algo._train_once(0, eps)
    logs = dict(zip(d[case + '_log_keys'], d[case + '_log_vals']))
    assert abs(ret - float(d[case + '_return'])) <= 1e-6 * abs(ret)
    assert np.allclose(policy.flat.cpu().numpy(), d[case + '_policy_flat_after'], rtol=TOL, atol=1e-6)
    assert np.allclose(vf.flat.cpu().numpy(), d[case + '_vf_flat_after'], rtol=TOL, atol=1e-6)
    upd = (d[case + '_policy_flat_after'] - d[case + '_policy_flat']).astype(np.float64)
    assert rel(policy.flat.cpu().numpy() - d[case + '_policy_flat'], upd) < 1e-2
    for k, want in logs.items():
        got = algo.last_stats[k]
        tol = 1e-4 * max(abs(want), 1e-3) if ('Loss' in k or 'KL' in k) else 1e-5 * max(abs(want), 1.0)
        assert abs(got - want) <= tol, (k, got, want)
    assert n == int(d[case + '_lengths'].sum())
