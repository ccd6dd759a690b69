"""Pin the oracle (oracle/np_oracle.py, oracle/port.py) against fixtures produced by
RUNNING the unmodified reference (oracle/gen_golden.py -> tests/golden/*.npz).  No GPU."""
import os

import numpy as np
import pytest
import torch

from oracle import np_oracle as O
from oracle import port

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


def load(name):
    return np.load(os.path.join(G, name), allow_pickle=False)


def rel(got, want):
    return float(np.abs(np.asarray(got, np.float64) - want).max() / max(np.abs(want).max(), 1e-30))


@pytest.mark.parametrize('case', 'abcd')
def test_gae_and_returns(case):
    d = load('gae.npz')
    g, l = d[case + '_cfg']
    b, r = d[case + '_baselines'], d[case + '_rewards']
    adv = O.compute_advantages(g, l, b.shape[1], b, r)
    assert np.allclose(adv, d[case + '_adv'], rtol=1e-5, atol=1e-5)
    ret = O.discount_cumsum(r.astype(np.float64), g)
    assert np.allclose(ret, d[case + '_ret'], rtol=1e-12, atol=1e-12)
    # the time-major restatement gives the same numbers (each row = one episode ending at T-1)
    st = np.full(b.T.shape, O.MID, np.uint8)
    st[-1] = O.TIMEOUT
    a2, r2 = O.gae_time_major(r.T.copy(), b.T.copy(), st, None, g, l)
    assert np.allclose(a2.T, d[case + '_adv'], rtol=2e-5, atol=2e-5)
    assert np.allclose(r2.T, d[case + '_ret'], rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize('case', 'abc')
def test_point_env_transitions(case):
    d = load('point_env.npz')
    gx, gy, arena, bonus, never, max_len = d[case + '_params']
    acts, obs = d[case + '_actions'], d[case + '_obs']
    point, o0 = O.point_reset(1, (gx, gy))
    assert np.array_equal(o0[0], obs[0])
    for t, a in enumerate(acts):
        point, o, rew, done, succ = O.point_step(point, a[None], goal=(gx, gy), arena_size=arena,
                                                 done_bonus=bonus, never_done=bool(never))
        assert np.allclose(o[0], obs[t + 1], rtol=1e-6, atol=1e-7)
        assert abs(rew[0] - d[case + '_rewards'][t]) <= 1e-6 * max(1, abs(rew[0]))
        assert bool(succ[0]) == bool(d[case + '_success'][t])
        assert O.get_step_type(t + 1, int(max_len), bool(done[0])) == d[case + '_step_types'][t]


@pytest.mark.parametrize('case', 'abcd')
def test_normalized_env(case):
    """oracle NormalizedEnvOracle (+ point_step) vs the reference's normalize(PointEnv) run."""
    d = load('normalized_env.npz')
    gx, gy, arena, bonus, never, max_len = d[case + '_env']
    scale_reward, n_obs, n_rew, es, oa, ra = d[case + '_norm']
    w = O.NormalizedEnvOracle(1, 3, [-0.1, -0.1], [0.1, 0.1], scale_reward=scale_reward, normalize_obs=bool(n_obs),
                              normalize_reward=bool(n_rew), expected_action_scale=es, obs_alpha=oa, reward_alpha=ra)
    acts, obs, first = d[case + '_actions'], d[case + '_obs'], list(d[case + '_first'])
    k = 0  # row of `obs`: one extra row per reset()
    cnt = 0
    for t, a in enumerate(acts):
        if t in first:
            point, o0 = O.point_reset(1, (gx, gy))
            assert np.allclose(w.obs(o0)[0], obs[k], rtol=1e-6, atol=1e-7)
            k += 1
            cnt = 0
        point, o, rew, done, succ = O.point_step(point, w.action(a[None]), goal=(gx, gy), arena_size=arena,
                                                 done_bonus=bonus, never_done=bool(never))
        cnt += 1
        assert np.allclose(w.obs(o)[0], obs[k], rtol=1e-5, atol=1e-6)
        assert abs(w.reward(rew)[0] - d[case + '_rewards'][t]) <= 1e-5 * max(1, abs(d[case + '_rewards'][t]))
        assert O.get_step_type(cnt, int(max_len), bool(done[0])) == d[case + '_step_types'][t]
        k += 1
    assert np.allclose(w.obs_mean[0], d[case + '_obs_mean'], rtol=1e-5, atol=1e-7)
    assert np.allclose(w.obs_var[0], d[case + '_obs_var'], rtol=1e-5, atol=1e-7)
    assert np.allclose([w.reward_mean[0], w.reward_var[0]], d[case + '_reward_stats'], rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize('case', ['p64', 'p32', 'c64'])
def test_losses_gradients_adam(case):
    d = load('losses.npz')
    O_, A, H, n = (int(v) for v in d[case + '_dims'])
    obs, act, adv = d[case + '_obs'], d[case + '_act'], d[case + '_adv']
    ret, olp = d[case + '_ret'], d[case + '_old_logp']
    pf, vfl = d[case + '_policy_flat'], d[case + '_vf_flat']
    for kind in ('ppo', 'vpg'):
        loss, grad = O.policy_loss_and_grad(pf, O_, H, A, obs, act, adv, olp, kind=kind)
        assert abs(loss - d['{}_{}_loss'.format(case, kind)]) <= 1e-5 * max(1, abs(loss))
        want = d['{}_{}_grad'.format(case, kind)].astype(np.float64)
        for k, v in O.unflatten_params(want, O_, H, A).items():
            assert rel(O.unflatten_params(grad, O_, H, A)[k], v) < 1e-5, (kind, k)
    loss, grad = O.vf_loss_and_grad(vfl, O_, H, obs, ret)
    assert abs(loss - d[case + '_vf_loss']) <= 1e-5 * max(1, abs(loss))
    want = d[case + '_vf_grad'].astype(np.float64)
    for k, v in O.unflatten_params(want, O_, H, 1).items():
        assert rel(O.unflatten_params(grad, O_, H, 1)[k], v) < 1e-5, k
    # KL / entropy
    p = O.unflatten_params(pf.astype(np.float64), O_, H, A)
    mean, _, _ = O.mlp_forward(p, obs)
    ls, _ = O.clamped_log_std(p)
    kl = O.kl_normal(d[case + '_old_mean'].astype(np.float64), float(d[case + '_old_log_std']), mean, ls).mean()
    assert abs(kl - d[case + '_kl']) <= 1e-4 * max(1e-3, abs(kl))
    assert abs(O.normal_entropy(ls, A) - d[case + '_entropy']) < 1e-5
    # Adam trajectory (policy and value function, 3 steps each)
    p_, m, v = pf.copy(), np.zeros_like(pf), np.zeros_like(pf)
    q_, mq, vq = vfl.copy(), np.zeros_like(vfl), np.zeros_like(vfl)
    for it, mb in enumerate(d[case + '_adam_idx'], start=1):
        _, g = O.policy_loss_and_grad(p_, O_, H, A, obs[mb], act[mb], adv[mb], olp[mb])
        p_, m, v = O.adam_step(p_, g.astype(np.float32), m, v, it, lr=2.5e-4)
        _, g = O.vf_loss_and_grad(q_, O_, H, obs[mb], ret[mb])
        q_, mq, vq = O.adam_step(q_, g.astype(np.float32), mq, vq, it, lr=2.5e-4)
    assert rel(p_ - pf, (d[case + '_policy_flat_after'] - pf).astype(np.float64)) < 5e-3
    assert rel(q_ - vfl, (d[case + '_vf_flat_after'] - vfl).astype(np.float64)) < 5e-3


@pytest.mark.parametrize('case', ['fixed', 'early'])
def test_port_train_once_matches_reference(case):
    """oracle/port.py (the CPU baseline) reproduces the reference's _train_once on the
    reference's own batch: same minibatch shuffles (np.random.seed) -> same parameters."""
    d = load('train_once.npz')
    max_len = int(d[case + '_cfg'][0])
    pol = port.GaussianMLP(3, 2, (32, 32))
    vf = port.GaussianMLP(3, 1, (32, 32), min_std=None)
    pol.load_flat(d[case + '_policy_flat'])
    vf.load_flat(d[case + '_vf_flat'])
    algo = port.PPOPort(pol, vf, max_len, kind='ppo', epochs=3, minibatch_size=64, gae_lambda=0.95)
    batch = dict(observations=d[case + '_obs'], actions=d[case + '_act'], rewards=d[case + '_rew'],
                 lengths=d[case + '_lengths'])
    np.random.seed(11)
    out = algo.train_once(batch)
    assert np.allclose(pol.flat(), d[case + '_policy_flat_after'], rtol=1e-5, atol=1e-6)
    assert np.allclose(vf.flat(), d[case + '_vf_flat_after'], rtol=1e-5, atol=1e-6)
    logs = dict(zip(d[case + '_log_keys'], d[case + '_log_vals']))
    assert abs(out['average_return'] - logs['Evaluation/AverageReturn']) < 1e-9
    assert abs(out['policy_loss_after'] - logs['GaussianMLPPolicy/LossAfter']) < 1e-6
    assert abs(out['vf_loss_after'] - logs['GaussianMLPValueFunction/LossAfter']) <= 1e-5 * abs(out['vf_loss_after'])
    assert abs(out['kl'] - logs['GaussianMLPPolicy/KL']) < 1e-6


@pytest.mark.parametrize('k', range(6))
def test_gridworld_vecworker_restatement(k):
    """oracle.gridworld_vec_rollout vs the reference's VecWorker + LocalSampler on GridWorldEnv
    with a ScriptedPolicy (fixture: oracle/gen_golden.py::gen_gridworld); integers, bit-exact."""
    d = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'gridworld.npz'))
    c = {name[len('c%d_' % k):]: d[name] for name in d.files if name.startswith('c%d_' % k)}
    got = O.gridworld_vec_rollout(c['cells'], c['table'], int(c['max_len']), int(c['n_envs']),
                                  int(c['num_samples']))
    for key in ('observations', 'actions', 'rewards', 'step_types', 'last_observations', 'lengths'):
        np.testing.assert_array_equal(got[key], c[key], err_msg=key)


# ------------------------------------------------------------------ TRPO --
TRPO_CASES = ['fixed', 'early', 'fixed_reg', 'early_reg']


def _trpo_case(case):
    d = load('trpo.npz')
    hid, reg = int(d[case + '_cfg'][3]), float(d[case + '_cfg'][4])
    flat, obs, act = d[case + '_policy_flat'], d[case + '_obs'], d[case + '_act']
    p = O.unflatten_params(flat.astype(np.float64), 3, hid, 2)
    mean, _, _ = O.mlp_forward(p, obs)
    ls, _ = O.clamped_log_std(p)
    old_logp = O.normal_log_prob(act.astype(np.float64), mean, ls)
    return d, hid, reg, flat, obs, act, mean, ls, old_logp


@pytest.mark.parametrize('case', TRPO_CASES)
def test_trpo_pieces_vs_reference(case):
    """Oracle restatement of the loss gradient (trpo.py:90-117) and of the Hessian-vector product of
    the KL constraint (conjugate_gradient_optimizer.py:18-66) against the reference's autograd
    (double backward) on its own batch: fixture oracle/gen_golden.py::gen_trpo."""
    d, hid, reg, flat, obs, act, mean, ls, old_logp = _trpo_case(case)
    _, g = O.policy_loss_and_grad(flat, 3, hid, 2, obs, act, d[case + '_adv'], old_logp, kind='ppo', clip=np.inf)
    want = d[case + '_loss_grad'].astype(np.float64)
    assert np.abs(g - want).max() <= 1e-5 * np.abs(want).max()
    for v, w in zip(d[case + '_hvp_vec'], d[case + '_hvp_out']):
        got = O.fisher_vector_product(flat, 3, hid, 2, obs, v, reg)
        assert np.abs(got - w).max() <= 1e-5 * np.abs(w).max()


@pytest.mark.parametrize('case', ['fixed_reg', 'early_reg'])
def test_trpo_cg_and_step_vs_reference(case):
    """Where the linear system is well conditioned (hvp_reg_coeff = 1) the reference's float32 CG, step
    size and line search are reproducible: CG solution and the policy after `_train_policy`."""
    d, hid, reg, flat, obs, act, mean, ls, old_logp = _trpo_case(case)
    f_Ax = lambda v: O.fisher_vector_product(flat, 3, hid, 2, obs, v, reg)
    x = O.conjugate_gradient(f_Ax, d[case + '_cg_b'])
    assert np.abs(x - d[case + '_cg_x']).max() <= 1e-5 * np.abs(d[case + '_cg_x']).max()
    new, info = O.trpo_policy_step(flat, 3, hid, 2, obs, act, d[case + '_adv'], old_logp, mean, ls,
                                   hvp_reg_coeff=reg)
    assert not info['rejected']
    assert np.allclose(new, d[case + '_policy_flat_after'], rtol=1e-5, atol=2e-6)


def test_trpo_default_config_is_ill_conditioned():
    """Documents why the default-configuration cases pin pieces only: a float64 replay of the SAME
    algorithm lands far from the reference's float32 step (CG with A ~ 1e-5 I in the null space of a
    500-sample Fisher matrix amplifies rounding noise)."""
    d, hid, reg, flat, obs, act, mean, ls, old_logp = _trpo_case('fixed')
    new, _ = O.trpo_policy_step(flat, 3, hid, 2, obs, act, d['fixed_adv'], old_logp, mean, ls, hvp_reg_coeff=reg)
    step = np.abs(d['fixed_policy_flat_after'] - flat).max()
    assert np.abs(new - d['fixed_policy_flat_after']).max() > 0.1 * step
