"""GPU parity of the fused training loop (b2g_mlp_train_loop: every minibatch step of
`_train`, both networks, one cooperative launch) against the step-wise path
(b2g_mlp_minibatch_grad + b2g_mlp_reduce per minibatch), which the other GPU tests pin to the
oracle and to the reference's golden fixtures.  Same data, same shuffles, same Adam state:
the parameters after all steps must agree within 1e-5 relative (fp32)."""
import numpy as np
import pytest
import torch

from tests.test_e2e_gpu import build

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _run(env_cls, algo, mode, n_envs, epochs, mb, seed_np=None, hidden=(64, 64), **kw):
    env = env_cls(max_episode_length=40)
    a, policy, vf, sampler = build(env, hidden=hidden, n_envs=n_envs, algo=algo, epochs=epochs, mb=mb,
                                   update_mode=mode, **kw)
    if seed_np is not None:
        np.random.seed(seed_np)
    for itr in range(2):
        eps = sampler.obtain_samples(itr, n_envs * 40, policy)
        a._train_once(itr, eps)
    torch.cuda.synchronize()
    steps = [int(o.state()[2].item()) for o in (a._policy_optimizer, a._vf_optimizer)]
    return policy.flat.clone(), vf.flat.clone(), steps, dict(a.last_stats)


@pytest.mark.parametrize('algo', ['ppo', 'vpg'])
@pytest.mark.parametrize('mb', [None, 512, 1000])
def test_fused_matches_stepwise_device_shuffle(algo, mb):
    from garage_b200.envs import PendulumEnv
    pf, vf_, sf, lf = _run(PendulumEnv, algo, 'fused', 64, 3, mb)
    ps, vs, ss, ls = _run(PendulumEnv, algo, 'stepwise', 64, 3, mb)
    assert sf == ss and sf[0] == 2 * 3 * (1 if mb is None else -(-2560 // mb))
    assert _rel(pf, ps) < 1e-5 and _rel(vf_, vs) < 1e-5
    for k in ('GaussianMLPPolicy/LossAfter', 'GaussianMLPPolicy/KL', 'GaussianMLPValueFunction/LossAfter'):
        assert abs(lf[k] - ls[k]) <= 1e-5 * max(1.0, abs(ls[k])) + 1e-7, k


def test_fused_matches_stepwise_numpy_shuffle_cartpole():
    """shuffle='numpy' (the reference's BatchDataset order) and ragged episodes (CartPole
    terminates early), obs_dim 4."""
    from garage_b200.envs import CartPoleEnv
    pf, vf_, sf, _ = _run(CartPoleEnv, 'ppo', 'fused', 128, 2, 700, seed_np=5, shuffle='numpy')
    ps, vs, ss, _ = _run(CartPoleEnv, 'ppo', 'stepwise', 128, 2, 700, seed_np=5, shuffle='numpy')
    assert sf == ss
    assert _rel(pf, ps) < 1e-5 and _rel(vf_, vs) < 1e-5


@pytest.mark.parametrize('kw', [
    dict(entropy_method='regularized', policy_ent_coeff=0.02, center_adv=True),
    dict(entropy_method='regularized', policy_ent_coeff=0.02, use_softplus_entropy=True, stop_entropy_gradient=True),
    dict(entropy_method='max', policy_ent_coeff=0.01, center_adv=False, stop_entropy_gradient=True),
    dict(positive_adv=True, center_adv=False),
])
def test_fused_matches_stepwise_entropy_and_advantage_options(kw):
    """The entropy configurations of vpg.py:109-125 / 324-347 / 408-432 and positive_adv (vpg.py:375-377)
    flow through the fused loop's arguments exactly as through the step-wise path."""
    from garage_b200.envs import PendulumEnv
    pf, vf_, sf, lf = _run(PendulumEnv, 'ppo', 'fused', 64, 2, 640, **kw)
    ps, vs, ss, ls = _run(PendulumEnv, 'ppo', 'stepwise', 64, 2, 640, **kw)
    assert sf == ss
    assert _rel(pf, ps) < 1e-5 and _rel(vf_, vs) < 1e-5
    assert abs(lf['GaussianMLPPolicy/Entropy'] - ls['GaussianMLPPolicy/Entropy']) < 1e-5


def test_fused_two_action_dims_pointenv():
    from garage_b200.envs import PointEnv
    pf, vf_, sf, _ = _run(PointEnv, 'ppo', 'fused', 64, 2, 640)
    ps, vs, ss, _ = _run(PointEnv, 'ppo', 'stepwise', 64, 2, 640)
    assert sf == ss
    assert _rel(pf, ps) < 1e-5 and _rel(vf_, vs) < 1e-5


def test_fused_refuses_mixed_hidden_sizes():
    """No persistent loop is built for networks of different hidden sizes: 'fused' fails loudly, 'auto'
    takes the step-wise path."""
    from garage_b200.envs import PendulumEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.algos import PPO
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    env = PendulumEnv(max_episode_length=40)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
    vf = GaussianMLPValueFunction(env.spec, hidden_sizes=(32, 32))
    sampler = GPUSampler(policy, env, max_episode_length=40, n_envs=16, seed=3)
    a = PPO(env.spec, policy, vf, sampler, update_mode='fused')
    eps = sampler.obtain_samples(0, 16 * 40, policy)
    with pytest.raises(NotImplementedError):
        a._train_once(0, eps)
    b = PPO(env.spec, policy, vf, sampler)
    assert not b._fused_applies()
    assert np.isfinite(b._train_once(0, eps))


@pytest.mark.parametrize('hidden', [(32, 32), (128, 128)])
@pytest.mark.parametrize('env_name,mb', [('pendulum', 64), ('pendulum', 1000), ('point', 64), ('cartpole', 700), ('pendulum', None)])
def test_general_fused_loop_matches_stepwise(hidden, env_name, mb):
    """The general-shape persistent loop (train_loop_general_*.cu: hidden 32 / 128, one CTA per network
    for 64-row minibatches, several with device-wide barriers for larger ones) against the step-wise
    kernels on the same data, shuffles and Adam state."""
    from garage_b200.envs import CartPoleEnv, PendulumEnv, PointEnv
    cls = dict(pendulum=PendulumEnv, point=PointEnv, cartpole=CartPoleEnv)[env_name]
    kw = dict(seed_np=5, shuffle='numpy') if env_name == 'cartpole' else {}
    pf, vf_, sf, lf = _run(cls, 'ppo', 'fused', 64, 2, mb, hidden=hidden, **kw)
    ps, vs, ss, ls = _run(cls, 'ppo', 'stepwise', 64, 2, mb, hidden=hidden, **kw)
    assert sf == ss and sf[0] > 0
    assert _rel(pf, ps) < 1e-5 and _rel(vf_, vs) < 1e-5
    for k in ('GaussianMLPPolicy/LossAfter', 'GaussianMLPPolicy/KL', 'GaussianMLPValueFunction/LossAfter'):
        assert abs(lf[k] - ls[k]) <= 1e-5 * max(1.0, abs(ls[k])) + 1e-7, k


@pytest.mark.parametrize('Od,H,A', [(6, 128, 3), (8, 32, 4), (5, 64, 2), (3, 32, 2)])
@pytest.mark.parametrize('mb', [64, 300, None])
def test_general_fused_loop_vs_oracle_adam_sequence(Od, H, A, mb):
    """Policy (PPO) + value function through the general persistent loop on supplied index lists vs the
    float64 oracle replaying forward / backward / Adam on the same rows (2 epochs)."""
    from garage_b200 import ops
    from oracle import np_oracle as O
    from tests.test_mlp_gpu import make_case
    import math
    B, E = 777, 2
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, B, seed=Od + H + A)
    vflat = make_case(Od, H, 1, 8, seed=H)[0]
    dev = torch.device('cuda:0')
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    rng = np.random.default_rng(1)
    ids = np.stack([rng.permutation(B) for _ in range(E)]).astype(np.int32)

    def net(sh, f0, policy):
        d = dict(shape=sh, params=t(f0.copy()), adam_m=torch.zeros(sh.P, device=dev), adam_v=torch.zeros(sh.P, device=dev),
                 step_counter=torch.zeros(1, dtype=torch.int64, device=dev), lr=2.5e-4, beta1=0.9, beta2=0.999, eps=1e-8,
                 obs=t(obs), minibatch=mb, epochs=E)
        if policy:
            d.update(act=t(act), adv=t(adv), old_logp=t(old_logp), loss_kind=ops.LOSS_PPO, min_log_std=math.log(1e-6))
        else:
            d.update(ret=t(ret), loss_kind=ops.LOSS_VF)
        if mb is not None:
            d.update(shuffle=ops.SHUFFLE_IDS, idx=t(ids))
        return d
    prev = ops.mlp_path()
    try:
        ops.set_mlp_path(3)  # (5, 64, 2) etc. must not fall to the tcgen05 loop
        pol, vf = net(ops.MlpShape(Od, H, A, dev), flat, True), net(ops.MlpShape(Od, H, 1, dev), vflat, False)
        assert ops.train_loop_kind([pol['shape'], vf['shape']]) == 2
        ops.mlp_train_loop([pol, vf])
        torch.cuda.synchronize()
    finally:
        ops.set_mlp_path(prev)
    f32 = np.float32
    op, om, ov = flat.copy(), np.zeros_like(flat), np.zeros_like(flat)
    vp, vm, vv = vflat.copy(), np.zeros_like(vflat), np.zeros_like(vflat)
    step = 0
    size = mb or B
    for e in range(E):
        for s0 in range(0, B, size):
            rows = ids[e, s0:s0 + size] if mb is not None else np.arange(B)
            step += 1
            _, g = O.policy_loss_and_grad(op, Od, H, A, obs[rows], act[rows], adv[rows], old_logp[rows])
            op, om, ov = O.adam_step(op, g.astype(f32), om, ov, step)
            _, g = O.vf_loss_and_grad(vp, Od, H, obs[rows], ret[rows])
            vp, vm, vv = O.adam_step(vp, g.astype(f32), vm, vv, step)
    assert int(pol['step_counter'].item()) == step == int(vf['step_counter'].item())
    for netd, wp in ((pol, op), (vf, vp)):
        got = netd['params'].cpu().numpy()
        close = np.isclose(got, wp, rtol=1e-5, atol=2e-6)
        assert close.mean() > 0.995 and np.abs(got - wp).max() <= 2 * 2.5e-4 * step


def test_fused_matches_oracle_adam_sequence():
    """Three full-batch value-function steps through the fused loop vs the float64 oracle
    (forward, Gaussian NLL, backward, Adam) on the same data."""
    from garage_b200 import ops
    from oracle import np_oracle as O
    rng = np.random.default_rng(3)
    dev = torch.device('cuda:0')
    Od, H, A, n = 3, 64, 1, 777
    shape = ops.MlpShape(Od, H, A, dev)
    flat = (rng.standard_normal(shape.P) * 0.5 / np.sqrt(H)).astype(np.float32)
    obs = rng.standard_normal((n, Od)).astype(np.float32)
    ret = (rng.standard_normal(n) * 3).astype(np.float32)
    t = lambda x: torch.from_numpy(x).to(dev)
    params, m, v = t(flat.copy()), torch.zeros(shape.P, device=dev), torch.zeros(shape.P, device=dev)
    step = torch.zeros(1, dtype=torch.int64, device=dev)
    ops.mlp_train_loop([dict(shape=shape, params=params, adam_m=m, adam_v=v, step_counter=step, lr=2.5e-4,
                             beta1=0.9, beta2=0.999, eps=1e-8, obs=t(obs), ret=t(ret), loss_kind=ops.LOSS_VF,
                             minibatch=None, epochs=3)])
    torch.cuda.synchronize()
    assert int(step.item()) == 3
    o_p, o_m, o_v = flat.copy(), np.zeros_like(flat), np.zeros_like(flat)
    for it in range(1, 4):
        _, g = O.vf_loss_and_grad(o_p, Od, H, obs, ret)
        o_p, o_m, o_v = O.adam_step(o_p, g.astype(np.float32), o_m, o_v, it, lr=2.5e-4)
    assert np.allclose(params.cpu().numpy(), o_p, rtol=1e-5, atol=1e-6)


def test_fused_loop_is_bit_reproducible():
    """Tiles -> CTAs, partial rows -> slices and every summation order are fixed, so repeating the same
    launch on the same inputs must give the same bits.  A race between warps or CTAs (the loop has
    device-wide barriers, named barriers and a software pipeline over tiles) shows up here as a
    run-to-run difference; 819,200 rows x 2 epochs x 8 repeats."""
    from garage_b200 import ops
    dev = torch.device('cuda:0')
    g = torch.Generator(device=dev).manual_seed(0)
    B = 819200
    shape = ops.MlpShape(3, 64, 1, dev)
    obs = torch.randn(B, 3, device=dev, generator=g)
    act = torch.randn(B, 1, device=dev, generator=g)
    adv = torch.randn(B, device=dev, generator=g)
    olp = torch.randn(B, device=dev, generator=g) * 0.1 - 1.0
    ret = torch.randn(B, device=dev, generator=g)
    p0 = torch.randn(shape.P, device=dev, generator=g) * 0.1
    v0 = torch.randn(shape.P, device=dev, generator=g) * 0.1

    def once():
        nets = []
        for policy, init in ((True, p0), (False, v0)):
            d = dict(shape=shape, params=init.clone(), adam_m=torch.zeros(shape.P, device=dev),
                     adam_v=torch.zeros(shape.P, device=dev), step_counter=torch.zeros(1, dtype=torch.int64, device=dev),
                     lr=2.5e-4, beta1=0.9, beta2=0.999, eps=1e-8, obs=obs, minibatch=32768, epochs=2,
                     shuffle=ops.SHUFFLE_DEVICE, shuffle_seed=3)
            if policy:
                d.update(act=act, adv=adv, old_logp=olp, loss_kind=ops.LOSS_PPO)
            else:
                d.update(ret=ret, loss_kind=ops.LOSS_VF)
            nets.append(d)
        ops.mlp_train_loop(nets)
        torch.cuda.synchronize()
        return nets[0]['params'], nets[1]['params']
    ref = once()
    assert not torch.equal(ref[0], p0) and bool(torch.isfinite(ref[0]).all()) and bool(torch.isfinite(ref[1]).all())
    for _ in range(7):
        got = once()
        assert torch.equal(got[0], ref[0]) and torch.equal(got[1], ref[1])
