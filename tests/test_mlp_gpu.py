"""GPU parity: fused MLP forward / loss / backward / Adam kernels vs the float64
numpy oracle (oracle/np_oracle.py), through the C ABI."""
import math

import numpy as np
import pytest
import torch

from oracle import np_oracle as O

pytestmark = pytest.mark.gpu
RTOL = 1e-5  # north_star: loss and gradients within 1e-5 relative (fp32)


def rel_err(got, want):
    """max |got - want| relative to the largest entry of the reference tensor group."""
    return float(np.abs(got - want).max() / max(np.abs(want).max(), 1e-30))


def make_case(Od, H, A, n, seed, scale=0.5):
    rng = np.random.default_rng(seed)
    P = O.mlp_param_count(Od, H, A)
    flat = (rng.standard_normal(P) * scale / math.sqrt(H)).astype(np.float32)
    flat[0] = rng.uniform(-0.7, 0.3)
    obs = rng.standard_normal((n, Od)).astype(np.float32)
    act = rng.standard_normal((n, A)).astype(np.float32)
    adv = rng.standard_normal(n).astype(np.float32)
    p = O.unflatten_params(flat.astype(np.float64), Od, H, A)
    mean, _, _ = O.mlp_forward(p, obs)
    ls, _ = O.clamped_log_std(p)
    ll = O.normal_log_prob(act.astype(np.float64), mean, ls)
    old_logp = (ll + rng.standard_normal(n) * 0.3).astype(np.float32)
    # The clipped surrogate is discontinuous in its gradient at ratio = 1 +- clip (ppo.py:113-132): a row
    # within float32 rounding of a boundary is clipped or not depending on the last bit of exp(), in the
    # reference as much as here.  Keep every row of the synthetic cases 1e-4 away from both boundaries.
    for _ in range(4):
        ratio = np.exp(ll - old_logp.astype(np.float64))
        near = np.minimum(np.abs(ratio - 0.8), np.abs(ratio - 1.2)) < 1e-4
        if not near.any():
            break
        old_logp[near] += np.float32(0.01)
    ret = rng.standard_normal(n).astype(np.float32) * 3
    return flat, obs, act, adv, old_logp, ret


def per_group(flat, Od, H, A):
    return O.unflatten_params(flat, Od, H, A)


def check_grads(got, want, Od, H, A, tol=RTOL):
    g, w = per_group(got, Od, H, A), per_group(want, Od, H, A)
    for k in w:
        e = rel_err(g[k], w[k])
        assert e < tol, 'gradient group {} rel err {:.3e}'.format(k, e)


SHAPES = [(3, 64, 1), (3, 64, 2), (4, 64, 1), (3, 32, 1), (3, 32, 2), (4, 32, 1)]


@pytest.mark.parametrize('Od,H,A', SHAPES)
@pytest.mark.parametrize('n', [1, 77, 128, 1000])
@pytest.mark.parametrize('kind', ['ppo', 'vpg'])
def test_policy_loss_and_grad(Od, H, A, n, kind):
    from garage_b200 import ops
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, n, seed=n + H + A)
    want_loss, want_g = O.policy_loss_and_grad(flat, Od, H, A, obs, act, adv, old_logp, kind=kind)
    dev = torch.device('cuda:0')
    sh = ops.MlpShape(Od, H, A, dev)
    t = lambda a: torch.from_numpy(a).to(dev)
    nb = ops.mlp_minibatch_grad(sh, t(flat), t(obs), act=t(act), adv=t(adv), old_logp=t(old_logp),
                                loss_kind=ops.LOSS_PPO if kind == 'ppo' else ops.LOSS_VPG,
                                min_log_std=math.log(1e-6))
    grad = torch.empty(sh.P, device=dev)
    stats = torch.empty(8, dtype=torch.float64, device=dev)
    ops.mlp_reduce(sh, nb, 1.0 / n, grad_out=grad, stats_out=stats)
    stats = stats.cpu().numpy()
    assert stats[4] == n
    assert abs(stats[0] / n - want_loss) <= RTOL * max(1.0, abs(want_loss))
    check_grads(grad.cpu().numpy().astype(np.float64), want_g, Od, H, A)


@pytest.mark.parametrize('Od,H', [(3, 64), (4, 64), (3, 32), (4, 32)])
@pytest.mark.parametrize('n', [5, 300, 4096])
def test_vf_loss_and_grad(Od, H, n):
    from garage_b200 import ops
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, 1, n, seed=7 * n + H)
    want_loss, want_g = O.vf_loss_and_grad(flat, Od, H, obs, ret)
    dev = torch.device('cuda:0')
    sh = ops.MlpShape(Od, H, 1, dev)
    t = lambda a: torch.from_numpy(a).to(dev)
    nb = ops.mlp_minibatch_grad(sh, t(flat), t(obs), ret=t(ret), loss_kind=ops.LOSS_VF)
    grad = torch.empty(sh.P, device=dev)
    stats = torch.empty(8, dtype=torch.float64, device=dev)
    ops.mlp_reduce(sh, nb, 1.0 / n, grad_out=grad, stats_out=stats)
    assert abs(stats[0].item() / n - want_loss) <= RTOL * max(1.0, abs(want_loss))
    check_grads(grad.cpu().numpy().astype(np.float64), want_g, Od, H, 1)


def test_minibatch_index_gather_and_entropy_flags():
    from garage_b200 import ops
    Od, H, A, B, mb = 3, 64, 2, 5000, 700
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, B, seed=11)
    rng = np.random.default_rng(0)
    idx = rng.permutation(B)[:mb].astype(np.int32)
    want_loss, want_g = O.policy_loss_and_grad(
        flat, Od, H, A, obs[idx], act[idx], adv[idx], old_logp[idx], kind='ppo',
        entropy_regularized=True, ent_coeff=0.01, use_softplus_entropy=True)
    dev = torch.device('cuda:0')
    sh = ops.MlpShape(Od, H, A, dev)
    t = lambda a: torch.from_numpy(a).to(dev)
    nb = ops.mlp_minibatch_grad(sh, t(flat), t(obs), idx=t(idx), act=t(act), adv=t(adv),
                                old_logp=t(old_logp), loss_kind=ops.LOSS_PPO,
                                flags=ops.FLAG_ENT_REG | ops.FLAG_SOFTPLUS_ENT, ent_coeff=0.01,
                                min_log_std=math.log(1e-6))
    grad = torch.empty(sh.P, device=dev)
    stats = torch.empty(8, dtype=torch.float64, device=dev)
    ops.mlp_reduce(sh, nb, 1.0 / mb, grad_out=grad, stats_out=stats)
    assert abs(stats[0].item() / mb - want_loss) <= RTOL * max(1.0, abs(want_loss))
    check_grads(grad.cpu().numpy().astype(np.float64), want_g, Od, H, A)


def test_forward_means_kl_entropy():
    from garage_b200 import ops
    Od, H, A, n = 3, 64, 1, 3000
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, n, seed=3)
    p = O.unflatten_params(flat.astype(np.float64), Od, H, A)
    mean, _, _ = O.mlp_forward(p, obs)
    ls, _ = O.clamped_log_std(p)
    rng = np.random.default_rng(1)
    old_mean = (mean + rng.standard_normal(mean.shape) * 0.1).astype(np.float32)
    old_ls = ls - 0.05
    dev = torch.device('cuda:0')
    sh = ops.MlpShape(Od, H, A, dev)
    t = lambda a: torch.from_numpy(a).to(dev)
    mean_out = torch.empty(n, A, device=dev)
    nb = ops.mlp_eval(sh, t(flat), t(obs), act=t(act), adv=t(adv), old_logp=t(old_logp),
                      old_mean=t(old_mean), old_log_std=torch.tensor([old_ls], dtype=torch.float32, device=dev), loss_kind=ops.LOSS_PPO,
                      min_log_std=math.log(1e-6), mean_out=mean_out)
    stats = torch.empty(8, dtype=torch.float64, device=dev)
    ops.mlp_reduce(sh, nb, 1.0, stats_out=stats)
    stats = stats.cpu().numpy()
    assert np.allclose(mean_out.cpu().numpy(), mean, rtol=1e-5, atol=1e-5)
    want_kl = O.kl_normal(old_mean.astype(np.float64), old_ls, mean, ls).mean()
    assert abs(stats[1] / n - want_kl) <= 1e-5 * max(1, abs(want_kl))
    assert abs(stats[2] / n - O.normal_entropy(ls, A)) <= 1e-5
    want_loss, _ = O.policy_loss_and_grad(flat, Od, H, A, obs, act, adv, old_logp)
    assert abs(stats[0] / n - want_loss) <= 1e-5 * max(1, abs(want_loss))


def test_adam_steps_match_oracle():
    from garage_b200 import ops
    Od, H, A, n = 3, 64, 1, 512
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, n, seed=5)
    dev = torch.device('cuda:0')
    sh = ops.MlpShape(Od, H, A, dev)
    t = lambda a: torch.from_numpy(a).to(dev)
    params = t(flat.copy())
    m = torch.zeros_like(params)
    v = torch.zeros_like(params)
    step = torch.zeros(1, dtype=torch.int64, device=dev)
    o_p, o_m, o_v = flat.copy(), np.zeros_like(flat), np.zeros_like(flat)
    d_obs, d_act, d_adv, d_olp = t(obs), t(act), t(adv), t(old_logp)
    for it in range(1, 6):
        _, g = O.policy_loss_and_grad(o_p, Od, H, A, obs, act, adv, old_logp)
        o_p, o_m, o_v = O.adam_step(o_p, g.astype(np.float32), o_m, o_v, it, lr=2.5e-4)
        nb = ops.mlp_minibatch_grad(sh, params, d_obs, act=d_act, adv=d_adv, old_logp=d_olp,
                                    loss_kind=ops.LOSS_PPO, min_log_std=math.log(1e-6),
                                    step_counter=step)
        ops.mlp_reduce(sh, nb, 1.0 / n, params=params, adam_m=m, adam_v=v, step_counter=step,
                       lr=2.5e-4)
    assert int(step.item()) == 5
    got = params.cpu().numpy()
    # Adam's first steps move every weight by ~lr regardless of gradient scale:
    # compare the UPDATE (param - initial), not the parameter, to 1e-4 relative
    upd_got, upd_want = got - flat, o_p - flat
    assert rel_err(upd_got, upd_want) < 2e-3
    assert np.allclose(got, o_p, rtol=1e-5, atol=1e-6)


def test_unsupported_shape_fails_loudly():
    from garage_b200 import ops
    with pytest.raises(NotImplementedError):
        ops.MlpShape(9, 64, 1, torch.device('cuda:0'))  # obs_dim > 8 (general kernels: obs 1..8, out 1..4, hidden 32/64/128)


@pytest.mark.parametrize('Od,A,n', [(3, 1, 64), (3, 2, 1000), (4, 1, 20001), (3, 1, 1)])
def test_forward_tensor_core_vs_oracle_and_cuda_core(Od, A, n):
    """Forward-only passes on the tcgen05 tile pipeline (hidden 64): means and value loss against the
    float64 oracle, every statistic against the CUDA-core kernel, with an index list and a ragged tail."""
    from garage_b200 import ops
    H = 64
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, n, seed=11 + n)
    rng = np.random.default_rng(n)
    idx = rng.permutation(n).astype(np.int32)
    p = O.unflatten_params(flat.astype(np.float64), Od, H, A)
    mean, _, _ = O.mlp_forward(p, obs)
    dev = torch.device('cuda:0')
    sh = ops.MlpShape(Od, H, A, dev)
    t = lambda a: torch.from_numpy(a).to(dev)
    old_mean = (mean + 0.1).astype(np.float32)
    old_ls = torch.tensor([float(flat[0]) - 0.05], dtype=torch.float32, device=dev)

    def run():
        out = {}
        mo = torch.empty(n, A, device=dev)
        nb = ops.mlp_eval(sh, t(flat), t(obs), idx=t(idx), loss_kind=ops.LOSS_NONE, mean_out=mo)
        out['mean'] = mo.cpu().numpy()
        st = torch.empty(8, dtype=torch.float64, device=dev)
        nb = ops.mlp_eval(sh, t(flat), t(obs), idx=t(idx), act=t(act), adv=t(adv), old_logp=t(old_logp),
                          old_mean=t(old_mean), old_log_std=old_ls, loss_kind=ops.LOSS_PPO, min_log_std=math.log(1e-6))
        ops.mlp_reduce(sh, nb, 1.0, stats_out=st)
        out['ppo'] = st.cpu().numpy().copy()
        if A == 1:
            nb = ops.mlp_eval(sh, t(flat), t(obs), ret=t(ret), loss_kind=ops.LOSS_VF)
            ops.mlp_reduce(sh, nb, 1.0, stats_out=st)
            out['vf'] = st.cpu().numpy().copy()
        return out
    prev = ops.mlp_path()
    try:
        ops.set_mlp_path(0)
        tcr = run()
        ops.set_mlp_path(1)
        ffr = run()
    finally:
        ops.set_mlp_path(prev)
    assert np.allclose(tcr['mean'], mean[idx], rtol=1e-5, atol=1e-5)
    for k in tcr:
        if k == 'mean':
            continue
        assert tcr[k][4] == n == ffr[k][4]  # rows counted once
        assert np.allclose(tcr[k][:5], ffr[k][:5], rtol=1e-5, atol=1e-5 * n), (k, tcr[k], ffr[k])
    if A == 1:
        want = O.vf_loss_and_grad(flat, Od, H, obs, ret)[0]
        assert abs(tcr['vf'][0] / n - want) <= 1e-5 * max(1, abs(want))
