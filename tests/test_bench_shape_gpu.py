"""GPU parity of the tcgen05 backward and of the fused training loop IN THE SHAPE THE BENCH RUNS
THEM: minibatches of 37,888 / 75,776 rows (and ragged sizes), where every CTA owns several 64-row
tiles and the software pipeline over tiles, the double-buffered row publishing and the TMEM
accumulation of dW across tiles are all exercised.  The referee is the float64 numpy oracle
(`np_oracle.policy_loss_and_grad` / `vf_loss_and_grad` / `adam_step`, restating
torch/algos/ppo.py:96-132, vpg.py:434-455, value_functions/gaussian_mlp_value_function.py:81-97 and
the autograd graph of torch/modules/multi_headed_mlp_module.py:136-151), plus the CUDA-core fp32
kernel as an independent second implementation.  Bar: 1e-5 relative per gradient group (north star).
"""
import math

import numpy as np
import pytest
import torch

from oracle import np_oracle as O
from tests.test_mlp_gpu import check_grads, make_case, rel_err

pytestmark = pytest.mark.gpu
RTOL = 1e-5
H = 64
# 37,888 = 148 CTAs x 4 tiles (the bench minibatch), 75,776 = 8 tiles per loop CTA (C4), 37,925 = ragged
# last tile (37 valid rows), 20,001 = uneven tile counts per CTA
SIZES = [37888, 75776, 37925, 20001]
OA = [(3, 1), (3, 2), (4, 1)]


def _dev():
    return torch.device('cuda:0')


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(_dev())


def _torch_fp32_grad(flat, Od, A, obs, kind, act=None, adv=None, old_logp=None, ret=None):
    """The reference's OWN arithmetic on this input: torch-CPU float32 autograd of the same loss
    (ppo.py:96-132 / vpg.py:434-455 / gaussian_mlp_value_function.py:81-97).  Its distance from the
    float64 oracle measures how ill-conditioned a gradient group is: with tens of thousands of rows
    and zero-mean advantages the bias gradients are sums with heavy cancellation, and float32
    autograd itself is then off by more than 1e-5 of the group's largest entry."""
    p = O.unflatten_params(flat, Od, H, A)
    t = {k: torch.tensor(v, dtype=torch.float32, requires_grad=True) for k, v in p.items()}
    x = torch.tensor(obs)
    h1 = torch.tanh(x @ t['W1'].T + t['b1'])
    h2 = torch.tanh(h1 @ t['W2'].T + t['b2'])
    mu = h2 @ t['W3'].T + t['b3']
    ls = t['log_std'] if kind == 'vf' else t['log_std'].clamp(min=math.log(1e-6))
    dist = torch.distributions.Independent(torch.distributions.Normal(mu, ls.exp()), 1)
    if kind == 'vf':
        loss = -dist.log_prob(torch.tensor(ret).reshape(-1, 1)).mean()
    else:
        ll = dist.log_prob(torch.tensor(act))
        a = torch.tensor(adv)
        if kind == 'ppo':
            r = (ll - torch.tensor(old_logp)).exp()
            loss = -torch.min(r * a, r.clamp(0.8, 1.2) * a).mean()
        else:
            loss = -(ll * a).mean()
    loss.backward()
    return O.flatten_params({k: v.grad.numpy().astype(np.float64) for k, v in t.items()})


def check_grads_fp32(got, want, Od, A, ref32):
    """Per gradient group: within 1e-5 of the float64 oracle (the north-star bar), widened by the
    error the reference's own float32 arithmetic (`ref32`) makes on the same group -- i.e. the
    kernel must be as close to the exact gradient as the reference is, and within 1e-5 wherever
    float32 allows it."""
    g, w, r = (O.unflatten_params(np.asarray(v, np.float64), Od, H, A) for v in (got, want, ref32))
    for k in w:
        e, e32 = rel_err(g[k], w[k]), rel_err(r[k], w[k])
        assert e < RTOL + e32, 'gradient group {} rel err {:.3e} (torch fp32 itself: {:.3e})'.format(k, e, e32)


def _grad(sh, flat, obs, n, path, **kw):
    """loss sum, row count and mean gradient of one launch on code path `path` (0 tcgen05, 1 FFMA)."""
    from garage_b200 import ops
    prev = ops.mlp_path()
    try:
        ops.set_mlp_path(path)
        nb = ops.mlp_minibatch_grad(sh, flat, obs, **kw)
        grad = torch.empty(sh.P, device=_dev())
        stats = torch.empty(8, dtype=torch.float64, device=_dev())
        ops.mlp_reduce(sh, nb, 1.0 / n, grad_out=grad, stats_out=stats)
        torch.cuda.synchronize()
    finally:
        ops.set_mlp_path(prev)
    return stats.cpu().numpy(), grad.cpu().numpy().astype(np.float64), nb


@pytest.mark.parametrize('n', SIZES)
@pytest.mark.parametrize('Od,A', OA)
@pytest.mark.parametrize('kind', ['ppo', 'vpg'])
def test_policy_grad_many_tiles_per_cta(Od, A, n, kind):
    from garage_b200 import ops
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, n, seed=n + 10 * Od + A)
    want_loss, want_g = O.policy_loss_and_grad(flat, Od, H, A, obs, act, adv, old_logp, kind=kind)
    sh = ops.MlpShape(Od, H, A, _dev())
    kw = dict(act=_t(act), adv=_t(adv), old_logp=_t(old_logp), min_log_std=math.log(1e-6),
              loss_kind=ops.LOSS_PPO if kind == 'ppo' else ops.LOSS_VPG)
    st, g, nb = _grad(sh, _t(flat), _t(obs), n, 0, **kw)
    assert nb == min(-(-n // 64), 2 * torch.cuda.get_device_properties(0).multi_processor_count)
    assert -(-n // 64) >= 2 * nb or n == 20001  # at least two tiles per CTA
    assert st[4] == n
    assert abs(st[0] / n - want_loss) <= RTOL * max(1.0, abs(want_loss))
    g32 = _torch_fp32_grad(flat, Od, A, obs, kind, act=act, adv=adv, old_logp=old_logp)
    check_grads_fp32(g, want_g, Od, A, g32)
    # independent second implementation on the same inputs
    st_f, g_f, _ = _grad(sh, _t(flat), _t(obs), n, 1, **kw)
    assert abs(st[0] - st_f[0]) <= RTOL * max(n, abs(st_f[0]))
    check_grads_fp32(g_f, want_g, Od, A, g32)


@pytest.mark.parametrize('n', SIZES)
@pytest.mark.parametrize('Od', [3, 4])
def test_vf_grad_many_tiles_per_cta(Od, n):
    from garage_b200 import ops
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, 1, n, seed=3 * n + Od)
    want_loss, want_g = O.vf_loss_and_grad(flat, Od, H, obs, ret)
    sh = ops.MlpShape(Od, H, 1, _dev())
    kw = dict(ret=_t(ret), loss_kind=ops.LOSS_VF)
    st, g, _ = _grad(sh, _t(flat), _t(obs), n, 0, **kw)
    assert st[4] == n
    assert abs(st[0] / n - want_loss) <= RTOL * max(1.0, abs(want_loss))
    g32 = _torch_fp32_grad(flat, Od, 1, obs, 'vf', ret=ret)
    check_grads_fp32(g, want_g, Od, 1, g32)
    st_f, g_f, _ = _grad(sh, _t(flat), _t(obs), n, 1, **kw)
    check_grads_fp32(g_f, want_g, Od, 1, g32)


def test_policy_grad_index_list_many_tiles():
    """Supplied index list (the reference's BatchDataset ids) larger than one tile per CTA."""
    from garage_b200 import ops
    Od, A, B, mb = 3, 1, 100000, 50000
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, B, seed=99)
    idx = np.random.default_rng(5).permutation(B)[:mb].astype(np.int32)
    want_loss, want_g = O.policy_loss_and_grad(flat, Od, H, A, obs[idx], act[idx], adv[idx], old_logp[idx])
    sh = ops.MlpShape(Od, H, A, _dev())
    st, g, _ = _grad(sh, _t(flat), _t(obs), mb, 0, idx=_t(idx), act=_t(act), adv=_t(adv),
                     old_logp=_t(old_logp), loss_kind=ops.LOSS_PPO, min_log_std=math.log(1e-6))
    assert st[4] == mb
    assert abs(st[0] / mb - want_loss) <= RTOL * max(1.0, abs(want_loss))
    g32 = _torch_fp32_grad(flat, Od, A, obs[idx], 'ppo', act=act[idx], adv=adv[idx], old_logp=old_logp[idx])
    check_grads_fp32(g, want_g, Od, A, g32)


def _params_close(got, want, grad, ref32, Od, A, lr=2.5e-4):
    """Parameters after the FIRST Adam step: p - lr * g / (|g| + eps).  The step is +-lr whatever
    |g| is, so an entry whose gradient is ~0 amplifies a gradient error dg into lr * dg / |g|; the
    bound is 1e-5 relative + 1e-6 absolute + that sensitivity for the dg `check_grads_fp32` allows
    (1e-5 + the reference's own float32 error, times the group's largest gradient), capped at the
    largest possible difference 2 lr."""
    w, r = (O.unflatten_params(np.asarray(v, np.float64), Od, H, A) for v in (grad, ref32))
    dg = np.concatenate([np.full(w[k].size, (RTOL + rel_err(r[k], w[k])) * np.abs(w[k]).max()) for k in w])
    tol = 1e-5 * np.abs(want) + 1e-6 + lr * np.minimum(2.0, dg / (np.abs(grad) + 1e-12))
    bad = np.abs(got.astype(np.float64) - want) > tol
    assert not bad.any(), 'parameters differ at {} entries, worst {:.3e}'.format(
        int(bad.sum()), float(np.abs(got - want).max()))


def _loop_net(sh, flat, obs, policy, data, mb, epochs, idx=None):
    from garage_b200 import ops
    d = dict(shape=sh, params=_t(flat.copy()), adam_m=torch.zeros(sh.P, device=_dev()),
             adam_v=torch.zeros(sh.P, device=_dev()),
             step_counter=torch.zeros(1, dtype=torch.int64, device=_dev()), lr=2.5e-4, beta1=0.9,
             beta2=0.999, eps=1e-8, obs=_t(obs), minibatch=mb, epochs=epochs)
    if policy:
        d.update(act=_t(data['act']), adv=_t(data['adv']), old_logp=_t(data['old_logp']),
                 loss_kind=data['loss_kind'], min_log_std=math.log(1e-6))
    else:
        d.update(ret=_t(data['ret']), loss_kind=ops.LOSS_VF)
    if idx is not None:
        d.update(shuffle=ops.SHUFFLE_IDS, idx=_t(idx))
    return d


@pytest.mark.parametrize('n', [37888, 75776, 37925])
@pytest.mark.parametrize('Od,A', OA)
@pytest.mark.parametrize('kind', ['ppo', 'vpg'])
def test_fused_loop_one_full_batch_step_vs_oracle(Od, A, n, kind):
    """One optimiser step of both networks over n rows (4 / 8 tiles per loop CTA, ragged tail).
    After the first Adam step exp_avg = (1 - beta1) * grad, so the moments referee the GRADIENT the
    loop kernel computed (the parameter itself moves by ~lr whatever the gradient's size)."""
    from garage_b200 import ops
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, n, seed=n + Od + A)
    vflat = make_case(Od, H, 1, 8, seed=n + 1)[0]
    lk = ops.LOSS_PPO if kind == 'ppo' else ops.LOSS_VPG
    pol = _loop_net(ops.MlpShape(Od, H, A, _dev()), flat, obs, True,
                    dict(act=act, adv=adv, old_logp=old_logp, loss_kind=lk), None, 1)
    vf = _loop_net(ops.MlpShape(Od, H, 1, _dev()), vflat, obs, False, dict(ret=ret), None, 1)
    ops.mlp_train_loop([pol, vf])
    torch.cuda.synchronize()
    _, gp = O.policy_loss_and_grad(flat, Od, H, A, obs, act, adv, old_logp, kind=kind)
    _, gv = O.vf_loss_and_grad(vflat, Od, H, obs, ret)
    g32s = (_torch_fp32_grad(flat, Od, A, obs, kind, act=act, adv=adv, old_logp=old_logp),
            _torch_fp32_grad(vflat, Od, 1, obs, 'vf', ret=ret))
    for net, f0, want, Aa in ((pol, flat, gp, A), (vf, vflat, gv, 1)):
        assert int(net['step_counter'].item()) == 1
        got_g = net['adam_m'].cpu().numpy().astype(np.float64) / (1.0 - 0.9)
        check_grads_fp32(got_g, want, Od, Aa, g32s[0 if net is pol else 1])
        p1, _, _ = O.adam_step(f0, want.astype(np.float32), np.zeros_like(f0), np.zeros_like(f0), 1)
        _params_close(net['params'].cpu().numpy(), p1, want, g32s[0 if net is pol else 1], Od, Aa)


@pytest.mark.parametrize('Od,A', [(3, 1), (4, 1)])
def test_fused_loop_minibatches_vs_oracle_sequence(Od, A):
    """B = 2 x 37,888 + 37 rows in minibatches of 37,888 (supplied ids, 2 epochs): 6 optimiser steps,
    the last of every epoch a 37-row remainder.  The oracle replays forward/backward/Adam in float64
    arithmetic with float32 state on the same row ids; Adam moments and parameters must agree."""
    from garage_b200 import ops
    B, mb, E = 2 * 37888 + 37, 37888, 2
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, B, seed=17 + Od)
    vflat = make_case(Od, H, 1, 8, seed=23)[0]
    rng = np.random.default_rng(3)
    ids = np.stack([rng.permutation(B) for _ in range(E)]).astype(np.int32)
    pol = _loop_net(ops.MlpShape(Od, H, A, _dev()), flat, obs, True,
                    dict(act=act, adv=adv, old_logp=old_logp, loss_kind=ops.LOSS_PPO), mb, E, idx=ids)
    vf = _loop_net(ops.MlpShape(Od, H, 1, _dev()), vflat, obs, False, dict(ret=ret), mb, E, idx=ids)
    ops.mlp_train_loop([pol, vf])
    torch.cuda.synchronize()
    f32 = np.float32
    op, om, ov = flat.copy(), np.zeros_like(flat), np.zeros_like(flat)
    vp, vm, vv = vflat.copy(), np.zeros_like(vflat), np.zeros_like(vflat)
    step = 0
    for e in range(E):
        for s in range(0, B, mb):
            rows = ids[e, s:s + mb]
            step += 1
            _, g = O.policy_loss_and_grad(op, Od, H, A, obs[rows], act[rows], adv[rows], old_logp[rows])
            op, om, ov = O.adam_step(op, g.astype(f32), om, ov, step)
            _, g = O.vf_loss_and_grad(vp, Od, H, obs[rows], ret[rows])
            vp, vm, vv = O.adam_step(vp, g.astype(f32), vm, vv, step)
    assert int(pol['step_counter'].item()) == step == 6
    for net, wp, wm, wv, f0, Aa in ((pol, op, om, ov, flat, A), (vf, vp, vm, vv, vflat, 1)):
        check_grads(net['adam_m'].cpu().numpy().astype(np.float64), wm.astype(np.float64), Od, H, Aa, tol=1e-4)
        got = net['params'].cpu().numpy()
        close = np.isclose(got, wp, rtol=1e-5, atol=2e-6)
        # entries whose gradient crosses zero in one of the 37-row remainder steps move by +-lr on a
        # sign change: allow a handful of them, bounded by the largest possible drift
        assert close.mean() > 0.995, 'only {:.4f} of the parameters within 1e-5'.format(close.mean())
        assert np.abs(got - wp).max() <= 2 * 2.5e-4 * step
        assert rel_err(got - f0, wp - f0) < 5e-2
