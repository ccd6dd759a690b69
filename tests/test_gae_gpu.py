"""GPU parity: libb2g GAE / returns kernels vs the numpy oracle (through the C ABI)."""
import numpy as np
import pytest
import torch

from oracle import np_oracle as O
from tests.test_oracle_gae import KAT_CASES, kat_advantage

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-5, 1e-5  # north_star: 1e-5 relative in fp32 (absolute floor for ~0 values)


def dev():
    return torch.device('cuda:0')


@pytest.mark.parametrize('discount,num_eps,gae_lambda,rewards_eps,baselines_eps', KAT_CASES)
def test_compute_advantages_reference_kat(discount, num_eps, gae_lambda, rewards_eps,
                                          baselines_eps):
    """The reference's own KAT, tests/garage/torch/test_functions.py:86-117, on the kernel."""
    from garage_b200.torch import compute_advantages
    rewards = torch.Tensor(np.repeat(rewards_eps[None], num_eps, 0))
    baselines = torch.Tensor(np.repeat(baselines_eps[None], num_eps, 0))
    expected = torch.Tensor(kat_advantage(discount, gae_lambda, rewards.numpy(), baselines.numpy()))
    got = compute_advantages(discount, gae_lambda, 6, baselines, rewards)
    assert got.device.type == 'cpu' and got.shape == rewards.shape
    assert torch.allclose(expected, got)
    got_dev = compute_advantages(discount, gae_lambda, 6, baselines.cuda(), rewards.cuda())
    assert got_dev.is_cuda and torch.allclose(expected, got_dev.cpu())


@pytest.mark.parametrize('N,P', [(1, 1), (3, 31), (5, 32), (4, 33), (7, 200), (2, 1000), (64, 500)])
def test_episode_major_vs_oracle(N, P):
    from garage_b200 import ops
    rng = np.random.default_rng(P * 1000 + N)
    b = rng.standard_normal((N, P)).astype(np.float32)
    r = rng.standard_normal((N, P)).astype(np.float32)
    adv, ret = ops.gae_episode_major(torch.from_numpy(b).cuda(), torch.from_numpy(r).cuda(),
                                     discount=0.99, gae_lambda=0.95, want_ret=True)
    want_a = O.compute_advantages(0.99, 0.95, P, b, r)
    want_r = O.discount_cumsum(r.astype(np.float64), 0.99)
    assert np.allclose(adv.cpu().numpy(), want_a, rtol=RTOL, atol=ATOL)
    assert np.allclose(ret.cpu().numpy(), want_r, rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize('T,N,p_done', [(1, 1, 0.0), (200, 37, 0.0), (200, 4096, 0.0),
                                        (128, 1000, 0.05), (77, 40000, 0.01), (33, 70000, 0.02),
                                        (9, 140000, 0.1)])
def test_time_major_vs_oracle(T, N, p_done):
    from garage_b200 import ops
    rng = np.random.default_rng(T + N)
    rew = rng.standard_normal((T, N)).astype(np.float32)
    val = rng.standard_normal((T, N)).astype(np.float32)
    st = np.full((T, N), O.MID, np.uint8)
    st[rng.random((T, N)) < p_done] = O.TERMINAL
    st[T - 1] = O.TIMEOUT
    adv, ret = ops.gae_time_major(torch.from_numpy(rew).cuda(), torch.from_numpy(val).cuda(),
                                  torch.from_numpy(st).cuda(), discount=0.99, gae_lambda=0.95)
    wa, wr = O.gae_time_major(rew, val, st, None, 0.99, 0.95)
    assert np.allclose(adv.cpu().numpy(), wa, rtol=RTOL, atol=ATOL)
    assert np.allclose(ret.cpu().numpy(), wr, rtol=RTOL, atol=ATOL)


def test_time_major_variants_and_bootstrap():
    from garage_b200 import ops
    rng = np.random.default_rng(5)
    T, N = 50, 300
    rew = rng.standard_normal((T, N)).astype(np.float32)
    val = rng.standard_normal((T, N)).astype(np.float32)
    lv = rng.standard_normal(N).astype(np.float32)
    r, v, l = (torch.from_numpy(a).cuda() for a in (rew, val, lv))
    wa, wr = O.gae_time_major(rew, val, None, lv, 0.9, 0.8)
    adv, ret = ops.gae_time_major(r, v, None, l, discount=0.9, gae_lambda=0.8)
    assert np.allclose(adv.cpu().numpy(), wa, rtol=RTOL, atol=ATOL)
    assert np.allclose(ret.cpu().numpy(), wr, rtol=RTOL, atol=ATOL)
    adv2, none = ops.gae_time_major(r, v, None, l, discount=0.9, gae_lambda=0.8, want_ret=False)
    assert none is None and torch.equal(adv2, adv)
    none, ret2 = ops.gae_time_major(r, None, None, None, discount=0.9, gae_lambda=0.8, want_adv=False)
    assert none is None and torch.equal(ret2, ret)


def test_full_size_linearity_property():
    """BASELINE config size (T=200, N=65,536): GAE is linear in (rew, val) and a
    constant reward r with V=0 gives the geometric closed form."""
    from garage_b200 import ops
    T, N = 200, 65536
    g = torch.Generator(device='cuda').manual_seed(0)
    r1 = torch.randn(T, N, device='cuda', generator=g)
    v1 = torch.randn(T, N, device='cuda', generator=g)
    r2 = torch.randn(T, N, device='cuda', generator=g)
    v2 = torch.randn(T, N, device='cuda', generator=g)
    a1, q1 = ops.gae_time_major(r1, v1, discount=0.99, gae_lambda=0.95)
    a2, q2 = ops.gae_time_major(r2, v2, discount=0.99, gae_lambda=0.95)
    a3, q3 = ops.gae_time_major(r1 + r2, v1 + v2, discount=0.99, gae_lambda=0.95)
    assert torch.allclose(a1 + a2, a3, rtol=1e-4, atol=1e-4)
    assert torch.allclose(q1 + q2, q3, rtol=1e-4, atol=1e-4)
    ones = torch.ones(T, N, device='cuda')
    a, q = ops.gae_time_major(ones, torch.zeros_like(ones), discount=0.99, gae_lambda=1.0)
    k = torch.arange(T, 0, -1, device='cuda', dtype=torch.float64)
    closed = ((1 - 0.99**k) / (1 - 0.99)).to(torch.float32)[:, None].expand(T, N)
    assert torch.allclose(q, closed, rtol=1e-5) and torch.allclose(a, closed, rtol=1e-5)


def test_discount_cumsum_dropin_host_buffers():
    from garage_b200.np import discount_cumsum
    rng = np.random.default_rng(3)
    x = rng.standard_normal(200)
    got = discount_cumsum(x, 0.99)
    assert got.dtype == np.float64 and got.shape == x.shape
    assert np.allclose(got, O.discount_cumsum(x, 0.99), rtol=1e-12, atol=1e-12)
    assert discount_cumsum(np.zeros(0), 0.9).shape == (0, )


@pytest.mark.parametrize('n', [2, 1000, 819200])
@pytest.mark.parametrize('center,positive', [(True, False), (True, True), (False, True)])
def test_normalize_adv(n, center, positive):
    from garage_b200 import ops
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) * 3 + 1.5).astype(np.float32)
    xd = torch.from_numpy(x).cuda()
    st = ops.moments(xd)
    y = ops.normalize_adv(xd, st, center, positive).cpu().numpy()
    want = O.normalize_advantages(x, center, positive)
    assert np.allclose(y, want, rtol=RTOL, atol=ATOL)
    t = torch.from_numpy(x)
    assert abs(st[0].item() - t.double().mean().item()) < 1e-9
    assert abs(st[1].item() - t.double().var().item()) < 1e-7 * max(1, t.var().item())


def test_no_cpu_fallback():
    from garage_b200 import _lib, ops
    with pytest.raises(_lib.B2GError):
        ops.gae_time_major(torch.zeros(2, 2), torch.zeros(2, 2), discount=0.9, gae_lambda=0.9)
