"""Pin the numpy oracle for GAE / returns / step types / padding against the
reference's own known-answer tests (no GPU)."""
import numpy as np
import pytest

from oracle import np_oracle as O

ONES = np.ones((6, ))
ZEROS = np.zeros((6, ))
ARRANGE = np.arange(6)
PI_DIGITS = np.array([3, 1, 4, 1, 5, 9])
FIBS = np.array([1, 1, 2, 3, 5, 8])


def kat_advantage(discount, gae_lambda, rewards, baselines):
    """The closed form of tests/garage/torch/test_functions.py:99-108."""
    adv = np.zeros(rewards.shape, np.float64)
    for i in range(rewards.shape[0]):
        acc = 0.
        for j in range(rewards.shape[1]):
            acc = acc * discount * gae_lambda
            acc += rewards[i][-j - 1] - baselines[i][-j - 1]
            acc += discount * baselines[i][-j] if j else 0
            adv[i][-j - 1] = acc
    return adv


KAT_CASES = [(d, n, l, r, b) for d in (1, 0.95) for n in (1, 5) for l in (0, 0.5, 1)
             for (r, b) in ((ONES, ZEROS), (PI_DIGITS, ARRANGE), (ONES, FIBS))]


@pytest.mark.parametrize('discount,num_eps,gae_lambda,rewards_eps,baselines_eps', KAT_CASES)
def test_compute_advantages_kat(discount, num_eps, gae_lambda, rewards_eps, baselines_eps):
    rewards = np.repeat(rewards_eps[None].astype(np.float32), num_eps, 0)
    baselines = np.repeat(baselines_eps[None].astype(np.float32), num_eps, 0)
    expected = kat_advantage(discount, gae_lambda, rewards, baselines)
    got = O.compute_advantages(discount, gae_lambda, 6, baselines, rewards)
    assert np.allclose(expected, got, rtol=1e-5, atol=1e-8)


def test_step_type_table():
    """tests/garage/test_dtypes.py:290-318."""
    assert O.get_step_type(1, 5, False) == O.FIRST
    assert O.get_step_type(2, 5, False) == O.MID
    assert O.get_step_type(2, None, False) == O.MID
    assert O.get_step_type(5, 5, False) == O.TIMEOUT
    assert O.get_step_type(5, 5, True) == O.TIMEOUT
    assert O.get_step_type(1, 5, True) == O.TERMINAL
    with pytest.raises(ValueError):
        O.get_step_type(0, 5, False)


def test_pad_batch_array_golden():
    """tests/garage/np/test_functions.py:81-88 (too-long case)."""
    result = O.pad_batch_array(np.ones(9), [5, 2, 2], 2)
    assert (result == np.asarray([[1., 1., 1., 1., 1.], [1., 1., 0., 0., 0.],
                                  [1., 1., 0., 0., 0.]])).all()


def test_discount_cumsum_closed_form():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((3, 17))
    g = 0.97
    y = O.discount_cumsum(x, g)
    for t in range(17):
        want = sum(g**k * x[:, t + k] for k in range(17 - t))
        assert np.allclose(y[:, t], want, rtol=1e-12)


def test_time_major_equals_padded_rows():
    """Cutting the carry at TERMINAL/TIMEOUT == one zero-padded row per episode."""
    rng = np.random.default_rng(1)
    T, N = 40, 7
    rew = rng.standard_normal((T, N)).astype(np.float32)
    val = rng.standard_normal((T, N)).astype(np.float32)
    st = np.full((T, N), O.MID, np.uint8)
    st[rng.random((T, N)) < 0.1] = O.TERMINAL
    st[T - 1] = O.TIMEOUT
    adv, ret = O.gae_time_major(rew, val, st, None, 0.99, 0.95)
    for n in range(N):
        ends = np.nonzero(st[:, n] >= O.TERMINAL)[0]
        start = 0
        for e in ends:
            L = e - start + 1
            r = np.zeros((1, T), np.float32)
            b = np.zeros((1, T), np.float32)
            r[0, :L] = rew[start:e + 1, n]
            b[0, :L] = val[start:e + 1, n]
            a = O.compute_advantages(0.99, 0.95, T, b, r)[0, :L]
            g = O.discount_cumsum(r.astype(np.float64), 0.99)[0, :L]
            assert np.allclose(adv[start:e + 1, n], a, rtol=2e-5, atol=2e-6)
            assert np.allclose(ret[start:e + 1, n], g, rtol=2e-5, atol=2e-6)
            start = e + 1


@pytest.mark.reference
def test_against_live_reference():
    from oracle import ref_shims
    ref_shims.install()
    import torch
    from garage.np import discount_cumsum
    from garage.torch import compute_advantages
    rng = np.random.default_rng(2)
    b = rng.standard_normal((9, 50)).astype(np.float32)
    r = rng.standard_normal((9, 50)).astype(np.float32)
    want = compute_advantages(0.99, 0.95, 50, torch.from_numpy(b), torch.from_numpy(r)).numpy()
    assert np.allclose(O.compute_advantages(0.99, 0.95, 50, b, r), want, rtol=1e-5, atol=1e-6)
    x = rng.standard_normal(200)
    assert np.array_equal(O.discount_cumsum(x, 0.99), discount_cumsum(x, 0.99))
