"""Small device utilities of the update path: the minibatch shuffle (b2g_permutation) and the
advantage moments / normalisation (vpg.py:371-377)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize('n', [1, 2, 3, 64, 1000, 4097, 819200, 1 << 20, (1 << 20) + 1])
def test_permutation_is_a_bijection(n):
    """The device shuffle replaces np.random.shuffle of BatchDataset (minibatch_dataset.py:35): every row
    id must appear exactly once (integer work: exact)."""
    from garage_b200 import ops
    dev = torch.device('cuda:0')
    p = ops.permutation(n, seed=7, counter=3, device=dev).cpu().numpy()
    assert p.dtype == np.int32 and p.shape == (n,)
    assert np.array_equal(np.sort(p), np.arange(n, dtype=np.int32))


def test_permutation_depends_on_seed_and_counter_and_is_reproducible():
    from garage_b200 import ops
    dev = torch.device('cuda:0')
    n = 50000
    a = ops.permutation(n, seed=1, counter=0, device=dev)
    assert torch.equal(a, ops.permutation(n, seed=1, counter=0, device=dev))
    b = ops.permutation(n, seed=1, counter=1, device=dev)
    c = ops.permutation(n, seed=2, counter=0, device=dev)
    ctr = torch.tensor([5], dtype=torch.int64, device=dev)
    d = ops.permutation(n, seed=1, counter=0, counter_dev=ctr, device=dev)
    assert torch.equal(d, ops.permutation(n, seed=1, counter=5, device=dev))  # device counter adds to the host one
    for x in (b, c, d):
        assert float((x == a).float().mean()) < 0.01


def test_permutation_looks_uniform():
    """Statistics of a shuffle: where an element lands is uniform over seeds, neighbours are scattered,
    and a minibatch (a contiguous slice of the permutation) samples the batch evenly."""
    from garage_b200 import ops
    dev = torch.device('cuda:0')
    n, trials = 4096, 400
    first = np.empty(trials)
    disp = []
    for s in range(trials):
        p = ops.permutation(n, seed=100 + s, device=dev).cpu().numpy().astype(np.int64)
        first[s] = p[0]
        if s < 20:
            disp.append(np.abs(np.diff(p)).mean())
            # a 512-row minibatch: mean row id close to n/2 (std of the mean ~ n/sqrt(12*512) = 52)
            assert abs(p[:512].mean() - (n - 1) / 2) < 5 * 52
    # p[0] ~ U[0, n): mean n/2 +- n/sqrt(12 trials), and all four quarters are hit about equally
    assert abs(first.mean() - n / 2) < 5 * n / np.sqrt(12 * trials)
    q = np.histogram(first, bins=4, range=(0, n))[0]
    assert q.min() > trials / 4 - 5 * np.sqrt(trials * 3 / 16)
    # E|p[i+1] - p[i]| = n/3 for a uniform permutation
    assert abs(np.mean(disp) - n / 3) < 0.03 * n


def test_moments_and_normalize_match_torch():
    from garage_b200 import ops
    dev = torch.device('cuda:0')
    g = torch.Generator(device=dev).manual_seed(0)
    for n in (2, 1000, 819200):
        x = torch.randn(n, device=dev, generator=g) * 3 + 1
        st = ops.moments(x).cpu().numpy()
        xd = x.double()
        assert abs(st[0] - float(xd.mean())) < 1e-9 and abs(st[1] - float(xd.var())) < 1e-8 * max(1, float(xd.var()))
        assert st[2] == float(x.min())
        got = ops.normalize_adv(x, ops.moments(x), True, False)
        want = ((xd - xd.mean()) / (xd.var() + 1e-8)).float()   # variance, not std: vpg.py:371-374
        assert torch.allclose(got, want, rtol=1e-5, atol=1e-6)
        got = ops.normalize_adv(x, ops.moments(x), False, True)
        assert torch.allclose(got, (xd - xd.min()).float(), rtol=1e-5, atol=1e-6)
