"""`b2g_moments_static`: the allocation-free form of `b2g_moments` (scratch in a file-scope device
array, include/b2g.h).  Not part of the reference's path: added
for callers that cannot call cudaMalloc.  Called through the C-ABI (ctypes) and compared
with the existing entry point and with float64 numpy, twice in a row: the second call finds the
first call's contents in the shared scratch.  Collected last on purpose (file name)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(x):
    from garage_b200 import _lib
    lib = _lib.load()
    dev = x.device
    stream = torch.cuda.current_stream(dev).cuda_stream
    want = torch.full((3, ), float('nan'), dtype=torch.float64, device=dev)
    got = torch.full((3, ), float('nan'), dtype=torch.float64, device=dev)
    ws = torch.empty(int(lib.b2g_moments_ws_bytes()), dtype=torch.uint8, device=dev)
    assert lib.b2g_moments(x.data_ptr(), x.numel(), want.data_ptr(), ws.data_ptr(), stream) == 0
    rc = lib.b2g_moments_static(x.data_ptr(), x.numel(), got.data_ptr(), stream)
    assert rc == 0, _lib.load().b2g_last_error()
    torch.cuda.synchronize()
    return got.cpu().numpy(), want.cpu().numpy()


def test_moments_static_equals_moments_twice_in_a_row():
    dev = torch.device('cuda:0')
    g = torch.Generator(device='cpu').manual_seed(7)
    # a large vector first, then a small one: the second call must not see the first one's partials
    for n, shift in ((819200, 3.0), (1000, -40.0), (1, 2.5)):
        x = (torch.randn(n, generator=g) * 2.0 + shift).to(dev)
        got, want = _both(x)
        assert np.allclose(got[[0, 2]], want[[0, 2]], rtol=1e-12, atol=0.0), (n, got, want)
        x64 = x.cpu().numpy().astype(np.float64)
        assert got[0] == pytest.approx(x64.mean(), rel=1e-12)
        assert got[2] == x64.min()
        if n > 1:
            assert got[1] == pytest.approx(want[1], rel=1e-12)
            assert got[1] == pytest.approx(x64.var(ddof=1), rel=1e-10)
        else:
            assert np.isnan(got[1]) and np.isnan(want[1])  # torch.var of one element


def test_moments_static_rejects_sizes_beyond_its_limit():
    from garage_b200 import _lib
    lib = _lib.load()
    dev = torch.device('cuda:0')
    x = torch.zeros(16, device=dev)
    out = torch.zeros(3, dtype=torch.float64, device=dev)
    limit = int(lib.b2g_moments_static_max_n())
    assert limit == 1 << 28
    rc = lib.b2g_moments_static(x.data_ptr(), limit + 1, out.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
    assert rc == -1  # B2G_ERR_INVALID, before any launch
    msg = lib.b2g_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else msg
    assert 'exceeds the built-in limit' in msg
    torch.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0
