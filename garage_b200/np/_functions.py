"""Drop-ins for `garage.np._functions` on the on-policy hot path."""
import numpy as np
import torch

from garage_b200 import ops


def discount_cumsum(x, discount):
    """Discounted cumulative sum along the last axis, reference
    np/_functions.py:111-128 (scipy.signal.lfilter on the reversed signal).

    numpy in -> numpy float64 out (host buffers: H2D, kernel, D2H);
    CUDA tensor in -> CUDA float64 tensor out.
    """
    if isinstance(x, torch.Tensor) and x.is_cuda:
        return ops.discount_cumsum_f64(x.to(torch.float64).contiguous(), discount)
    arr = np.asarray(x, dtype=np.float64)
    if arr.size == 0:
        return arr.copy()
    dev = torch.device('cuda', torch.cuda.current_device())
    t = torch.from_numpy(np.ascontiguousarray(arr)).to(dev)
    if t.dim() == 1:
        t = t.unsqueeze(0)
    out = ops.discount_cumsum_f64(t, discount)
    return out.cpu().numpy().reshape(arr.shape)
