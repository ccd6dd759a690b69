"""Drop-ins for `garage.torch._functions` on the on-policy hot path.

Same names, argument meaning and return shapes as the reference
(torch/_functions.py), computed by libb2g kernels on the GPU.
"""
import torch

from garage_b200 import ops


def _to_cuda_f32(x, device=None):
    if isinstance(x, torch.Tensor):
        t = x
    else:
        t = torch.as_tensor(x)
    if not t.is_cuda:
        t = t.to(device or torch.device('cuda', torch.cuda.current_device()))
    return t.to(torch.float32).contiguous()


def compute_advantages(discount, gae_lambda, max_episode_length, baselines, rewards):
    """GAE advantages, reference torch/_functions.py:25-85.

    Args and shapes as the reference: baselines, rewards `(N, T)` zero padded,
    `max_episode_length == T`.  Inputs may be CPU tensors / arrays (copied to
    the current CUDA device; the result comes back on the input's device) or
    CUDA tensors (stay on device).
    """
    src_device = baselines.device if isinstance(baselines, torch.Tensor) else torch.device('cpu')
    b = _to_cuda_f32(baselines)
    r = _to_cuda_f32(rewards, b.device)
    if b.dim() != 2 or b.shape != r.shape:
        raise ValueError('baselines and rewards must both have shape (N, T)')
    if int(max_episode_length) != b.shape[1]:
        # the reference's conv2d/reshape only works when they agree
        raise ValueError('max_episode_length ({}) must equal T ({})'.format(
            max_episode_length, b.shape[1]))
    adv = ops.gae_episode_major(b, r, discount=discount, gae_lambda=gae_lambda)
    return adv.to(src_device) if src_device.type != 'cuda' else adv


def filter_valids(tensor, valids):
    """torch/_functions.py:119-132 (views, no copy)."""
    return [tensor[i][:valid] for i, valid in enumerate(valids)]
