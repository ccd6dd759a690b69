"""Thin typed wrappers: torch CUDA tensors in, libb2g C-ABI calls out.

Every function validates device / dtype / contiguity, then passes raw device
pointers and the current CUDA stream.  No arithmetic happens in Python.
"""
import torch

from garage_b200 import _lib
from garage_b200._lib import call, ptr, require_cuda


def _chk(t, dtype, name):
    if t is None:
        return
    if not t.is_cuda:
        raise _lib.B2GError('{} must be a CUDA tensor (no CPU path); got {}'.format(name, t.device))
    if t.dtype != dtype:
        raise TypeError('{} must be {}, got {}'.format(name, dtype, t.dtype))
    if not t.is_contiguous():
        raise ValueError('{} must be contiguous'.format(name))


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


# ------------------------------------------------------------------- GAE --
def gae_time_major(rew, val, step_type=None, last_val=None, *, discount, gae_lambda,
                   want_adv=True, want_ret=True, adv_out=None, ret_out=None):
    """Fused GAE + returns over time-major [T,N] buffers (b2g_gae_time_major)."""
    _chk(rew, torch.float32, 'rew')
    _chk(val, torch.float32, 'val')
    _chk(step_type, torch.uint8, 'step_type')
    _chk(last_val, torch.float32, 'last_val')
    if rew.dim() != 2:
        raise ValueError('rew must be [T, N]')
    T, N = rew.shape
    if want_adv and (val is None or val.shape != rew.shape):
        raise ValueError('val must have the shape of rew')
    if step_type is not None and step_type.shape != rew.shape:
        raise ValueError('step_type must have the shape of rew')
    if last_val is not None and last_val.shape != (N, ):
        raise ValueError('last_val must be [N]')
    adv = (adv_out if adv_out is not None else torch.empty_like(rew)) if want_adv else None
    ret = (ret_out if ret_out is not None else torch.empty_like(rew)) if want_ret else None
    _chk(adv, torch.float32, 'adv_out')
    _chk(ret, torch.float32, 'ret_out')
    with torch.cuda.device(rew.device):
        call('b2g_gae_time_major', ptr(rew), ptr(val), ptr(step_type), ptr(last_val),
             float(discount), float(gae_lambda), ptr(adv), ptr(ret), int(T), int(N), _stream(rew))
    return adv, ret


def gae_episode_major(baselines, rewards, *, discount, gae_lambda, want_ret=False):
    """GAE (+returns) over garage's [N,P] zero-padded layout (b2g_gae_episode_major)."""
    _chk(baselines, torch.float32, 'baselines')
    _chk(rewards, torch.float32, 'rewards')
    if baselines.dim() != 2 or baselines.shape != rewards.shape:
        raise ValueError('baselines and rewards must both be [N, P]')
    N, P = rewards.shape
    adv = torch.empty_like(rewards)
    ret = torch.empty_like(rewards) if want_ret else None
    with torch.cuda.device(rewards.device):
        call('b2g_gae_episode_major', ptr(baselines), ptr(rewards), float(discount),
             float(gae_lambda), ptr(adv), ptr(ret), int(N), int(P), _stream(rewards))
    return (adv, ret) if want_ret else adv


def discount_cumsum_f64(x, discount):
    _chk(x, torch.float64, 'x')
    x2 = x.reshape(-1, x.shape[-1]) if x.dim() != 2 else x
    out = torch.empty_like(x2)
    with torch.cuda.device(x.device):
        call('b2g_discount_cumsum_f64', ptr(x2), float(discount), ptr(out), int(x2.shape[0]),
             int(x2.shape[1]), _stream(x))
    return out.reshape(x.shape)


_moments_ws = {}


def _ws_for(device, key, nbytes):
    k = (device, key)
    ws = _moments_ws.get(k)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        _moments_ws[k] = ws
    return ws


def moments(x, out=None):
    """mean, unbiased variance, min of a float vector as a double[3] device tensor."""
    _chk(x, torch.float32, 'x')
    n = x.numel()
    if n == 0:
        raise ValueError('moments of an empty tensor')
    if out is None:
        out = torch.empty(3, dtype=torch.float64, device=x.device)
    ws = _ws_for(x.device, 'moments', _lib.load().b2g_moments_ws_bytes())
    with torch.cuda.device(x.device):
        call('b2g_moments', ptr(x), int(n), ptr(out), ptr(ws), _stream(x))
    return out


def normalize_adv(x, stats, center_adv=True, positive_adv=False, out=None):
    _chk(x, torch.float32, 'x')
    _chk(stats, torch.float64, 'stats')
    if out is None:
        out = torch.empty_like(x)
    with torch.cuda.device(x.device):
        call('b2g_normalize_adv', ptr(x), ptr(stats), int(bool(center_adv)), int(bool(positive_adv)),
             ptr(out), int(x.numel()), _stream(x))
    return out
