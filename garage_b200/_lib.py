"""ctypes binding of libb2g.so (the C ABI declared in include/b2g.h).

There is NO fallback: if the library is missing, or a call fails, an exception is
raised.  Tensors cross the boundary as raw device pointers (`tensor.data_ptr()`)
plus sizes; PyTorch only owns the memory and the stream.
"""
import ctypes
import os
from ctypes import (POINTER, c_char_p, c_double, c_float, c_int, c_int64, c_uint32,
                    c_uint64, c_void_p)

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libb2g.so')

P = c_void_p  # every device pointer / stream


class TrainNet(ctypes.Structure):
    """`b2g_train_net` of include/b2g.h (one network of the fused training loop)."""
    _fields_ = [('obs_dim', c_int), ('hidden', c_int), ('out_dim', c_int),
                ('params', P), ('adam_m', P), ('adam_v', P), ('step_counter', P),
                ('lr', c_double), ('beta1', c_double), ('beta2', c_double), ('eps', c_double),
                ('obs', P), ('act', P), ('adv', P), ('old_logp', P), ('ret', P),
                ('loss_kind', c_int), ('flags', c_int),
                ('clip', c_float), ('ent_coeff', c_float), ('min_log_std', c_float), ('max_log_std', c_float),
                ('n_rows', c_int64), ('minibatch', c_int64),
                ('epochs', c_int), ('shuffle', c_int),
                ('shuffle_seed', c_uint64), ('idx', P)]


MAX_RANKS = 8


class TrainComm(ctypes.Structure):
    """`b2g_train_comm` of include/b2g.h (peer-mapped exchange buffers of the multi-GPU loop)."""
    _fields_ = [('world', c_int), ('rank', c_int), ('peer_base', P * MAX_RANKS)]


class Norm(ctypes.Structure):
    """`b2g_norm` of include/b2g.h (the normalize(env) wrapper fused into the step)."""
    _fields_ = [('flags', c_int), ('act_low', c_float * 2), ('act_high', c_float * 2),
                ('expected_action_scale', c_float),
                ('scale_reward', c_double), ('obs_alpha', c_double), ('reward_alpha', c_double),
                ('stats', P)]


# name -> (restype, argtypes); mirrors include/b2g.h one to one
SIGNATURES = {
    'b2g_version': (c_int, []),
    'b2g_last_error': (c_char_p, []),
    'b2g_device_sm_count': (c_int, []),
    'b2g_gae_time_major': (c_int, [P, P, P, P, c_float, c_float, P, P, c_int, c_int64, P]),
    'b2g_gae_episode_major': (c_int, [P, P, c_float, c_float, P, P, c_int64, c_int, P]),
    'b2g_discount_cumsum_f64': (c_int, [P, c_double, P, c_int64, c_int, P]),
    'b2g_moments_ws_bytes': (c_int64, []),
    'b2g_moments': (c_int, [P, c_int64, P, P, P]),
    'b2g_moments_static_max_n': (c_int64, []),
    'b2g_moments_static': (c_int, [P, c_int64, P, P]),
    'b2g_gridworld_step': (c_int, [P, c_int, c_int, c_int, P, P, P, P, P, P, P, c_int64, P]),
    'b2g_gridworld_rollout': (c_int, [P, c_int, c_int, c_int, c_int, P, P, P, c_int, c_int64, P, P, P, P, P, P, P]),
    'b2g_episode_stats': (c_int, [P, c_int, P, P, P, P, c_int64, c_double, c_int64, P, P, P]),
    'b2g_normalize_adv': (c_int, [P, P, c_int, c_int, P, c_int64, P]),
    'b2g_permutation': (c_int, [c_uint64, P, c_uint64, c_int64, P, P]),
    'b2g_env_obs_dim': (c_int, [c_int]),
    'b2g_env_act_dim': (c_int, [c_int]),
    'b2g_env_state_dim': (c_int, [c_int]),
    'b2g_env_reset': (c_int, [c_int, POINTER(c_double), c_uint64, c_int64, P, P, P, P, c_int64, P]),
    'b2g_env_step': (c_int, [c_int, POINTER(c_double), c_int, P, P, P, P, P, P, P, P, c_int64, P]),
    'b2g_mlp_param_count': (c_int64, [c_int, c_int, c_int]),
    'b2g_counter_add': (c_int, [P, c_uint64, P]),
    'b2g_rollout': (c_int, [c_int, POINTER(c_double), c_int, P, c_int, c_float, c_float, c_uint64,
                            c_int64, c_uint32, P, P, P, P, P, c_int, c_int64, P, P, P, P, P, P, P, P,
                            P, P]),
    'b2g_env_reset_norm': (c_int, [c_int, POINTER(c_double), POINTER(Norm), c_uint64, c_int64, P, P, P, P, c_int64, P]),
    'b2g_env_step_norm': (c_int, [c_int, POINTER(c_double), POINTER(Norm), c_int, P, P, P, P, P, P, P, P, c_int64, P]),
    'b2g_rollout_norm': (c_int, [c_int, POINTER(c_double), POINTER(Norm), c_int, P, c_int, c_float, c_float,
                                 c_uint64, c_int64, c_uint32, P, P, P, P, P, c_int, c_int64, P, P, P, P, P, P,
                                 P, P, P, P]),
    'b2g_policy_sample': (c_int, [P, c_int, c_float, c_uint64, c_int64, c_uint32, P, P, c_int64, P]),
    'b2g_pack_ws_bytes': (c_int64, [c_int]),
    'b2g_pack_plan': (c_int, [P, P, c_int, c_int64, c_int64, P, P, P]),
    'b2g_pack_index': (c_int, [P, P, P, c_int, c_int, c_int, c_int64, P, P, P, P, P, P, P]),
    'b2g_pack_fields': (c_int, [P, P, P, c_int, c_int64, c_int, POINTER(c_void_p),
                                POINTER(c_void_p), POINTER(c_int), P]),
    'b2g_pack_rowmap': (c_int, [P, P, P, c_int, c_int64, P, P]),
    'b2g_pack_fields_mapped': (c_int, [P, c_int, c_int64, c_int, POINTER(c_void_p), POINTER(c_void_p),
                                       POINTER(c_int), P]),
    'b2g_mlp_num_partials': (c_int, [c_int, c_int, c_int, c_int, c_int64]),
    'b2g_set_mlp_path': (c_int, [c_int]),
    'b2g_get_mlp_path': (c_int, []),
    'b2g_mlp_debug_timing': (c_int, [P]),
    'b2g_mlp_ws_floats': (c_int64, [c_int, c_int, c_int]),
    'b2g_mlp_minibatch_grad': (c_int, [c_int, c_int, c_int, P, P, c_int64, c_int64, P, P, P, P, P,
                                       c_int, c_int, c_float, c_float, c_float, c_float, P, P, P]),
    'b2g_mlp_eval': (c_int, [c_int, c_int, c_int, P, P, c_int64, c_int64, P, P, P, P, P, P,
                             P, c_int, c_int, c_float, c_float, c_float, c_float, P, P, P]),
    'b2g_mlp_reduce': (c_int, [P, c_int, c_int, c_float, P, P, P, P, P, P, c_double, c_double,
                               c_double, c_double, P]),
    'b2g_adam_step': (c_int, [P, P, c_float, P, P, P, c_double, c_double, c_double, c_double,
                              c_int, P]),
    'b2g_mlp_train_loop_ws_floats': (c_int64, []),
    'b2g_mlp_train_loop_debug_timing': (c_int, [P, c_int]),
    'b2g_mlp_train_loop': (c_int, [POINTER(TrainNet), c_int, P, c_int64, P]),
    'b2g_train_comm_bytes': (c_int64, []),
    'b2g_mlp_train_loop_kind': (c_int, [POINTER(TrainNet), c_int, c_int]),
    'b2g_mlp_train_loop_dist': (c_int, [POINTER(TrainNet), c_int, POINTER(TrainComm), P, c_int64, P]),
    'b2g_mlp_fvp': (c_int, [c_int, c_int, c_int, P, P, P, c_int64, c_int64, P, c_float, c_float, c_float, P, P, P, P]),
    'b2g_cg_init': (c_int, [P, P, P, P, P, c_int, P]),
    'b2g_cg_step': (c_int, [P, P, P, P, P, c_int, c_double, P]),
    'b2g_cg_sanitize': (c_int, [P, c_int, P]),
    'b2g_cg_descent': (c_int, [P, P, c_double, P, P, c_int, P]),
    'b2g_line_search_point': (c_int, [P, P, c_float, P, c_int, P]),
    'b2g_peer_reduce_f64': (c_int, [POINTER(TrainComm), c_int, P, c_int, c_double, c_double, P, P]),
    'b2g_comm_alloc': (c_int, [c_int64, POINTER(c_void_p), ctypes.c_char_p]),
    'b2g_comm_open': (c_int, [ctypes.c_char_p, POINTER(c_void_p)]),
    'b2g_comm_close': (c_int, [P]),
    'b2g_comm_free': (c_int, [P]),
}


class B2GError(RuntimeError):
    pass


_lib = None


def load():
    """Load libb2g.so once; raise if it was not built (no CPU path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B2GError(
            'libb2g.so not found at {}: build it with `python -m garage_b200.csrc.build` '
            '(garage_b200 has no CPU fallback)'.format(LIB_PATH))
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(code, what=''):
    if code != 0:
        msg = load().b2g_last_error()
        raise B2GError('{} failed with code {}: {}'.format(
            what or 'libb2g call', code, msg.decode() if msg else ''))


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr(device=None):
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise B2GError('libb2g needs CUDA tensors (there is no CPU path); got device {}'.format(t.device))


def call(name, *args):
    lib = load()
    code = getattr(lib, name)(*args)
    check(code, name)
