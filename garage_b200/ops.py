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


def episode_stats(rewards, discount, step_types=None, success=None, lengths=None, fixed_len=None, out=None):
    """b2g_episode_stats over a packed episode-major batch.  Returns (per_episode [4, n_eps] f64:
    undiscounted return, discounted return, terminated, success; summary [7] f64)."""
    if rewards.dtype not in (torch.float32, torch.float64):
        raise TypeError('rewards must be float32 or float64')
    _chk(rewards, rewards.dtype, 'rewards')
    _chk(step_types, torch.uint8, 'step_types')
    _chk(success, torch.uint8, 'success')
    _chk(lengths, torch.int64, 'lengths')
    B = rewards.numel()
    if fixed_len:
        if B % int(fixed_len):
            raise ValueError('fixed_len does not divide the batch')
        n_eps, offsets = B // int(fixed_len), None
    else:
        if lengths is None:
            raise ValueError('need lengths or fixed_len')
        n_eps = lengths.numel()
        offsets = torch.cumsum(lengths, 0) - lengths
    if n_eps == 0:
        raise ValueError('no episodes')
    per = torch.empty(4, n_eps, dtype=torch.float64, device=rewards.device)
    if out is None:
        out = torch.empty(7, dtype=torch.float64, device=rewards.device)
    _chk(out, torch.float64, 'out')
    with torch.cuda.device(rewards.device):
        call('b2g_episode_stats', ptr(rewards), int(rewards.dtype == torch.float64), ptr(step_types),
             ptr(success), ptr(lengths), ptr(offsets), int(fixed_len or 0), float(discount), int(n_eps),
             ptr(per), ptr(out), _stream(rewards))
    return per, out


def normalize_adv(x, stats, center_adv=True, positive_adv=False, out=None):
    _chk(x, torch.float32, 'x')
    _chk(stats, torch.float64, 'stats')
    if out is None:
        out = torch.empty_like(x)
    with torch.cuda.device(x.device):
        call('b2g_normalize_adv', ptr(x), ptr(stats), int(bool(center_adv)), int(bool(positive_adv)),
             ptr(out), int(x.numel()), _stream(x))
    return out


# ------------------------------------------------------------------ envs --
import ctypes  # noqa: E402
import math  # noqa: E402

ENV_POINT, ENV_PENDULUM, ENV_CARTPOLE = 0, 1, 2
ENV_NPARAMS = 10


def env_params_array(values):
    arr = (ctypes.c_double * ENV_NPARAMS)()
    for i, v in enumerate(values):
        arr[i] = float(v)
    return arr


def env_dims(env_id):
    lib = _lib.load()
    return lib.b2g_env_obs_dim(env_id), lib.b2g_env_act_dim(env_id), lib.b2g_env_state_dim(env_id)


NORM_ACTION, NORM_OBS, NORM_REWARD = 1, 2, 4


class NormConfig:
    """Host description of `normalize(env, ...)` (reference envs/normalized_env.py:29-55); the
    per-env running statistics live in `EnvState.norm_stats`."""

    def __init__(self, flags, act_low, act_high, expected_action_scale=1., scale_reward=1.,
                 obs_alpha=0.001, reward_alpha=0.001):
        self.flags = int(flags)
        self.act_low = [float(v) for v in act_low]
        self.act_high = [float(v) for v in act_high]
        self.expected_action_scale = float(expected_action_scale)
        self.scale_reward = float(scale_reward)
        self.obs_alpha, self.reward_alpha = float(obs_alpha), float(reward_alpha)

    @property
    def needs_stats(self):
        return bool(self.flags & (NORM_OBS | NORM_REWARD))

    def struct(self, es):
        """ctypes `b2g_norm` bound to the statistics of EnvState `es` (allocated on first use)."""
        q = _lib.Norm()
        q.flags = self.flags
        for i in range(2):
            q.act_low[i] = self.act_low[i] if i < len(self.act_low) else 0.
            q.act_high[i] = self.act_high[i] if i < len(self.act_high) else 0.
        q.expected_action_scale = self.expected_action_scale
        q.scale_reward, q.obs_alpha, q.reward_alpha = self.scale_reward, self.obs_alpha, self.reward_alpha
        if self.needs_stats:
            if es.norm_stats is None:
                es.norm_stats = new_norm_stats(es.env_id, es.n, es.device)
            q.stats = es.norm_stats.data_ptr()
        else:
            q.stats = None
        return q


def new_norm_stats(env_id, n, device):
    """[2*O + 2, n] float64: obs_mean = 0, obs_var = 1, reward_mean = 0, reward_var = 1
    (normalized_env.py:50-55)."""
    O = env_dims(env_id)[0]
    st = torch.zeros(2 * O + 2, n, dtype=torch.float64, device=device)
    st[O:2 * O] = 1.
    st[2 * O + 1] = 1.
    return st


def _norm_arg(es, norm):
    return None if norm is None else ctypes.byref(norm.struct(es))


class EnvState:
    """Device-resident state of N batched environments (caller-owned tensors)."""

    def __init__(self, env_id, n, device):
        O, A, S = env_dims(env_id)
        self.env_id, self.n, self.device = env_id, n, device
        self.norm_stats = None  # running statistics of normalize(env), see NormConfig
        self.state = torch.zeros(S, n, dtype=torch.float64, device=device)
        self.obs = torch.zeros(n, O, dtype=torch.float32, device=device)
        self.step_cnt = torch.full((n, ), -1, dtype=torch.int32, device=device)
        self.episode = torch.zeros(n, dtype=torch.int32, device=device)


def env_reset(es, params, seed, env_id_offset=0, norm=None):
    with torch.cuda.device(es.device):
        call('b2g_env_reset_norm', es.env_id, params, _norm_arg(es, norm), int(seed) & (2**64 - 1),
             int(env_id_offset), ptr(es.state), ptr(es.obs), ptr(es.step_cnt), ptr(es.episode), es.n,
             _stream(es.state))


def env_step(es, params, max_episode_length, actions, want_success=False, norm=None):
    """One Environment.step() for all envs. Returns (obs, reward f64, step_type u8, success|None)."""
    _chk(actions, torch.float32, 'actions')
    O, A, S = env_dims(es.env_id)
    if actions.shape != (es.n, A):
        raise ValueError('actions must be [{}, {}]'.format(es.n, A))
    dev = es.device
    obs = torch.empty(es.n, O, dtype=torch.float32, device=dev)
    rew = torch.empty(es.n, dtype=torch.float64, device=dev)
    st = torch.empty(es.n, dtype=torch.uint8, device=dev)
    succ = torch.empty(es.n, dtype=torch.uint8, device=dev) if want_success else None
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        call('b2g_env_step_norm', es.env_id, params, _norm_arg(es, norm), int(max_episode_length or 0), ptr(es.state),
             ptr(actions), ptr(obs), ptr(rew), ptr(st), ptr(succ), ptr(es.step_cnt), ptr(err),
             es.n, _stream(actions))
    bad = int(err.item())
    if bad:
        # the reference raises RuntimeError (envs/point_env.py:113-114)
        raise RuntimeError('reset() must be called before step()! (env {})'.format(bad - 1))
    es.obs = obs
    return obs, rew, st, succ


# --------------------------------------------------------------- rollout --
def mlp_param_count(obs_dim, hidden, out_dim):
    return int(_lib.load().b2g_mlp_param_count(obs_dim, hidden, out_dim))


class RolloutBuffers:
    """Time-major [T,N,...] rollout buffers (caller-owned)."""

    def __init__(self, env_id, T, n, device, want_mean=True, want_success=False):
        O, A, S = env_dims(env_id)
        f32 = dict(dtype=torch.float32, device=device)
        self.T, self.n = T, n
        self.obs = torch.empty(T, n, O, **f32)
        self.act = torch.empty(T, n, A, **f32)
        self.rew = torch.empty(T, n, **f32)
        self.logp = torch.empty(T, n, **f32)
        self.step_type = torch.empty(T, n, dtype=torch.uint8, device=device)
        self.mean = torch.empty(T, n, A, **f32) if want_mean else None
        self.success = torch.empty(T, n, dtype=torch.uint8, device=device) if want_success else None
        self.last_obs = torch.empty(T, n, O, **f32)
        self.ep_len = torch.empty(T, n, dtype=torch.int32, device=device)


def counter_add(counter, delta):
    with torch.cuda.device(counter.device):
        call('b2g_counter_add', ptr(counter), int(delta), _stream(counter))


def rollout(es, params, max_episode_length, policy_flat, hidden, min_log_std, max_log_std, seed,
            env_id_offset, t_counter, buf, t0, n_steps, t_counter_dev=None, norm=None):
    """b2g_rollout for rows t0..t0+n_steps of `buf`."""
    _chk(policy_flat, torch.float32, 'policy_flat')
    if t0 + n_steps > buf.T:
        raise ValueError('rollout rows exceed the buffer')

    def row(t):
        return None if t is None else t[t0:]

    with torch.cuda.device(es.device):
        call('b2g_rollout_norm', es.env_id, params, _norm_arg(es, norm), int(max_episode_length or 0), ptr(policy_flat),
             int(hidden), float(min_log_std), float(max_log_std), int(seed) & (2**64 - 1),
             int(env_id_offset), int(t_counter) & 0xffffffff, ptr(t_counter_dev), ptr(es.state), ptr(es.obs),
             ptr(es.step_cnt), ptr(es.episode), int(n_steps), es.n, ptr(row(buf.obs)),
             ptr(row(buf.act)), ptr(row(buf.rew)), ptr(row(buf.logp)), ptr(row(buf.step_type)),
             ptr(row(buf.mean)), ptr(row(buf.success)), ptr(row(buf.last_obs)),
             ptr(row(buf.ep_len)), _stream(policy_flat))


def policy_sample(mean, log_std, seed, env_id_offset, t_counter, want_logp=True):
    _chk(mean, torch.float32, 'mean')
    n, A = mean.shape
    act = torch.empty_like(mean)
    logp = torch.empty(n, dtype=torch.float32, device=mean.device) if want_logp else None
    with torch.cuda.device(mean.device):
        call('b2g_policy_sample', ptr(mean), int(A), float(log_std), int(seed) & (2**64 - 1),
             int(env_id_offset), int(t_counter) & 0xffffffff, ptr(act), ptr(logp), int(n),
             _stream(mean))
    return act, logp


# ------------------------------------------------------------------ pack --
def pack_plan(step_type, ep_len, T, num_samples, sync=True):
    """Row statistics + cut row.  sync=True returns (ws, (t_star, n_eps, n_steps, reached)) after
    one small D2H copy; sync=False returns (ws, summary_device_tensor) without synchronising."""
    _chk(step_type, torch.uint8, 'step_type')
    _chk(ep_len, torch.int32, 'ep_len')
    N = step_type.shape[1]
    dev = step_type.device
    ws = torch.empty(int(_lib.load().b2g_pack_ws_bytes(T)), dtype=torch.uint8, device=dev)
    summary = torch.empty(4, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        call('b2g_pack_plan', ptr(step_type), ptr(ep_len), int(T), int(N), int(num_samples), ptr(ws),
             ptr(summary), _stream(step_type))
    if not sync:
        return ws, summary
    t_star, n_eps, n_steps, reached = (int(v) for v in summary.tolist())
    return ws, (t_star, n_eps, n_steps, bool(reached))


def pack_index(step_type, ep_len, last_obs, ws, T_ws, t_star, n_eps):
    N = step_type.shape[1]
    dev = step_type.device
    O = last_obs.shape[-1]
    lengths = torch.empty(n_eps, dtype=torch.int64, device=dev)
    ep_off = torch.empty(t_star + 1, N, dtype=torch.int64, device=dev)
    last_observations = torch.empty(n_eps, O, dtype=torch.float32, device=dev)
    ep_env = torch.empty(n_eps, dtype=torch.int32, device=dev)
    ep_end = torch.empty(n_eps, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        call('b2g_pack_index', ptr(step_type), ptr(ep_len), ptr(last_obs), int(O), int(t_star + 1),
             int(T_ws), int(N), ptr(ws), ptr(lengths), ptr(ep_off), ptr(last_observations),
             ptr(ep_env), ptr(ep_end), _stream(step_type))
    return lengths, ep_off, last_observations, ep_env, ep_end


def pack_rowmap(step_type, ep_len, ep_off, t_star):
    """int32 [t_star+1, N]: packed row of every step, -1 for steps of unfinished episodes."""
    N = step_type.shape[1]
    row_map = torch.empty((int(t_star) + 1, N), dtype=torch.int32, device=step_type.device)
    with torch.cuda.device(step_type.device):
        call('b2g_pack_rowmap', ptr(step_type), ptr(ep_len), ptr(ep_off), int(t_star + 1), int(N), ptr(row_map),
             _stream(step_type))
    return row_map


def pack_fields(step_type, ep_len, ep_off, t_star, n_steps, fields, row_map=None):
    """fields: list of time-major tensors [T,N,...]; returns packed [n_steps,...] tensors.
    `row_map` (pack_rowmap) can be passed to share it between several calls on one rollout."""
    N = step_type.shape[1]
    dev = step_type.device
    outs = []
    for f in fields:
        if not f.is_contiguous() or not f.is_cuda:
            raise ValueError('pack fields must be contiguous CUDA tensors')
        outs.append(torch.empty((n_steps, ) + tuple(f.shape[2:]), dtype=f.dtype, device=dev))
    if row_map is None:
        row_map = pack_rowmap(step_type, ep_len, ep_off, t_star)
    for i in range(0, len(fields), 12):
        fs, os_ = fields[i:i + 12], outs[i:i + 12]
        k = len(fs)
        src = (ctypes.c_void_p * k)(*[f.data_ptr() for f in fs])
        dst = (ctypes.c_void_p * k)(*[o.data_ptr() for o in os_])
        nbytes = (ctypes.c_int * k)(*[f.element_size() * int(math.prod(f.shape[2:])) for f in fs])
        with torch.cuda.device(dev):
            call('b2g_pack_fields_mapped', ptr(row_map), int(t_star + 1), int(N), k, src, dst, nbytes,
                 _stream(step_type))
    return outs


def pack_fields_scatter(step_type, ep_len, ep_off, t_star, n_steps, fields):
    """Single-kernel form (thread per env, scattered stores): the first implementation, kept as the
    cross-check of the tiled path."""
    N = step_type.shape[1]
    dev = step_type.device
    outs = [torch.empty((n_steps, ) + tuple(f.shape[2:]), dtype=f.dtype, device=dev) for f in fields]
    for i in range(0, len(fields), 12):
        fs, os_ = fields[i:i + 12], outs[i:i + 12]
        k = len(fs)
        src = (ctypes.c_void_p * k)(*[f.data_ptr() for f in fs])
        dst = (ctypes.c_void_p * k)(*[o.data_ptr() for o in os_])
        nbytes = (ctypes.c_int * k)(*[f.element_size() * int(math.prod(f.shape[2:])) for f in fs])
        with torch.cuda.device(dev):
            call('b2g_pack_fields', ptr(step_type), ptr(ep_len), ptr(ep_off), int(t_star + 1), int(N),
                 k, src, dst, nbytes, _stream(step_type))
    return outs


# ------------------------------------------------------------- MLP update --
LOSS_PPO, LOSS_VPG, LOSS_VF, LOSS_NONE = 0, 1, 2, 3
FLAG_ENT_REG, FLAG_SOFTPLUS_ENT, FLAG_STOP_ENT_GRAD = 1, 2, 4
MLP_NSTATS = 8
INF = float('inf')


class MlpShape:
    """(obs_dim, hidden, out_dim) of a two-hidden-layer Gaussian MLP + its workspace."""

    def __init__(self, obs_dim, hidden, out_dim, device):
        self.O, self.H, self.A = int(obs_dim), int(hidden), int(out_dim)
        lib = _lib.load()
        if lib.b2g_mlp_num_partials(self.O, self.H, self.A, 1, 1) < 0:
            raise NotImplementedError(lib.b2g_last_error().decode())
        self.P = int(lib.b2g_mlp_param_count(self.O, self.H, self.A))
        self.device = device
        self.ws = torch.empty(int(lib.b2g_mlp_ws_floats(self.O, self.H, self.A)),
                              dtype=torch.float32, device=device)

    def num_partials(self, backward, rows):
        return int(_lib.load().b2g_mlp_num_partials(self.O, self.H, self.A, int(backward), int(rows)))


def mlp_minibatch_grad(shape, params, obs, *, idx=None, start=0, mb=None, act=None, adv=None,
                       old_logp=None, ret=None, loss_kind, flags=0, clip=0.2, ent_coeff=0.0,
                       min_log_std=-INF, max_log_std=INF, step_counter=None):
    """Launch the fused fwd/loss/bwd kernel for one minibatch; partials land in shape.ws."""
    _chk(params, torch.float32, 'params')
    _chk(obs, torch.float32, 'obs')
    _chk(idx, torch.int32, 'idx')
    for name, t in (('act', act), ('adv', adv), ('old_logp', old_logp), ('ret', ret)):
        _chk(t, torch.float32, name)
    if params.numel() != shape.P:
        raise ValueError('params has {} elements, expected {}'.format(params.numel(), shape.P))
    if mb is None:
        mb = idx.numel() if idx is not None else obs.shape[0] - start
    with torch.cuda.device(shape.device):
        call('b2g_mlp_minibatch_grad', shape.O, shape.H, shape.A, ptr(params), ptr(idx), int(start),
             int(mb), ptr(obs), ptr(act), ptr(adv), ptr(old_logp), ptr(ret), int(loss_kind),
             int(flags), float(clip), float(ent_coeff), float(min_log_std), float(max_log_std),
             ptr(shape.ws), ptr(step_counter), _stream(params))
    return shape.num_partials(True, mb)


def mlp_eval(shape, params, obs, *, idx=None, start=0, n=None, act=None, adv=None, old_logp=None,
             ret=None, old_mean=None, old_log_std=None, loss_kind=LOSS_NONE, flags=0, clip=0.2,
             ent_coeff=0.0, min_log_std=-INF, max_log_std=INF, mean_out=None):
    _chk(params, torch.float32, 'params')
    _chk(obs, torch.float32, 'obs')
    _chk(idx, torch.int32, 'idx')
    _chk(mean_out, torch.float32, 'mean_out')
    _chk(old_log_std, torch.float32, 'old_log_std')
    if n is None:
        n = idx.numel() if idx is not None else obs.shape[0] - start
    with torch.cuda.device(shape.device):
        call('b2g_mlp_eval', shape.O, shape.H, shape.A, ptr(params), ptr(idx), int(start), int(n),
             ptr(obs), ptr(act), ptr(adv), ptr(old_logp), ptr(ret), ptr(old_mean),
             ptr(old_log_std), int(loss_kind), int(flags), float(clip), float(ent_coeff),
             float(min_log_std), float(max_log_std), ptr(mean_out), ptr(shape.ws), _stream(params))
    return shape.num_partials(False, n)


def mlp_reduce(shape, n_partials, scale, *, grad_out=None, stats_out=None, params=None, adam_m=None,
               adam_v=None, step_counter=None, lr=0.0, beta1=0.9, beta2=0.999, eps=1e-8):
    _chk(grad_out, torch.float32, 'grad_out')
    _chk(stats_out, torch.float64, 'stats_out')
    with torch.cuda.device(shape.device):
        call('b2g_mlp_reduce', ptr(shape.ws), int(n_partials), shape.P, float(scale), ptr(grad_out),
             ptr(stats_out), ptr(params), ptr(adam_m), ptr(adam_v), ptr(step_counter), float(lr),
             float(beta1), float(beta2), float(eps), _stream(shape.ws))


def mlp_path():
    """0: tcgen05 for hidden 64 (default); 1: CUDA-core fp32 everywhere (B2G_MLP_PATH=ffma)."""
    return int(_lib.load().b2g_get_mlp_path())


def set_mlp_path(path):
    _lib.check(_lib.load().b2g_set_mlp_path(int(path)), 'b2g_set_mlp_path')


SHUFFLE_NONE, SHUFFLE_DEVICE, SHUFFLE_IDS = 0, 1, 2
_loop_ws = {}


def mlp_train_loop(nets, comm=None):
    """All minibatch steps of `_train` for one or two networks in one cooperative launch
    (b2g_mlp_train_loop).  `nets`: dicts with shape, params, adam_m, adam_v, step_counter, lr,
    beta1, beta2, eps, obs, [act, adv, old_logp | ret], loss_kind, flags, clip, ent_coeff,
    min_log_std, max_log_std, minibatch (None = full batch), epochs, shuffle, shuffle_seed, idx.
    `comm`: a `_dist.PeerComm` when several ranks train one model (every rank calls this)."""
    lib = _lib.load()
    arr = (_lib.TrainNet * len(nets))()
    dev = nets[0]['params'].device
    for t, n in zip(arr, nets):
        shape = n['shape']
        for name in ('params', 'adam_m', 'adam_v', 'obs', 'act', 'adv', 'old_logp', 'ret'):
            _chk(n.get(name), torch.float32, name)
        _chk(n.get('idx'), torch.int32, 'idx')
        _chk(n['step_counter'], torch.int64, 'step_counter')
        if n['params'].numel() != shape.P:
            raise ValueError('params has {} elements, expected {}'.format(n['params'].numel(), shape.P))
        B = n['obs'].shape[0]
        t.obs_dim, t.hidden, t.out_dim = shape.O, shape.H, shape.A
        t.params, t.adam_m, t.adam_v = ptr(n['params']), ptr(n['adam_m']), ptr(n['adam_v'])
        t.step_counter = ptr(n['step_counter'])
        t.lr, t.beta1, t.beta2, t.eps = float(n['lr']), float(n['beta1']), float(n['beta2']), float(n['eps'])
        t.obs, t.act, t.adv = ptr(n['obs']), ptr(n.get('act')), ptr(n.get('adv'))
        t.old_logp, t.ret = ptr(n.get('old_logp')), ptr(n.get('ret'))
        t.loss_kind, t.flags = int(n['loss_kind']), int(n.get('flags', 0))
        t.clip, t.ent_coeff = float(n.get('clip', 0.2)), float(n.get('ent_coeff', 0.0))
        t.min_log_std, t.max_log_std = float(n.get('min_log_std', -INF)), float(n.get('max_log_std', INF))
        t.n_rows = B
        t.minibatch = int(n['minibatch']) if n.get('minibatch') else 0
        t.epochs, t.shuffle = int(n['epochs']), int(n.get('shuffle', SHUFFLE_NONE))
        t.shuffle_seed = int(n.get('shuffle_seed', 0)) & (2**64 - 1)
        idx = n.get('idx')
        if idx is not None and idx.numel() != t.epochs * B:
            raise ValueError('idx must hold epochs x n_rows ids')
        t.idx = ptr(idx)
    need = int(lib.b2g_mlp_train_loop_ws_floats())
    ws = _loop_ws.get(dev)
    if ws is None or ws.numel() < need:
        ws = _loop_ws[dev] = torch.empty(need, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        if comm is None:
            call('b2g_mlp_train_loop', arr, len(nets), ptr(ws), ws.numel(), _stream(ws))
        else:  # multi-GPU: gradients all-reduced inside the kernel over peer-mapped memory
            call('b2g_mlp_train_loop_dist', arr, len(nets), comm.struct(), ptr(ws), ws.numel(), _stream(ws))


def adam_step(params, grad, grad_scale, adam_m, adam_v, step_counter, lr, beta1=0.9, beta2=0.999,
              eps=1e-8):
    with torch.cuda.device(params.device):
        call('b2g_adam_step', ptr(params), ptr(grad), float(grad_scale), ptr(adam_m), ptr(adam_v),
             ptr(step_counter), float(lr), float(beta1), float(beta2), float(eps),
             int(params.numel()), _stream(params))


def permutation(n, seed, counter=0, counter_dev=None, out=None, device=None):
    """Random permutation of 0..n-1 (int32) from (seed, counter [+ *counter_dev])."""
    if out is None:
        out = torch.empty(n, dtype=torch.int32, device=device)
    _chk(out, torch.int32, 'out')
    with torch.cuda.device(out.device):
        call('b2g_permutation', int(seed) & (2**64 - 1), ptr(counter_dev), int(counter), int(n),
             ptr(out), _stream(out))
    return out


# ------------------------------------------------------------- GridWorld --
class GridState:
    """Device state of N GridWorld env copies: cells u8 [n_row*n_col], state / step_cnt int32 [N]."""

    def __init__(self, cells, n_row, n_col, start_state, n, device):
        self.n_row, self.n_col, self.start, self.n, self.device = int(n_row), int(n_col), int(start_state), int(n), device
        self.cells = torch.as_tensor(cells, dtype=torch.uint8).reshape(-1).to(device)
        self.state = torch.full((n, ), self.start, dtype=torch.int32, device=device)
        self.step_cnt = torch.full((n, ), -1, dtype=torch.int32, device=device)

    def reset(self):
        self.state.fill_(self.start)
        self.step_cnt.zero_()


class GridBuffers:
    """Time-major [T, N] rollout buffers of b2g_gridworld_rollout."""

    def __init__(self, T, n, device):
        self.T = int(T)
        i32 = dict(dtype=torch.int32, device=device)
        self.obs = torch.empty(T, n, 1, **i32)
        self.act = torch.empty(T, n, 1, **i32)
        self.rew = torch.empty(T, n, dtype=torch.float32, device=device)
        self.step_type = torch.empty(T, n, dtype=torch.uint8, device=device)
        self.last_obs = torch.zeros(T, n, 1, **i32)
        self.ep_len = torch.zeros(T, n, **i32)


def gridworld_step(gs, max_episode_length, actions):
    """Environment.step for all envs; returns (obs int32 [N], reward f32 [N], step_type u8 [N])."""
    _chk(actions, torch.int32, 'actions')
    if actions.shape != (gs.n, ):
        raise ValueError('actions must be int32 [{}]'.format(gs.n))
    obs = torch.empty(gs.n, dtype=torch.int32, device=gs.device)
    rew = torch.empty(gs.n, dtype=torch.float32, device=gs.device)
    st = torch.empty(gs.n, dtype=torch.uint8, device=gs.device)
    err = torch.zeros(1, dtype=torch.int32, device=gs.device)
    with torch.cuda.device(gs.device):
        call('b2g_gridworld_step', ptr(gs.cells), gs.n_row, gs.n_col, int(max_episode_length or 0), ptr(gs.state),
             ptr(gs.step_cnt), ptr(actions), ptr(obs), ptr(rew), ptr(st), ptr(err), gs.n, _stream(actions))
    bad = int(err.item())
    if bad:
        raise RuntimeError('reset() must be called before step()! (env {}, or an action outside 0..3)'.format(bad - 1))
    return obs, rew, st


def gridworld_rollout(gs, max_episode_length, table, buf, t0, n_steps):
    """b2g_gridworld_rollout into rows t0..t0+n_steps of `buf`."""
    _chk(table, torch.int32, 'table')
    if table.numel() != gs.cells.numel():
        raise ValueError('scripted_actions needs one action per grid cell')
    if t0 + n_steps > buf.T:
        raise ValueError('rollout rows exceed the buffer')
    with torch.cuda.device(gs.device):
        call('b2g_gridworld_rollout', ptr(gs.cells), gs.n_row, gs.n_col, gs.start, int(max_episode_length or 0),
             ptr(table), ptr(gs.state), ptr(gs.step_cnt), int(n_steps), gs.n, ptr(buf.obs[t0:]), ptr(buf.act[t0:]),
             ptr(buf.rew[t0:]), ptr(buf.step_type[t0:]), ptr(buf.last_obs[t0:]), ptr(buf.ep_len[t0:]),
             _stream(table))


# ------------------------------------------------------------------ TRPO --
def mlp_fvp(shape, params, vec, obs, *, idx=None, start=0, n=None, min_log_std=-INF, max_log_std=INF,
            reg_coeff=1e-5, out=None, scratch=None):
    """Fisher-vector product of the Gaussian MLP policy over `n` rows (b2g_mlp_fvp): the Hessian of
    mean KL(old || new) at new == old times `vec`, plus reg_coeff * vec."""
    _chk(params, torch.float32, 'params')
    _chk(vec, torch.float32, 'vec')
    _chk(obs, torch.float32, 'obs')
    _chk(idx, torch.int32, 'idx')
    if params.numel() != shape.P or vec.numel() != shape.P:
        raise ValueError('params / vec must have {} elements'.format(shape.P))
    if n is None:
        n = idx.numel() if idx is not None else obs.shape[0] - start
    if out is None:
        out = torch.empty(shape.P, dtype=torch.float32, device=params.device)
    if scratch is None or scratch.numel() < n * shape.A:
        scratch = torch.empty(n * shape.A, dtype=torch.float32, device=params.device)
    with torch.cuda.device(shape.device):
        call('b2g_mlp_fvp', shape.O, shape.H, shape.A, ptr(params), ptr(vec), ptr(idx), int(start), int(n),
             ptr(obs), float(min_log_std), float(max_log_std), float(reg_coeff), ptr(scratch), ptr(shape.ws),
             ptr(out), _stream(params))
    return out


def cg_init(b, x, r, p, scal):
    with torch.cuda.device(b.device):
        call('b2g_cg_init', ptr(b), ptr(x), ptr(r), ptr(p), ptr(scal), int(b.numel()), _stream(b))


def cg_step(z, x, r, p, scal, residual_tol=1e-10):
    with torch.cuda.device(z.device):
        call('b2g_cg_step', ptr(z), ptr(x), ptr(r), ptr(p), ptr(scal), int(z.numel()), float(residual_tol), _stream(z))


def cg_sanitize(x):
    with torch.cuda.device(x.device):
        call('b2g_cg_sanitize', ptr(x), int(x.numel()), _stream(x))


def cg_descent(x, Ax, max_constraint_value, descent, scal):
    with torch.cuda.device(x.device):
        call('b2g_cg_descent', ptr(x), ptr(Ax), float(max_constraint_value), ptr(descent), ptr(scal),
             int(x.numel()), _stream(x))


def line_search_point(prev, descent, ratio, params):
    with torch.cuda.device(prev.device):
        call('b2g_line_search_point', ptr(prev), ptr(descent), float(ratio), ptr(params), int(prev.numel()),
             _stream(prev))


def train_loop_kind(shapes, world=1):
    """b2g_mlp_train_loop_kind for MlpShape objects: 1 tcgen05 loop, 2 general CUDA-core loop, 0 none."""
    arr = (_lib.TrainNet * len(shapes))()
    for t, sh in zip(arr, shapes):
        t.obs_dim, t.hidden, t.out_dim = sh.O, sh.H, sh.A
    return int(_lib.load().b2g_mlp_train_loop_kind(arr, len(shapes), int(world)))


def loop_shape_built(shape):
    """True when the tcgen05 fused training loop is compiled for this (obs, out) pair (hidden 64)."""
    return shape.H == 64 and (shape.O, shape.A) in ((3, 1), (3, 2), (4, 1))
