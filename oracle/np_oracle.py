"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of the reference algorithms
on the on-policy hot path of garage v2021.03.0.

Nothing under `garage_b200/` may import this module: it is the checker for
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline leg, never the
product.  Each function cites the reference file:line (under
/root/reference/src/garage) whose arithmetic it restates.

Pinning: `tests/test_oracle_golden.py` checks every function here against
 (a) the reference's own known-answer tests (compute_advantages KAT
     tests/garage/torch/test_functions.py:86-117, StepType table
     tests/garage/test_dtypes.py:290-318, pad_batch_array golden
     tests/garage/np/test_functions.py:81-88, PointEnv tests
     tests/garage/envs/test_point_env.py:9-81), and
 (b) fixtures in tests/golden/*.npz produced by RUNNING the unmodified
     reference in the build container (oracle/gen_golden.py, committed).
Not pinnable (said here and in DESIGN.md): Pendulum-v0 / CartPole-v1 dynamics
live in gym==0.17.2 (setup.py:8,38), absent from /root/reference and from this
image -> "parity unpinned" for those two step functions; they are restated from
the published gym 0.17.2 source (gym/envs/classic_control/{pendulum,cartpole}.py).
"""
import math

import numpy as np

FIRST, MID, TERMINAL, TIMEOUT = 0, 1, 2, 3


# --------------------------------------------------------------- StepType --
def get_step_type(step_cnt, max_episode_length, done):
    """_dtypes.py:44-68 StepType.get_step_type."""
    if max_episode_length is not None and step_cnt >= max_episode_length:
        return TIMEOUT
    if done:
        return TERMINAL
    if step_cnt == 1:
        return FIRST
    if step_cnt < 1:
        raise ValueError('Expect step_cnt to be >= 1, but got {} instead.'.format(step_cnt))
    return MID


# -------------------------------------------------------------- GAE / returns --
def compute_advantages(discount, gae_lambda, max_episode_length, baselines, rewards):
    """torch/_functions.py:25-85: filter = cumprod([1, g*l, g*l, ...]) in float32,
    deltas = r + g * shift(V) - V with V_P := 0, advantages = correlation of the
    right-zero-padded deltas with the filter (what F.conv2d computes)."""
    baselines = np.asarray(baselines, dtype=np.float32)
    rewards = np.asarray(rewards, dtype=np.float32)
    P = int(max_episode_length)
    filt = np.cumprod(np.concatenate(
        [[np.float32(1.0)], np.full(P - 1, np.float32(discount * gae_lambda), dtype=np.float32)]),
        dtype=np.float32)
    nxt = np.concatenate([baselines[:, 1:], np.zeros((baselines.shape[0], 1), np.float32)], axis=1)
    deltas = (rewards + np.float32(discount) * nxt - baselines).astype(np.float32)
    padded = np.concatenate([deltas, np.zeros((deltas.shape[0], P - 1), np.float32)], axis=1)
    adv = np.zeros_like(deltas)
    for k in range(P):  # O(N P^2), exactly the reference's cost model
        adv += filt[k] * padded[:, k:k + deltas.shape[1]]
    return adv


def discount_cumsum(x, discount):
    """np/_functions.py:111-128: scipy.signal.lfilter([1],[1,-discount], x[::-1])[::-1],
    i.e. the direct-form recurrence y[t] = x[t] + discount * y[t+1] in x's dtype
    (float64 for garage rewards)."""
    x = np.asarray(x)
    y = np.zeros(x.shape, dtype=np.result_type(x.dtype, np.float64))
    acc = np.zeros(x.shape[:-1], dtype=y.dtype)
    for t in range(x.shape[-1] - 1, -1, -1):
        acc = x[..., t] + float(discount) * acc
        y[..., t] = acc
    return y


def gae_time_major(rew, val, step_type, last_val, discount, gae_lambda):
    """Time-major restatement of the two recurrences above with episode
    boundaries given by step_type (TERMINAL/TIMEOUT cut the carry and set
    V_{t+1} := 0, the reference's zero padding: torch/_functions.py:80, Q4).
    Equivalent to padding every episode of column n to its own row and calling
    compute_advantages / discount_cumsum on the rows (checked in tests)."""
    rew = np.asarray(rew, np.float32)
    val = np.asarray(val, np.float32)
    T, N = rew.shape
    adv = np.zeros((T, N), np.float32)
    ret = np.zeros((T, N), np.float32)
    g = np.float32(discount)
    gl = np.float32(np.float32(discount) * np.float32(gae_lambda))
    nv = np.zeros(N, np.float32) if last_val is None else np.asarray(last_val, np.float32).copy()
    ca = np.zeros(N, np.float32)
    cr = np.zeros(N, np.float64)  # returns: float64 lfilter then cast (vpg.py:149-153)
    for t in range(T - 1, -1, -1):
        if step_type is not None:
            cut = np.asarray(step_type[t]) >= TERMINAL
            ca[cut] = 0
            cr[cut] = 0
            nv[cut] = 0
        delta = (rew[t] + g * nv - val[t]).astype(np.float32)
        ca = (delta + gl * ca).astype(np.float32)
        cr = rew[t].astype(np.float64) + float(discount) * cr
        nv = val[t].copy()
        adv[t] = ca
        ret[t] = cr
    return adv, ret


def normalize_advantages(adv_flat, center_adv=True, positive_adv=False):
    """torch/algos/vpg.py:371-377: divides by VARIANCE (unbiased) + 1e-8 (Q3)."""
    a = np.asarray(adv_flat, np.float32)
    if center_adv:
        mean = np.float32(a.astype(np.float64).mean())
        var = np.float32(a.astype(np.float64).var(ddof=1))
        a = ((a - mean) / (var + np.float32(1e-8))).astype(np.float32)
    if positive_adv:
        a = a - a.min()
    return a


# ------------------------------------------------------------ padding utils --
def pad_batch_array(array, lengths, max_length=None):
    """np/_functions.py:375-406."""
    array = np.asarray(array)
    assert array.shape[0] == sum(lengths)
    if max_length is None or max_length < max(lengths):
        max_length = max(lengths)
    out = np.zeros((len(lengths), max_length) + array.shape[1:], dtype=array.dtype)
    start = 0
    for i, n in enumerate(lengths):
        out[i, :n] = array[start:start + n]
        start += n
    return out


def filter_valids(padded, lengths):
    """torch/_functions.py:119-132, concatenated."""
    return np.concatenate([padded[i][:n] for i, n in enumerate(lengths)])


# ------------------------------------------------------------ Philox RNG --
# Not reference arithmetic (the reference samples with torch's global CPU
# generator, stochastic_policy.py:86, which a counter RNG cannot reproduce):
# this restates garage_b200's documented stream so that tests can check
# `action == mean + std * eps(seed, env, t)` exactly and the statistics of eps.
_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)


def philox4x32_10(k0, k1, c0, c1, c2, c3):
    """Philox4x32-10 (Salmon et al. 2011); all args broadcastable uint32 arrays."""
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint32) for c in np.broadcast_arrays(c0, c1, c2, c3))
    k0 = np.uint32(k0)
    k1 = np.uint32(k1)
    with np.errstate(over='ignore'):
        for _ in range(10):
            p0 = _M0 * c0.astype(np.uint64)
            p1 = _M1 * c2.astype(np.uint64)
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            k0 = np.uint32(k0 + _W0)
            k1 = np.uint32(k1 + _W1)
    return c0, c1, c2, c3


def u01f(u):
    return ((u >> np.uint32(8)).astype(np.float32) + np.float32(1.0)) * np.float32(1.0 / 16777216.0)


def u01d(hi, lo):
    v = ((hi.astype(np.uint64) << np.uint64(32)) | lo.astype(np.uint64)) >> np.uint64(11)
    return (v.astype(np.float64) + 0.5) * (1.0 / 9007199254740992.0)


def box_muller(a, b):
    u1, u2 = u01f(a), u01f(b)
    r = np.sqrt(np.float32(-2.0) * np.log(u1)).astype(np.float32)
    ang = (np.float32(2.0) * u2).astype(np.float64) * np.pi
    return (r * np.cos(ang).astype(np.float32)).astype(np.float32), \
           (r * np.sin(ang).astype(np.float32)).astype(np.float32)


def action_noise(seed, gid, t, act_dim):
    """eps[len(gid), act_dim] of garage_b200's action stream (stream id 0)."""
    gid = np.asarray(gid, dtype=np.uint32)
    lo, hi = np.uint32(seed & 0xffffffff), np.uint32((seed >> 32) & 0xffffffff)
    out = np.zeros((gid.shape[0], act_dim), np.float32)
    for q in range((act_dim + 3) // 4):
        x, y, z, w = philox4x32_10(lo, hi, gid, np.uint32(t), np.uint32(0), np.uint32(q))
        n0, n1 = box_muller(x, y)
        n2, n3 = box_muller(z, w)
        for i, nn in enumerate((n0, n1, n2, n3)):
            if q * 4 + i < act_dim:
                out[:, q * 4 + i] = nn
    return out


def reset_uniforms(seed, gid, episode, k):
    """k in {0,1}: two (0,1) doubles per Philox call of the reset stream (id 1)."""
    lo, hi = np.uint32(seed & 0xffffffff), np.uint32((seed >> 32) & 0xffffffff)
    x, y, z, w = philox4x32_10(lo, hi, np.asarray(gid, np.uint32), np.asarray(episode, np.uint32),
                               np.uint32(1), np.uint32(k))
    return u01d(x, y), u01d(z, w)


# ------------------------------------------------------------------- envs --
POINT_DEFAULTS = dict(goal=(1., 1.), arena_size=5., done_bonus=0., never_done=False)


def point_reset(n, goal=(1., 1.)):
    """envs/point_env.py:79-98."""
    goal = np.asarray(goal, np.float32)
    point = np.zeros((n, 2), np.float32)
    dist = np.linalg.norm(point - goal, axis=1).astype(np.float32)
    return point, np.concatenate([point, dist[:, None]], axis=1).astype(np.float32)


def point_step(point, action, goal=(1., 1.), arena_size=5., done_bonus=0., never_done=False):
    """envs/point_env.py:100-162, vectorised over envs; float32 like numpy on
    float32 arrays.  Returns (point', obs, reward float64, done, success)."""
    goal = np.asarray(goal, np.float32)
    a = np.clip(np.asarray(action, np.float32), np.float32(-0.1), np.float32(0.1))
    point = np.clip(np.asarray(point, np.float32) + a, np.float32(-arena_size), np.float32(arena_size))
    d = point - goal
    dist = np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]).astype(np.float32)).astype(np.float32)
    low = np.array([-0.1, -0.1], np.float32)
    thr = np.sqrt(np.float32(low[0] * low[0] + low[1] * low[1]))
    succ = dist < thr
    reward = -dist.astype(np.float64) + np.where(succ, float(done_bonus), 0.0)
    done = succ & (not never_done)
    obs = np.concatenate([point, dist[:, None]], axis=1).astype(np.float32)
    return point, obs, reward, done, succ


PENDULUM_PARAMS = dict(max_speed=8., max_torque=2., dt=.05, g=10.0, m=1., l=1.)


def pendulum_obs(state):
    th, thdot = state[:, 0], state[:, 1]
    return np.stack([np.cos(th), np.sin(th), thdot], axis=1).astype(np.float32)


def pendulum_step(state, action, max_speed=8., max_torque=2., dt=.05, g=10.0, m=1., l=1.):
    """gym==0.17.2 gym/envs/classic_control/pendulum.py step() (third-party, NOT in
    /root/reference; restated from the published source -> parity unpinned).
    state float64 [N,2] = (theta, theta_dot); action float32 [N,1]."""
    th, thdot = state[:, 0].astype(np.float64), state[:, 1].astype(np.float64)
    uf = np.clip(np.asarray(action, np.float32)[:, 0], np.float32(-max_torque), np.float32(max_torque))
    u = uf.astype(np.float64)
    ang = ((th + np.pi) % (2 * np.pi)) - np.pi
    costs = ang**2 + .1 * thdot**2 + .001 * (uf * uf).astype(np.float64)
    newthdot = thdot + (-3 * g / (2 * l) * np.sin(th + np.pi) + 3. / (m * l**2) * u) * dt
    newth = th + newthdot * dt
    newthdot = np.clip(newthdot, -max_speed, max_speed)
    new_state = np.stack([newth, newthdot], axis=1)
    return new_state, pendulum_obs(new_state), -costs, np.zeros(th.shape, bool)


CARTPOLE_PARAMS = dict(gravity=9.8, masscart=1.0, masspole=0.1, length=0.5, force_mag=10.0,
                       tau=0.02, x_threshold=2.4, theta_threshold=12 * 2 * math.pi / 360,
                       act_mode=0)


def cartpole_step(state, action, gravity=9.8, masscart=1.0, masspole=0.1, length=0.5,
                  force_mag=10.0, tau=0.02, x_threshold=2.4,
                  theta_threshold=12 * 2 * math.pi / 360, act_mode=0):
    """gym==0.17.2 gym/envs/classic_control/cartpole.py step(), euler integrator
    (third-party, restated; parity unpinned).  "CartPole-style": the continuous
    action a picks gym's action 1 (push right) iff a > 0 (act_mode 0), or the
    force is force_mag * clip(a, -1, 1) (act_mode 1)."""
    x, x_dot, theta, theta_dot = (state[:, i].astype(np.float64) for i in range(4))
    a = np.asarray(action, np.float32)[:, 0]
    if act_mode == 0:
        force = np.where(a > 0, force_mag, -force_mag)
    else:
        force = force_mag * np.clip(a, np.float32(-1), np.float32(1)).astype(np.float64)
    total_mass = masspole + masscart
    polemass_length = masspole * length
    costheta, sintheta = np.cos(theta), np.sin(theta)
    temp = (force + polemass_length * theta_dot**2 * sintheta) / total_mass
    thetaacc = (gravity * sintheta - costheta * temp) / (
        length * (4.0 / 3.0 - masspole * costheta**2 / total_mass))
    xacc = temp - polemass_length * thetaacc * costheta / total_mass
    x = x + tau * x_dot
    x_dot = x_dot + tau * xacc
    theta = theta + tau * theta_dot
    theta_dot = theta_dot + tau * thetaacc
    new_state = np.stack([x, x_dot, theta, theta_dot], axis=1)
    done = (x < -x_threshold) | (x > x_threshold) | (theta < -theta_threshold) | (theta > theta_threshold)
    return new_state, new_state.astype(np.float32), np.ones(x.shape, np.float64), done
