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
