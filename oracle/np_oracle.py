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


class NormalizedEnvOracle:
    """envs/normalized_env.py:8-171 for n independent env copies (each copy owns its running
    estimates, as every worker's env object does); float64 like the reference's numpy state,
    float32 for the action arithmetic (float32 action, float32 bounds).
    Pinned by tests/golden/normalized_env.npz (the reference wrapper around its PointEnv)."""

    def __init__(self, n, obs_dim, act_low, act_high, scale_reward=1., normalize_obs=False,
                 normalize_reward=False, expected_action_scale=1., obs_alpha=0.001, reward_alpha=0.001):
        self.lb = np.asarray(act_low, np.float32)
        self.ub = np.asarray(act_high, np.float32)
        self.scale_reward, self.normalize_obs, self.normalize_reward = scale_reward, normalize_obs, normalize_reward
        self.es, self.oa, self.ra = expected_action_scale, obs_alpha, reward_alpha
        self.obs_mean = np.zeros((n, obs_dim))      # :50-52
        self.obs_var = np.ones((n, obs_dim))
        self.reward_mean = np.zeros(n)              # :54-55
        self.reward_var = np.ones(n)
        # :91-94, including the reference's `ub != -inf`
        self.scale_action = bool(np.all(self.lb != -np.inf) and np.all(self.ub != -np.inf))

    def action(self, action):
        """:91-101: what the wrapped env receives (float32)."""
        a = np.asarray(action, np.float32)
        if not self.scale_action:
            return a
        es = np.float32(self.es)
        k = (np.float32(0.5) * (self.ub - self.lb)) / es
        return np.clip(self.lb + (a + es) * k, self.lb, self.ub).astype(np.float32)

    def obs(self, obs, rows=None):
        """:117-146 (_update_obs_estimate, _apply_normalize_obs) for the env copies `rows`."""
        obs = np.asarray(obs)
        if not self.normalize_obs:
            return obs
        r = slice(None) if rows is None else rows
        o = obs.astype(np.float64)
        self.obs_mean[r] = (1 - self.oa) * self.obs_mean[r] + self.oa * o
        self.obs_var[r] = (1 - self.oa) * self.obs_var[r] + self.oa * np.square(o - self.obs_mean[r])
        return (o - self.obs_mean[r]) / (np.sqrt(self.obs_var[r]) + 1e-8)

    def reward(self, reward):
        """:124-131, 148-160, then scale_reward (:112)."""
        rew = np.asarray(reward, np.float64)
        if self.normalize_reward:
            self.reward_mean = (1 - self.ra) * self.reward_mean + self.ra * rew
            self.reward_var = (1 - self.ra) * self.reward_var + self.ra * np.square(rew - self.reward_mean)
            rew = rew / (np.sqrt(self.reward_var) + 1e-8)
        return rew * self.scale_reward


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


# ------------------------------------------------- Gaussian MLP (policy / VF) --
LOG_SQRT_2PI = math.log(math.sqrt(2 * math.pi))


def mlp_param_count(O, H, A):
    return 1 + H * O + H + H * H + H + A * H + A


def unflatten_params(flat, O, H, A):
    """garage_b200's flat layout == GaussianMLPPolicy.state_dict() order
    (torch/policies/gaussian_mlp_policy.py:53-87, hidden_sizes=(H,H)):
    [log_std | W1[H,O] | b1[H] | W2[H,H] | b2[H] | W3[A,H] | b3[A]]."""
    flat = np.asarray(flat)
    i = 0
    out = {}
    for name, shape in (('log_std', (1, )), ('W1', (H, O)), ('b1', (H, )), ('W2', (H, H)),
                        ('b2', (H, )), ('W3', (A, H)), ('b3', (A, ))):
        n = int(np.prod(shape))
        out[name] = flat[i:i + n].reshape(shape)
        i += n
    assert i == flat.shape[0]
    return out


def flatten_params(p):
    return np.concatenate([np.asarray(p[k]).ravel() for k in ('log_std', 'W1', 'b1', 'W2', 'b2', 'W3', 'b3')])


def mlp_forward(p, obs, dtype=np.float64):
    """torch/modules/multi_headed_mlp_module.py:136-151 with tanh hidden layers
    (gaussian_mlp_policy.py:56): returns (mean, h1, h2)."""
    x = np.asarray(obs, dtype)
    h1 = np.tanh(x @ p['W1'].astype(dtype).T + p['b1'].astype(dtype))
    h2 = np.tanh(h1 @ p['W2'].astype(dtype).T + p['b2'].astype(dtype))
    mean = h2 @ p['W3'].astype(dtype).T + p['b3'].astype(dtype)
    return mean, h1, h2


def clamped_log_std(p, min_std=1e-6, max_std=None):
    """gaussian_mlp_module.py:171-176 (clamp only; returns (log_std, grad passes))."""
    ls = float(p['log_std'][0])
    lo = -np.inf if min_std is None else float(np.float32(np.log(np.float32(min_std))))
    hi = np.inf if max_std is None else float(np.float32(np.log(np.float32(max_std))))
    return min(max(ls, lo), hi), (lo <= ls <= hi)


def normal_log_prob(x, mean, log_std):
    """Independent(Normal(mean, exp(log_std)), 1).log_prob(x), torch.distributions.normal."""
    var = math.exp(log_std)**2
    return (-((x - mean)**2) / (2 * var) - log_std - LOG_SQRT_2PI).sum(-1)


def normal_entropy(log_std, act_dim):
    return act_dim * (0.5 + 0.5 * math.log(2 * math.pi) + log_std)


def kl_normal(mean_p, log_std_p, mean_q, log_std_q):
    """torch.distributions.kl._kl_normal_normal summed over action dims (Independent)."""
    vr = (math.exp(log_std_p) / math.exp(log_std_q))**2
    t1 = ((mean_p - mean_q) / math.exp(log_std_q))**2
    return (0.5 * (vr + t1 - 1 - math.log(vr))).sum(-1)


def policy_loss_and_grad(flat, O, H, A, obs, act, adv, old_logp, kind='ppo', clip=0.2,
                         min_std=1e-6, max_std=None, ent_coeff=0.0, entropy_regularized=False,
                         use_softplus_entropy=False, stop_entropy_gradient=False):
    """loss = -mean(objective) and d loss / d flat, in float64.
    PPO objective: torch/algos/ppo.py:96-132; VPG: vpg.py:434-455; entropy term:
    vpg.py:324-347, 408-432.  Backward written out by hand (tanh MLP, nn.Linear)."""
    p = unflatten_params(np.asarray(flat, np.float64), O, H, A)
    obs = np.asarray(obs, np.float64)
    act = np.asarray(act, np.float64)
    adv = np.asarray(adv, np.float64)
    n = obs.shape[0]
    mean, h1, h2 = mlp_forward(p, obs)
    ls, ls_pass = clamped_log_std(p, min_std, max_std)
    var = math.exp(ls)**2
    ll = normal_log_prob(act, mean, ls)
    if kind == 'ppo':
        ratio = np.exp(ll - np.asarray(old_logp, np.float64))
        surr = ratio * adv
        rc = np.clip(ratio, 1 - clip, 1 + clip)
        surr_c = rc * adv
        obj = np.minimum(surr, surr_c)
        inside = (ratio >= 1 - clip) & (ratio <= 1 + clip)
        dobj_dll = np.where(inside | (surr < surr_c), ratio * adv, 0.0)
    else:
        obj = ll * adv
        dobj_dll = adv.copy()
    dobj_dls_ent = 0.0
    if entropy_regularized:
        ent = normal_entropy(ls, A)
        dent = float(A)
        if use_softplus_entropy:
            dent *= 1.0 / (1.0 + math.exp(-ent))
            ent = math.log1p(math.exp(ent))
        obj = obj + ent_coeff * ent
        if not stop_entropy_gradient:
            dobj_dls_ent = ent_coeff * dent
    loss = -obj.mean()
    d = act - mean
    dmean = -(dobj_dll[:, None] * d / var) / n
    dls = (-(dobj_dll * ((d**2).sum(-1) / var - A)).sum() / n - dobj_dls_ent) if ls_pass else 0.0
    return loss, _mlp_backward(p, obs, h1, h2, dmean, dls)


def vf_loss_and_grad(flat, O, H, obs, ret):
    """value_functions/gaussian_mlp_value_function.py:81-97: -mean log N(ret; V(obs), exp(s));
    min_std=None -> no clamp (:76-77)."""
    p = unflatten_params(np.asarray(flat, np.float64), O, H, 1)
    obs = np.asarray(obs, np.float64)
    ret = np.asarray(ret, np.float64).reshape(-1, 1)
    n = obs.shape[0]
    mean, h1, h2 = mlp_forward(p, obs)
    ls = float(p['log_std'][0])
    var = math.exp(ls)**2
    loss = -normal_log_prob(ret, mean, ls).mean()
    d = ret - mean
    dmean = -(d / var) / n
    dls = (1.0 - (d**2).sum(-1) / var).sum() / n
    return loss, _mlp_backward(p, obs, h1, h2, dmean, dls)


def _mlp_backward(p, obs, h1, h2, dmean, dls):
    g = {}
    g['log_std'] = np.array([dls])
    g['W3'] = dmean.T @ h2
    g['b3'] = dmean.sum(0)
    dz2 = (dmean @ p['W3']) * (1 - h2**2)
    g['W2'] = dz2.T @ h1
    g['b2'] = dz2.sum(0)
    dz1 = (dz2 @ p['W2']) * (1 - h1**2)
    g['W1'] = dz1.T @ obs
    g['b1'] = dz1.sum(0)
    return flatten_params(g)


def adam_step(param, grad, m, v, step, lr=2.5e-4, beta1=0.9, beta2=0.999, eps=1e-8):
    """torch.optim.Adam single-tensor update (amsgrad=False, weight_decay=0), float32
    state like the reference's CPU tensors; step counts from 1."""
    f = np.float32
    param, grad, m, v = (np.asarray(a, f) for a in (param, grad, m, v))
    m = (m + f(1 - beta1) * (grad - m)).astype(f)
    v = (v * f(beta2) + (f(1 - beta2) * grad) * grad).astype(f)
    bc1 = 1 - beta1**step
    bc2 = 1 - beta2**step
    step_size = f(lr / bc1)
    denom = (np.sqrt(v) / f(math.sqrt(bc2)) + f(eps)).astype(f)
    param = (param - step_size * (m / denom)).astype(f)
    return param, m, v


# ------------------------------------------------ vectorised rollout replay --
def env_reset_states(env, seed, gid, episode, goal=(1., 1.)):
    """Initial (state, obs) of `episode` of each global env id: PointEnv resets to
    the origin (point_env.py:79-98); Pendulum/CartPole draw U(-high, high) like gym
    0.17.2's reset() but from garage_b200's Philox reset stream."""
    n = len(gid)
    if env == 'point':
        return point_reset(n, goal)
    if env == 'pendulum':
        u0, u1 = reset_uniforms(seed, gid, episode, 0)
        st = np.stack([(2 * u0 - 1) * np.pi, (2 * u1 - 1) * 1.0], axis=1)
        return st, pendulum_obs(st)
    u0, u1 = reset_uniforms(seed, gid, episode, 0)
    u2, u3 = reset_uniforms(seed, gid, episode, 1)
    st = np.stack([(2 * u - 1) * 0.05 for u in (u0, u1, u2, u3)], axis=1)
    return st, st.astype(np.float32)


def env_step_any(env, state, action, kw):
    if env == 'point':
        p, obs, rew, done, succ = point_step(state, action, **kw)
        return p, obs, rew, done, succ
    if env == 'pendulum':
        s, obs, rew, done = pendulum_step(state, action, **kw)
    else:
        s, obs, rew, done = cartpole_step(state, action, **kw)
    return s, obs, rew, done, np.zeros(done.shape, bool)


def replay_rollout(env, actions, max_episode_length, seed, gid0=0, first_episode=0, env_kw=None, norm=None):
    """Replay recorded actions [T,N,A] through N vectorised envs with the
    bookkeeping of sampler/vec_worker.py:176-204: per-env step counters,
    StepType.get_step_type (_dtypes.py:44-68), last observation kept at the
    finishing step (Q10), immediate reset of a finished env (:171).
    norm: a NormalizedEnvOracle wrapped around every env copy (envs/normalized_env.py)."""
    env_kw = dict(env_kw or {})
    actions = np.asarray(actions, np.float32)
    T, N, A = actions.shape
    gid = np.arange(gid0, gid0 + N)
    ep = np.full(N, first_episode, np.int64)
    goal = env_kw.get('goal', (1., 1.))
    state, obs = env_reset_states(env, seed, gid, ep, goal)
    if norm is not None:
        obs = norm.obs(obs)
    ep += 1
    O_ = obs.shape[1]
    out = dict(obs=np.zeros((T, N, O_), np.float32), rew=np.zeros((T, N), np.float64),
               step_type=np.zeros((T, N), np.uint8), success=np.zeros((T, N), bool),
               last_obs=np.zeros((T, N, O_), np.float32), ep_len=np.zeros((T, N), np.int32))
    cnt = np.zeros(N, np.int64)
    for t in range(T):
        out['obs'][t] = obs
        act_t = actions[t] if norm is None else norm.action(actions[t])
        state, obs, rew, done, succ = env_step_any(env, state, act_t, env_kw)
        if norm is not None:
            obs, rew = norm.obs(obs), norm.reward(rew)
        cnt += 1
        st = np.where(cnt == 1, FIRST, MID)
        st = np.where(done, TERMINAL, st)
        if max_episode_length:
            st = np.where(cnt >= max_episode_length, TIMEOUT, st)
        out['rew'][t] = rew
        out['step_type'][t] = st
        out['success'][t] = succ
        last = st >= TERMINAL
        if last.any():
            out['last_obs'][t][last] = obs[last]
            out['ep_len'][t][last] = cnt[last]
            rs, ro = env_reset_states(env, seed, gid[last], ep[last], goal)
            if norm is not None:
                ro = norm.obs(ro, rows=last)
            state = np.array(state)
            obs = np.array(obs)
            state[last] = rs
            obs[last] = ro
            ep[last] += 1
            cnt[last] = 0
    return out


def pack_episodes(step_type, ep_len, num_samples, fields):
    """Episode-major packing with the reference's ordering and cut:
    episodes in completion order, ties by env index (vec_worker.py:185-201 loops
    envs in index order inside a step), stop after the first step at which
    completed samples >= num_samples (local_sampler.py:156-166).
    Returns (lengths, [packed fields], [(t_end, n)])."""
    T, N = step_type.shape
    eps = []
    total = 0
    for t in range(T):
        row = np.nonzero(step_type[t] >= TERMINAL)[0]
        for n in row:
            eps.append((t, int(n), int(ep_len[t, n])))
            total += int(ep_len[t, n])
        if total >= num_samples:
            break
    lengths = np.array([L for _, _, L in eps], dtype=np.int64)
    packed = []
    for f in fields:
        parts = [f[t - L + 1:t + 1, n] for t, n, L in eps]
        packed.append(np.concatenate(parts) if parts else f[:0, 0])
    return lengths, packed, [(t, n) for t, n, _ in eps]


# ------------------------------------------------------------------ GridWorld --
GW_CODES = {'S': 0, 'F': 0, '.': 0, 'W': 1, 'x': 1, 'H': 2, 'o': 2, 'G': 3}


def gridworld_step(cells, state, action):
    """One transition of GridWorldEnv (reference envs/grid_world_env.py:129-176, 227-250).
    cells: 2-D array of characters.  Returns (next_state, reward, done)."""
    n_row, n_col = cells.shape
    x, y = divmod(int(state), n_col)
    dx, dy = [(0, -1), (1, 0), (0, 1), (-1, 0)][int(action)]
    nx, ny = min(max(x + dx, 0), n_row - 1), min(max(y + dy, 0), n_col - 1)
    here, there = GW_CODES[cells[x, y]], GW_CODES[cells[nx, ny]]
    nxt = state if (there == 1 or here in (2, 3)) else nx * n_col + ny
    kind = GW_CODES[cells[divmod(int(nxt), n_col)]]
    return int(nxt), float(kind == 3), kind in (2, 3)


def gridworld_vec_rollout(cells, table, max_episode_length, n_envs, num_samples):
    """VecWorker + LocalSampler over n_envs copies of a GridWorldEnv with a ScriptedPolicy
    (sampler/vec_worker.py:107-219, sampler/local_sampler.py:134-166): whole batches of lockstep
    steps until at least num_samples steps of completed episodes exist.  Returns dict of
    observations, actions, rewards, step_types, last_observations, lengths (episode-major, in
    completion order, ties in env order)."""
    start = int(np.argwhere(cells == 'S')[0] @ np.array([cells.shape[1], 1]))
    state = [start] * n_envs
    cnt = [0] * n_envs
    cur = [dict(o=[], a=[], r=[], s=[]) for _ in range(n_envs)]
    out = dict(observations=[], actions=[], rewards=[], step_types=[], last_observations=[], lengths=[])
    total = 0
    while total < num_samples:
        for n in range(n_envs):
            a = int(table[state[n]])
            nxt, rew, done = gridworld_step(cells, state[n], a)
            cnt[n] += 1
            st = get_step_type(cnt[n], max_episode_length, done)
            e = cur[n]
            e['o'].append(state[n]); e['a'].append(a); e['r'].append(rew); e['s'].append(int(st))
            if int(st) >= 2:
                out['observations'] += e['o']; out['actions'] += e['a']; out['rewards'] += e['r']
                out['step_types'] += e['s']; out['last_observations'].append(nxt); out['lengths'].append(cnt[n])
                total += cnt[n]
                cur[n] = dict(o=[], a=[], r=[], s=[])
                state[n], cnt[n] = start, 0
            else:
                state[n] = nxt
    return {k: np.asarray(v) for k, v in out.items()}


# ------------------------------------------------------------------- TRPO --
def mlp_jvp(p, v, obs):
    """Forward-mode derivative of the mean network: d mu / d theta . v for a tangent `v` given as
    an unflattened parameter dict (what autograd's double-backward evaluates inside
    torch/optimizers/conjugate_gradient_optimizer.py:18-66 for the mean parameters)."""
    x = np.asarray(obs, np.float64)
    mean, h1, h2 = mlp_forward(p, x)
    dz1 = x @ v['W1'].T + v['b1']
    dh1 = (1 - h1**2) * dz1
    dz2 = dh1 @ p['W2'].T + h1 @ v['W2'].T + v['b2']
    dh2 = (1 - h2**2) * dz2
    return dh2 @ p['W3'].T + h2 @ v['W3'].T + v['b3'], h1, h2


def fisher_vector_product(flat, O, H, A, obs, vec, reg_coeff=1e-5, min_std=1e-6, max_std=None):
    """`_build_hessian_vector_product(f_constraint, params, reg_coeff)(vec)`
    (conjugate_gradient_optimizer.py:18-66) with f_constraint = mean KL(old || new)
    (torch/algos/vpg.py:381-406) evaluated where new == old (TRPO's full-batch step starts there,
    trpo.py:119-144): the Hessian of the KL of two diagonal Gaussians with a state-independent
    log-std is J^T diag(1/var) J / n on the mean parameters (the network's second-order term
    multiplies mu_old - mu_new = 0) and 2 A on log_std (0 when the clamp blocks the gradient),
    no cross term; plus reg_coeff * vec."""
    p = unflatten_params(np.asarray(flat, np.float64), O, H, A)
    v = unflatten_params(np.asarray(vec, np.float64), O, H, A)
    n = np.asarray(obs).shape[0]
    dmu, h1, h2 = mlp_jvp(p, v, obs)
    ls, ls_pass = clamped_log_std(p, min_std, max_std)
    var = math.exp(ls)**2
    out = _mlp_backward(p, np.asarray(obs, np.float64), h1, h2, dmu / var / n, 0.0)
    out[0] = 2.0 * A * float(v['log_std'][0]) if ls_pass else 0.0
    return out + reg_coeff * np.asarray(vec, np.float64)


def conjugate_gradient(f_Ax, b, cg_iters=10, residual_tol=1e-10):
    """conjugate_gradient_optimizer.py:69-110 (Demmel p. 312), float64."""
    b = np.asarray(b, np.float64)
    p, r, x = b.copy(), b.copy(), np.zeros_like(b)
    rdotr = r @ r
    for _ in range(cg_iters):
        z = f_Ax(p)
        v = rdotr / (p @ z)
        x += v * p
        r -= v * z
        newrdotr = r @ r
        mu = newrdotr / rdotr
        p = r + mu * p
        rdotr = newrdotr
        if rdotr < residual_tol:
            break
    return x


def trpo_policy_step(flat, O, H, A, obs, act, adv, old_logp, old_mean, old_log_std,
                     max_constraint_value=0.01, cg_iters=10, max_backtracks=15, backtrack_ratio=0.8,
                     hvp_reg_coeff=1e-5, accept_violation=False):
    """One `TRPO._train_policy` (trpo.py:119-144) = ConjugateGradientOptimizer.step
    (conjugate_gradient_optimizer.py:141-260): loss gradient of -mean(ratio * adv)
    (trpo.py:90-117), CG on the Fisher-vector product, step size sqrt(2 delta / (x^T A x + 1e-8)),
    backtracking line search on (loss, mean KL(old || new)).  Returns (new flat, info)."""
    flat = np.asarray(flat, np.float64)

    def f_loss(theta):
        return policy_loss_and_grad(theta, O, H, A, obs, act, adv, old_logp, kind='ppo', clip=np.inf)

    def f_kl(theta):
        p = unflatten_params(theta, O, H, A)
        mean, _, _ = mlp_forward(p, obs)
        ls, _ = clamped_log_std(p)
        return float(kl_normal(np.asarray(old_mean, np.float64), float(old_log_std), mean, ls).mean())

    loss_before, g = f_loss(flat)
    f_Ax = lambda v: fisher_vector_product(flat, O, H, A, obs, v, hvp_reg_coeff)
    step_dir = conjugate_gradient(f_Ax, g, cg_iters)
    step_dir[np.isnan(step_dir)] = 0.0
    step_size = np.sqrt(2.0 * max_constraint_value * (1.0 / (step_dir @ f_Ax(step_dir) + 1e-8)))
    if np.isnan(step_size):
        step_size = 1.0
    descent = step_size * step_dir
    new, loss, kl, used = flat, loss_before, 0.0, None
    for k in range(max_backtracks):
        ratio = backtrack_ratio**k
        new = flat - ratio * descent
        loss, kl = f_loss(new)[0], f_kl(new)
        used = k
        if loss < loss_before and kl <= max_constraint_value:
            break
    violated = (np.isnan(loss) or np.isnan(kl) or loss >= loss_before or kl >= max_constraint_value)
    if violated and not accept_violation:
        new = flat
    return new, dict(loss_before=loss_before, loss=loss, kl=kl, backtracks=used, rejected=bool(violated and not accept_violation),
                     grad=g, step_dir=step_dir, step_size=step_size)
