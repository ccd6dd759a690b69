"""TEST INFRASTRUCTURE ONLY -- CPU port (numpy + torch-CPU, like the reference itself)
of one full on-policy iteration of garage v2021.03.0:

    LocalSampler/VecWorker rollout  (sampler/local_sampler.py:134-166, sampler/vec_worker.py:107-219)
 -> EpisodeBatch assembly           (_dtypes.py:591-632, np/_functions.py:375-406)
 -> GAE / returns                   (torch/_functions.py:25-85, np/_functions.py:111-128)
 -> PPO/VPG._train_once             (torch/algos/vpg.py:136-293, torch/algos/ppo.py:96-132)

Used ONLY as the checker in tests/ and as the timed CPU baseline of bench.py
(`cpu_baseline`, `--impl reference`): /root/reference cannot travel to the GPU box
(pure Python + missing third-party deps), so this port stands in for it there.
`tests/test_port_vs_reference.py` pins it against the UNMODIFIED reference in the
build container (same seeds -> same batches, losses and parameters).
Nothing under garage_b200/ imports this module.
"""
import copy
import math

import numpy as np
import scipy.signal
import torch
import torch.nn.functional as F
from torch import nn
from torch.distributions import Normal
from torch.distributions.independent import Independent

from oracle import np_oracle as O


# ------------------------------------------------------------------ envs --
class PointNp:
    """envs/point_env.py:25-162, one env, numpy float32."""
    obs_dim, act_dim = 3, 2

    def __init__(self, goal=(1., 1.), arena_size=5., done_bonus=0., never_done=False,
                 max_episode_length=math.inf):
        self.goal = np.array(goal, dtype=np.float32)
        self.kw = dict(goal=goal, arena_size=arena_size, done_bonus=done_bonus, never_done=never_done)
        self.max_episode_length = max_episode_length
        self._step_cnt = None

    def reset(self):
        self._point, obs = O.point_reset(1, self.goal)
        self._step_cnt = 0
        return obs[0]

    def step(self, action):
        if self._step_cnt is None:
            raise RuntimeError('reset() must be called before step()!')
        self._point, obs, rew, done, succ = O.point_step(self._point, action[None], **self.kw)
        self._step_cnt += 1
        st = O.get_step_type(self._step_cnt, None if math.isinf(self.max_episode_length)
                             else self.max_episode_length, bool(done[0]))
        if st >= O.TERMINAL:
            self._step_cnt = None
        return obs[0], float(rew[0]), st


class PendulumNp:
    """gym 0.17.2 Pendulum-v0 behind GymEnv (envs/gym_env.py:201-270), float64 state."""
    obs_dim, act_dim = 3, 1

    def __init__(self, max_episode_length=200, seed=0):
        self.max_episode_length = max_episode_length
        self.rng = np.random.RandomState(seed)
        self._step_cnt = None

    def reset(self):
        high = np.array([np.pi, 1])
        self.state = self.rng.uniform(low=-high, high=high)[None]
        self._step_cnt = 0
        return O.pendulum_obs(self.state)[0]

    def step(self, action):
        self.state, obs, rew, done = O.pendulum_step(self.state, np.asarray(action, np.float32)[None])
        self._step_cnt += 1
        st = O.get_step_type(self._step_cnt, self.max_episode_length, False)
        return obs[0], float(rew[0]), st


class CartPoleNp:
    obs_dim, act_dim = 4, 1

    def __init__(self, max_episode_length=500, seed=0, act_mode=0):
        self.max_episode_length = max_episode_length
        self.rng = np.random.RandomState(seed)
        self.act_mode = act_mode

    def reset(self):
        self.state = self.rng.uniform(low=-0.05, high=0.05, size=(4, ))[None]
        self._step_cnt = 0
        return self.state[0].astype(np.float32)

    def step(self, action):
        self.state, obs, rew, done = O.cartpole_step(self.state, np.asarray(action, np.float32)[None],
                                                     act_mode=self.act_mode)
        self._step_cnt += 1
        st = O.get_step_type(self._step_cnt, self.max_episode_length, bool(done[0]))
        return obs[0], float(rew[0]), st


# ------------------------------------------------------------- networks --
class GaussianMLP(nn.Module):
    """GaussianMLPModule (torch/modules/gaussian_mlp_module.py:195-305) on MultiHeadedMLPModule
    (multi_headed_mlp_module.py:49-151): tanh hidden layers, xavier_uniform weights, zero
    biases, one scalar learned log-std, optional clamp at log(min_std)."""

    def __init__(self, in_dim, out_dim, hidden_sizes=(32, 32), init_std=1.0, min_std=1e-6):
        super().__init__()
        self.log_std = nn.Parameter(torch.Tensor([init_std]).log())
        self.layers = nn.ModuleList()
        prev = in_dim
        for h in hidden_sizes:
            lin = nn.Linear(prev, h)
            nn.init.xavier_uniform_(lin.weight)
            nn.init.zeros_(lin.bias)
            self.layers.append(lin)
            prev = h
        self.out = nn.Linear(prev, out_dim)
        nn.init.xavier_uniform_(self.out.weight)
        nn.init.zeros_(self.out.bias)
        self.min_log_std = None if min_std is None else float(torch.Tensor([min_std]).log())
        self.out_dim = out_dim

    def forward(self, x):
        for lin in self.layers:
            x = torch.tanh(lin(x))
        mean = self.out(x)
        log_std = torch.zeros(*mean.shape) + self.log_std
        if self.min_log_std:
            log_std = log_std.clamp(min=self.min_log_std)
        return Independent(Normal(mean, log_std.exp()), 1)

    def flat(self):
        """garage_b200's flat parameter order."""
        parts = [self.log_std]
        for lin in self.layers:
            parts += [lin.weight, lin.bias]
        parts += [self.out.weight, self.out.bias]
        return torch.cat([p.detach().reshape(-1) for p in parts]).numpy().copy()

    def load_flat(self, flat):
        flat = torch.as_tensor(np.asarray(flat, np.float32))
        i = 0
        with torch.no_grad():
            for p in [self.log_std] + [q for lin in self.layers for q in (lin.weight, lin.bias)] + \
                    [self.out.weight, self.out.bias]:
                n = p.numel()
                p.copy_(flat[i:i + n].reshape(p.shape))
                i += n

    def flat_grad(self):
        parts = [self.log_std] + [q for lin in self.layers for q in (lin.weight, lin.bias)] + \
            [self.out.weight, self.out.bias]
        return torch.cat([p.grad.reshape(-1) for p in parts]).numpy().copy()


# --------------------------------------------------------------- sampler --
def vec_rollout(policy, envs, max_episode_length, num_samples, action_fn=None):
    """LocalSampler.obtain_samples with ONE VecWorker over `envs`
    (local_sampler.py:155-166 + vec_worker.py:107-219): all envs reset, then per time step
    one batched get_actions and a Python loop over envs; a finished episode is gathered
    and its env reset at once; stop when completed samples >= num_samples after a step.
    `action_fn(t, prev_obs)` replaces the policy (replayed actions).  Returns a dict of
    numpy arrays laid out like EpisodeBatch."""
    n = len(envs)
    prev_obs = np.asarray([e.reset() for e in envs])
    lists = [dict(obs=[], act=[], rew=[], st=[], mean=[], log_std=[]) for _ in range(n)]
    lens = [0] * n
    done_eps = []
    completed = 0
    t = 0
    while True:
        if action_fn is None:
            with torch.no_grad():
                dist = policy(torch.from_numpy(prev_obs.astype(np.float32)))
                actions = dist.sample().numpy()
                mean = dist.mean.numpy()
                log_std = (dist.variance**.5).log().numpy()
        else:
            actions = action_fn(t, prev_obs)
            mean = log_std = np.zeros_like(actions)
        finished = False
        for i in range(n):
            obs, rew, st = envs[i].step(actions[i])
            L = lists[i]
            L['obs'].append(prev_obs[i].copy()); L['rew'].append(rew); L['act'].append(actions[i])
            L['st'].append(st); L['mean'].append(mean[i]); L['log_std'].append(log_std[i])
            lens[i] += 1
            prev_obs[i] = obs
            if lens[i] >= max_episode_length or st >= O.TERMINAL:
                done_eps.append(dict(observations=np.asarray(L['obs']), last_observation=np.asarray(obs),
                                     actions=np.asarray(L['act']), rewards=np.asarray(L['rew']),
                                     step_types=np.asarray(L['st'], dtype=np.uint8),
                                     mean=np.asarray(L['mean']), log_std=np.asarray(L['log_std']),
                                     length=lens[i], env=i, t_end=t))
                completed += lens[i]
                lists[i] = dict(obs=[], act=[], rew=[], st=[], mean=[], log_std=[])
                lens[i] = 0
                prev_obs[i] = envs[i].reset()
                finished = True
        t += 1
        if finished and completed >= num_samples:
            break
    cat = lambda k: np.concatenate([e[k] for e in done_eps])
    return dict(observations=cat('observations').astype(np.float32), actions=cat('actions').astype(np.float32),
                rewards=cat('rewards').astype(np.float64), step_types=cat('step_types'),
                mean=cat('mean'), log_std=cat('log_std'),
                last_observations=np.stack([e['last_observation'] for e in done_eps]),
                lengths=np.array([e['length'] for e in done_eps], dtype=np.int64),
                order=[(e['t_end'], e['env']) for e in done_eps])


# ------------------------------------------------------------- algorithm --
def compute_advantages_torch(discount, gae_lambda, max_episode_length, baselines, rewards):
    """torch/_functions.py:25-85 restated with the same torch ops (pad + conv2d)."""
    adv_filter = torch.full((1, 1, 1, max_episode_length - 1), discount * gae_lambda, dtype=torch.float)
    adv_filter = torch.cumprod(F.pad(adv_filter, (1, 0), value=1), dim=-1)
    deltas = (rewards + discount * F.pad(baselines, (0, 1))[:, 1:] - baselines)
    deltas = F.pad(deltas, (0, max_episode_length - 1)).unsqueeze(0).unsqueeze(0)
    return F.conv2d(deltas, adv_filter, stride=1).reshape(rewards.shape)


def discount_cumsum(x, discount):
    """np/_functions.py:111-128."""
    return scipy.signal.lfilter([1], [1, float(-discount)], x[::-1], axis=-1)[::-1]


class PPOPort:
    """torch/algos/vpg.py + ppo.py on torch-CPU autograd, with the reference's defaults
    (ppo.py:47-76) unless overridden.  kind: 'ppo' | 'vpg'."""

    def __init__(self, policy, vf, max_episode_length, kind='ppo', lr=2.5e-4, epochs=10,
                 minibatch_size=64, discount=0.99, gae_lambda=0.97, clip=0.2, center_adv=True,
                 positive_adv=False):
        self.policy, self.vf = policy, vf
        self.old_policy = copy.deepcopy(policy)
        self.P = max_episode_length
        self.kind, self.epochs, self.mbs = kind, epochs, minibatch_size
        self.discount, self.lam, self.clip = discount, gae_lambda, clip
        self.center_adv, self.positive_adv = center_adv, positive_adv
        self.opt_p = torch.optim.Adam(policy.parameters(), lr=lr)
        self.opt_v = torch.optim.Adam(vf.parameters(), lr=lr)

    # vpg.py:349-379
    def advantages(self, rewards_p, lengths, baselines_p):
        adv = compute_advantages_torch(self.discount, self.lam, self.P, baselines_p, rewards_p)
        flat = torch.cat([adv[i][:n] for i, n in enumerate(lengths)])
        if self.center_adv:
            flat = (flat - flat.mean()) / (flat.var() + 1e-8)
        if self.positive_adv:
            flat = flat - flat.min()
        return flat

    # ppo.py:96-132 / vpg.py:434-455, 324-347
    def policy_loss(self, obs, act, adv):
        new_ll = self.policy(obs).log_prob(act)
        if self.kind == 'ppo':
            with torch.no_grad():
                old_ll = self.old_policy(obs).log_prob(act)
            ratio = (new_ll - old_ll).exp()
            obj = torch.min(ratio * adv, torch.clamp(ratio, min=1 - self.clip, max=1 + self.clip) * adv)
        else:
            obj = new_ll * adv
        return -obj.mean()

    def vf_loss(self, obs, ret):
        return -self.vf(obs).log_prob(ret.reshape(-1, 1)).mean()

    @staticmethod
    def minibatches(n, mbs, epochs, inputs):
        """OptimizerWrapper.get_minibatch over BatchDataset (optimizer_wrapper.py:31-49,
        minibatch_dataset.py:4-35): numpy global RNG shuffles."""
        if mbs is None:
            for _ in range(epochs):
                yield inputs
            return
        ids = np.arange(n)
        np.random.shuffle(ids)
        for _ in range(epochs):
            for s in range(0, n, mbs):
                sel = ids[s:s + mbs]
                yield [d[sel] for d in inputs]
            np.random.shuffle(ids)

    def prepare(self, batch):
        """vpg.py:147-166."""
        lengths = batch['lengths']
        obs_p = torch.Tensor(O.pad_batch_array(batch['observations'], lengths, self.P))
        rew_p64 = O.pad_batch_array(batch['rewards'], lengths, self.P)
        rewards_p = torch.Tensor(rew_p64)
        returns_p = torch.Tensor(np.stack([discount_cumsum(r, self.discount) for r in rew_p64]))
        with torch.no_grad():
            baselines_p = self.vf(obs_p).mean.flatten(-2)
        obs = torch.Tensor(batch['observations'])
        act = torch.Tensor(batch['actions'])
        ret = torch.cat([returns_p[i][:n] for i, n in enumerate(lengths)])
        adv = self.advantages(rewards_p, lengths, baselines_p)
        return obs, act, ret, adv, baselines_p

    def train_once(self, batch, log=True):
        """vpg.py:136-206; returns the dict of dowel values."""
        obs, act, ret, adv, _ = self.prepare(batch)
        # KL and entropy are logged over the PADDED observations (vpg.py:172-173, 181-184)
        obs_p = torch.Tensor(O.pad_batch_array(batch['observations'], batch['lengths'], self.P))
        out = {}
        if log:
            with torch.no_grad():
                out['policy_loss_before'] = float(self.policy_loss(obs, act, adv))
                out['vf_loss_before'] = float(self.vf_loss(obs, ret))
        n = obs.shape[0]
        for o, a, d in self.minibatches(n, self.mbs, self.epochs, [obs, act, adv]):
            self.opt_p.zero_grad()
            loss = self.policy_loss(o, a, d)
            loss.backward()
            self.opt_p.step()
        for o, r in self.minibatches(n, self.mbs, self.epochs, [obs, ret]):
            self.opt_v.zero_grad()
            loss = self.vf_loss(o, r)
            loss.backward()
            self.opt_v.step()
        if log:
            with torch.no_grad():
                out['policy_loss_after'] = float(self.policy_loss(obs, act, adv))
                out['vf_loss_after'] = float(self.vf_loss(obs, ret))
                old, new = self.old_policy(obs_p), self.policy(obs_p)
                out['kl'] = float(torch.distributions.kl.kl_divergence(old, new).mean())
                out['entropy'] = float(self.policy(obs_p).entropy().mean())
        self.old_policy.load_state_dict(self.policy.state_dict())
        starts = np.cumsum(batch['lengths']) - batch['lengths']
        und = np.add.reduceat(batch['rewards'], starts)
        out['average_return'] = float(np.mean(und))
        return out


def time_iteration(env_name='pendulum', n_envs=64, horizon=200, hidden=64, epochs=10,
                   minibatch_size=None, iters=1, seed=1, kind='ppo', gae_lambda=0.95):
    """One or more full CPU iterations (rollout + train_once); returns a dict with the
    env-steps processed and the wall-clock split -- the timed CPU baseline of bench.py."""
    import time
    torch.manual_seed(seed)
    np.random.seed(seed)
    mk = dict(pendulum=lambda i: PendulumNp(horizon, seed + i),
              cartpole=lambda i: CartPoleNp(horizon, seed + i),
              point=lambda i: PointNp(max_episode_length=horizon))[env_name]
    envs = [mk(i) for i in range(n_envs)]
    pol = GaussianMLP(envs[0].obs_dim, envs[0].act_dim, (hidden, hidden))
    vf = GaussianMLP(envs[0].obs_dim, 1, (hidden, hidden), min_std=None)
    algo = PPOPort(pol, vf, horizon, kind=kind, epochs=epochs, minibatch_size=minibatch_size,
                   gae_lambda=gae_lambda)
    steps = 0
    t_roll = t_train = 0.0
    for _ in range(iters):
        t0 = time.perf_counter()
        batch = vec_rollout(pol, envs, horizon, n_envs * horizon)
        t1 = time.perf_counter()
        algo.train_once(batch)
        t2 = time.perf_counter()
        steps += int(batch['lengths'].sum())
        t_roll += t1 - t0
        t_train += t2 - t1
    return dict(env_steps=steps, rollout_s=t_roll, train_s=t_train, total_s=t_roll + t_train)
