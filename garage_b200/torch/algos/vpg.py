"""Vanilla Policy Gradient (REINFORCE) with GAE -- drop-in for the reference's
`garage.torch.algos.VPG` (torch/algos/vpg.py:19-455): same constructor, `train(trainer)`,
`_train_once(itr, eps)`, same dowel keys; the arithmetic runs in libb2g kernels:

  baselines        b2g_mlp_eval            (vpg.py:155-156  value_function(obs))
  GAE + returns    b2g_gae_time_major      (vpg.py:149-153, 349-369; torch/_functions.py:25-85)
  normalisation    b2g_moments/_normalize  (vpg.py:371-377)
  minibatch steps  b2g_mlp_minibatch_grad + b2g_mlp_reduce(+Adam)   (vpg.py:230-293)
  logging passes   b2g_mlp_eval            (vpg.py:168-184)
"""
import collections

import numpy as np
import torch

from garage_b200 import ops
from garage_b200._dtypes import DeviceEpisodeBatch, StepType
from garage_b200.tabular import tabular
from garage_b200.torch.algos import _dist
from garage_b200.torch.optimizers import OptimizerWrapper

_INF = float('inf')


class RLAlgorithm:
    """Reference np/algos/rl_algorithm.py:5-25."""

    def train(self, trainer):
        raise NotImplementedError


class _Flat:
    """Episode-major packed training arrays on the device."""
    # obs [B,O], act [B,A], rew/ret/adv/old_logp [B], old_mean [B,A], old_log_std: device
    # float32[1], lengths [n_eps] int64, step_types [B] u8, success [B] u8|None, fixed_len int|None


class VPG(RLAlgorithm):
    """See the reference docstring (vpg.py:19-55) for the shared arguments.

    Extra keyword-only arguments.  The defaults reproduce the reference for fixed-length batches;
    for batches with early-terminated episodes the default `pad_value_mode='mask'` is a DELIBERATE
    deviation (see below) and `pad_value_mode='reference'` is the bit-for-bit parity mode:
        shuffle (str): 'device' (counter-based Feistel permutation kernel) or 'numpy'
            (np.random.shuffle on the host exactly like BatchDataset -- parity mode).
        pad_value_mode (str): 'mask' (default) cuts GAE at episode ends (what the reference's
            docstring describes, NOT what its code computes for ragged batches); 'reference'
            reproduces quirk Q5 (baselines evaluated on zero-padded observations leak V(0) into
            early-terminated episodes) and is what the golden `train_once` fixtures use.
        process_group: torch.distributed group for the gradient all-reduce (default: the
            world group when torch.distributed is initialised with world_size > 1).
        update_mode (str): 'fused' runs every minibatch step of `_train` (both networks) in one
            persistent cooperative launch (b2g_mlp_train_loop; hidden sizes (64, 64), one rank);
            'stepwise' launches one gradient + one Adam kernel per minibatch step; 'auto'
            (default) picks 'fused' whenever it applies.
    """
    _loss_kind = ops.LOSS_VPG
    _clip = 0.0

    def __init__(self, env_spec, policy, value_function, sampler, policy_optimizer=None,
                 vf_optimizer=None, num_train_per_epoch=1, discount=0.99, gae_lambda=1,
                 center_adv=True, positive_adv=False, policy_ent_coeff=0.0,
                 use_softplus_entropy=False, stop_entropy_gradient=False,
                 entropy_method='no_entropy', *, shuffle='device', pad_value_mode='mask',
                 shuffle_seed=0, process_group=None, log_stats=True, update_mode='auto'):
        self._discount = discount
        self.policy = policy
        self.max_episode_length = env_spec.max_episode_length
        self._value_function = value_function
        self._gae_lambda = gae_lambda
        self._center_adv = center_adv
        self._positive_adv = positive_adv
        self._policy_ent_coeff = policy_ent_coeff
        self._use_softplus_entropy = use_softplus_entropy
        self._stop_entropy_gradient = stop_entropy_gradient
        self._entropy_method = entropy_method
        self._n_samples = num_train_per_epoch
        self._env_spec = env_spec
        self._maximum_entropy = (entropy_method == 'max')
        self._entropy_regularzied = (entropy_method == 'regularized')
        self._check_entropy_configuration(entropy_method, center_adv, stop_entropy_gradient,
                                          policy_ent_coeff)
        self._episode_reward_mean = collections.deque(maxlen=100)
        self._sampler = sampler
        self._policy_optimizer = policy_optimizer or OptimizerWrapper(torch.optim.Adam, policy)
        self._vf_optimizer = vf_optimizer or OptimizerWrapper(torch.optim.Adam, value_function)
        if shuffle not in ('device', 'numpy'):
            raise ValueError("shuffle must be 'device' or 'numpy'")
        if pad_value_mode not in ('mask', 'reference'):
            raise ValueError("pad_value_mode must be 'mask' or 'reference'")
        self._shuffle = shuffle
        self._shuffle_seed = int(shuffle_seed)
        self._pad_value_mode = pad_value_mode
        self._pg = process_group
        self._log_stats = log_stats
        if update_mode not in ('auto', 'fused', 'stepwise'):
            raise ValueError("update_mode must be 'auto', 'fused' or 'stepwise'")
        self._update_mode = update_mode
        self._comm = None
        self._use_peer = None
        self._dist_B = None
        self.last_stats = {}

    @staticmethod
    def _check_entropy_configuration(entropy_method, center_adv, stop_entropy_gradient,
                                     policy_ent_coeff):
        """Reference vpg.py:109-125 (same errors)."""
        if entropy_method not in ('max', 'regularized', 'no_entropy'):
            raise ValueError('Invalid entropy_method')
        if entropy_method == 'max':
            if center_adv:
                raise ValueError('center_adv should be False when entropy_method is max')
            if not stop_entropy_gradient:
                raise ValueError('stop_gradient should be True when entropy_method is max')
        if entropy_method == 'no_entropy':
            if policy_ent_coeff != 0.0:
                raise ValueError('policy_ent_coeff should be zero when there is no entropy method')

    @property
    def discount(self):
        return self._discount

    def __getstate__(self):
        """Snapshots (trainer.py:263-293) carry parameters, optimiser state and hyper-parameters;
        the peer-memory communicator and the process group belong to the running job and are
        re-created on first use after `Trainer.restore`."""
        state = self.__dict__.copy()
        state['_comm'] = None
        state['_use_peer'] = None
        state['_pg'] = None
        state['_dist_B'] = None
        return state

    # ------------------------------------------------------------ distributed --
    def _world(self):
        return _dist.world_size(self._pg)

    def _peer_comm(self):
        """The peer-mapped exchange buffers of this job (one node, CUDA IPC), created on first use by
        a COLLECTIVE call -- every rank reaches its first use at the same point of the same
        iteration.  None when B2G_PEER_STATS=0 asks for the NCCL path of the small exchanges."""
        if self._comm is None and self._use_peer is None:
            import os
            self._use_peer = os.environ.get('B2G_PEER_STATS', '1') != '0'
        if self._comm is None and self._use_peer:
            self._comm = _dist.PeerComm(self.policy.flat.device, self._pg)
        return self._comm if self._use_peer else None

    # ---------------------------------------------------------------- prepare --
    def _entropy_value(self, log_std):
        """Entropy of the (obs independent) Gaussian: sum_a 0.5 + 0.5 log 2pi + log std."""
        A = self.policy.mlp_shape.A
        ent = A * (0.5 + 0.5 * np.log(2 * np.pi)) + A * log_std
        if self._use_softplus_entropy:
            ent = torch.nn.functional.softplus(ent)
        return ent

    def _prepare(self, eps):
        """EpisodeBatch -> flat device arrays with returns and (normalised) advantages."""
        if isinstance(eps, DeviceEpisodeBatch) and eps.dev is not None:
            return self._prepare_device(eps)
        return self._prepare_host(eps)

    def _prepare_device(self, eps):
        dev, tm = eps.dev, eps.tm
        f = _Flat()
        T = eps.t_star + 1
        N = tm.n
        fixed = (eps.n_steps == T * N)  # every env contributed every row: no padding anywhere
        vf = self._value_function
        rew_tm = tm.rew[:T]
        ls_dev = self._log_std_dev(eps, rew_tm.device)
        if self._pad_value_mode == 'reference' and not fixed:
            return self._prepare_padded(eps, dev, ls_dev, int(self.max_episode_length), True)
        kw = dict(discount=self._discount, gae_lambda=self._gae_lambda)
        f.vf_before = None
        if fixed and self._log_stats and not self._maximum_entropy:
            # Every rollout row is in the batch, so the baselines pass can also produce the value
            # function's LossBefore (vpg.py:170-173): returns first (they need no baselines), then ONE
            # forward pass writes V(obs) and sums (V - return)^2, then the advantages.
            _, ret_tm = ops.gae_time_major(rew_tm, None, tm.step_type[:T], want_adv=False, **kw)
            vals = torch.empty(T * N, 1, dtype=torch.float32, device=rew_tm.device)
            nb = ops.mlp_eval(vf.mlp_shape, vf.flat, tm.obs[:T].reshape(T * N, -1), ret=ret_tm.reshape(-1),
                              loss_kind=ops.LOSS_VF, mean_out=vals)
            f.vf_before = torch.empty(8, dtype=torch.float64, device=rew_tm.device)
            ops.mlp_reduce(vf.mlp_shape, nb, 1.0, stats_out=f.vf_before)
            values = vals.reshape(T, N)
            adv_tm, _ = ops.gae_time_major(rew_tm, values, tm.step_type[:T], want_ret=False, **kw)
        else:
            values = vf.module.mean(tm.obs[:T].reshape(T * N, -1)).reshape(T, N)
            if self._maximum_entropy:
                # vpg.py:149-160: returns from the raw rewards, advantages from rewards + coeff * entropy
                rew_ent = rew_tm + (self._policy_ent_coeff * self._entropy_value(ls_dev)).to(torch.float32)
                adv_tm, _ = ops.gae_time_major(rew_ent, values, tm.step_type[:T], want_ret=False, **kw)
                _, ret_tm = ops.gae_time_major(rew_tm, None, tm.step_type[:T], want_adv=False, **kw)
            else:
                adv_tm, ret_tm = ops.gae_time_major(rew_tm, values, tm.step_type[:T], **kw)
        adv, ret = ops.pack_fields(tm.step_type, tm.ep_len, eps.ep_off, eps.t_star, eps.n_steps,
                                   [adv_tm, ret_tm], row_map=getattr(eps, 'row_map', None))
        f.obs, f.act, f.rew, f.ret = dev.observations, dev.actions, dev.rewards, ret
        f.old_logp, f.old_mean = dev.logp, dev.mean
        f.old_log_std = ls_dev
        f.B, f.lengths, f.n_eps = eps.n_steps, dev.lengths, eps.n_episodes
        f.step_types, f.success = dev.step_types, dev.success
        f.fixed_len = (eps.n_steps // eps.n_episodes) if fixed else None
        f.adv = self._normalize(adv)
        return f

    @staticmethod
    def _log_std_dev(eps, device):
        ls = eps._log_std
        if hasattr(ls, 'device_value'):
            return ls.device_value
        return torch.tensor([float(ls)], dtype=torch.float32, device=device)

    def _prepare_padded(self, eps, dev, ls_dev, P, pad_with_v0):
        """The reference's own data flow on padded [N,P] arrays (vpg.py:147-166).
        pad_with_v0=True evaluates the value function on the zero-padded observations
        too (quirk Q5); False leaves 0 beyond the end of each episode."""
        lengths = dev.lengths
        n_eps, B = eps.n_episodes, eps.n_steps
        device = dev.observations.device
        rows = torch.repeat_interleave(torch.arange(n_eps, device=device), lengths)
        starts = torch.cumsum(lengths, 0) - lengths
        cols = torch.arange(B, device=device) - torch.repeat_interleave(starts, lengths)
        rew_p = torch.zeros(n_eps, P, device=device)
        rew_p[rows, cols] = dev.rewards
        vmod = self._value_function.module
        if pad_with_v0:
            obs_p = torch.zeros(n_eps, P, dev.observations.shape[1], device=device)
            obs_p[rows, cols] = dev.observations
            base_p = vmod.mean(obs_p.reshape(n_eps * P, -1)).reshape(n_eps, P)
        else:
            base_p = torch.zeros(n_eps, P, device=device)
            base_p[rows, cols] = vmod.mean(dev.observations).reshape(-1)
        kw = dict(discount=self._discount, gae_lambda=self._gae_lambda)
        if self._maximum_entropy:
            # the reference adds the entropy to EVERY padded entry (vpg.py:158-160)
            rew_ent = rew_p + (self._policy_ent_coeff * self._entropy_value(ls_dev)).to(torch.float32)
            if not pad_with_v0:
                rew_ent = torch.zeros_like(rew_p).index_put_((rows, cols), rew_ent[rows, cols])
            adv_p = ops.gae_episode_major(base_p, rew_ent.contiguous(), **kw)
            _, ret_p = ops.gae_episode_major(base_p, rew_p, want_ret=True, **kw)
        else:
            adv_p, ret_p = ops.gae_episode_major(base_p, rew_p, want_ret=True, **kw)
        f = _Flat()
        f.obs, f.act, f.rew = dev.observations, dev.actions, dev.rewards
        f.ret = ret_p[rows, cols].contiguous()
        f.old_logp, f.old_mean, f.old_log_std = dev.logp, dev.mean, ls_dev
        f.B, f.lengths, f.n_eps = B, lengths, n_eps
        f.step_types, f.success, f.fixed_len = dev.step_types, dev.success, None
        f.adv = self._normalize(adv_p[rows, cols].contiguous())
        return f

    def _prepare_host(self, eps):
        """A host EpisodeBatch (e.g. made by the reference's sampler): upload, then the
        padded data flow with libb2g kernels.  The old policy is the current policy at the
        start of _train_once (vpg.py:107, 201; Q8)."""
        device = self.policy.flat.device
        n = len(eps.rewards)
        dev = DeviceEpisodeBatch._Dev()
        dev.observations = torch.as_tensor(np.asarray(eps.observations, np.float32), device=device).reshape(n, -1).contiguous()
        dev.actions = torch.as_tensor(np.asarray(eps.actions, np.float32), device=device).reshape(n, -1).contiguous()
        dev.rewards = torch.as_tensor(np.asarray(eps.rewards, np.float32), device=device).contiguous()
        dev.lengths = torch.as_tensor(np.asarray(eps.lengths, dtype=np.int64), device=device)
        dev.step_types = torch.as_tensor(np.asarray(eps.step_types, dtype=np.uint8), device=device)
        succ = eps.env_infos.get('success') if hasattr(eps, 'env_infos') else None
        dev.success = None if succ is None else torch.as_tensor(np.asarray(succ, np.uint8), device=device)
        mod = self.policy._module
        dev.mean = mod.mean(dev.observations)
        ls = mod.log_std_value()
        ls_dev = torch.tensor([ls], dtype=torch.float32, device=device)
        dist = torch.distributions.Independent(
            torch.distributions.Normal(dev.mean, float(np.exp(np.float32(ls)))), 1)
        dev.logp = dist.log_prob(dev.actions).contiguous()
        shim = type('E', (), {})()
        shim.n_episodes, shim.n_steps = len(eps.lengths), n
        if self._pad_value_mode == 'reference':
            return self._prepare_padded(shim, dev, ls_dev, int(self.max_episode_length), True)
        return self._prepare_padded(shim, dev, ls_dev, int(np.max(eps.lengths)), False)

    def _normalize(self, adv):
        """vpg.py:371-377 (variance, not std: quirk Q3); global statistics across ranks."""
        if not (self._center_adv or self._positive_adv):
            return adv
        stats = ops.moments(adv)
        if self._world() > 1:
            comm = self._peer_comm()
            if comm is not None:  # 4 doubles per rank through peer memory, combined in one small kernel
                stats = comm.reduce_f64(comm.PEER_MOMENTS, stats, torch.empty_like(stats), scalar0=adv.numel())
            else:
                stats = _dist.combine_moments(stats, adv.numel(), self._pg)
        return ops.normalize_adv(adv, stats, self._center_adv, self._positive_adv)

    # ------------------------------------------------------------ loss set-up --
    def _flags(self):
        fl = 0
        if self._entropy_regularzied:
            fl |= ops.FLAG_ENT_REG
        if self._use_softplus_entropy:
            fl |= ops.FLAG_SOFTPLUS_ENT
        if self._stop_entropy_gradient:
            fl |= ops.FLAG_STOP_ENT_GRAD
        return fl

    def _policy_kwargs(self):
        mod = self.policy._module
        return dict(loss_kind=self._loss_kind, flags=self._flags(), clip=self._clip,
                    ent_coeff=self._policy_ent_coeff, min_log_std=mod.min_log_std,
                    max_log_std=mod.max_log_std)

    def _evaluate(self, f, out, vf_stats=None, before=False):
        """The logging passes of vpg.py:168-173 / 178-184 in two forward launches:
        out[0:8] policy statistics (loss, KL(old||new), entropy sums), out[8:16] value loss
        (`vf_stats`: already computed by the baselines pass, see _prepare_device)."""
        pshape, vshape = self.policy.mlp_shape, self._value_function.mlp_shape
        if before and self._loss_kind == ops.LOSS_PPO and self._entropy_method == 'no_entropy':
            # Before the update the reference's old policy IS the policy (ppo.py:96-132 with
            # vpg.py:168-173: deepcopy, re-synchronised after every _train_once), so the likelihood
            # ratio is exactly 1 on every row: LossBefore = -mean(advantages), KLBefore = 0 -- no
            # forward pass needed (out is zero-initialised; only [0] and [1] of the "before" block
            # are logged).
            out[0] = -f.adv.sum(dtype=torch.float64)
        else:
            nb = ops.mlp_eval(pshape, self.policy.flat, f.obs, act=f.act, adv=f.adv,
                              old_logp=f.old_logp, old_mean=f.old_mean, old_log_std=f.old_log_std,
                              **self._policy_kwargs())
            ops.mlp_reduce(pshape, nb, 1.0, stats_out=out[0:8])
        if vf_stats is not None:
            out[8:16].copy_(vf_stats)
            return
        nb = ops.mlp_eval(vshape, self._value_function.flat, f.obs, ret=f.ret, loss_kind=ops.LOSS_VF)
        ops.mlp_reduce(vshape, nb, 1.0, stats_out=out[8:16])

    # ------------------------------------------------------------------ train --
    def _epoch_ids(self, B, opt, host_ids, step_counter):
        if opt.minibatch_size is None:
            return None
        if self._shuffle == 'numpy':
            return torch.from_numpy(host_ids.astype(np.int32)).to(self.policy.flat.device)
        return ops.permutation(B, self._shuffle_seed, counter_dev=step_counter,
                               device=self.policy.flat.device)

    def _run_minibatches(self, opt, module_owner, f, policy):
        """`for epoch: for minibatch: zero_grad; loss; backward; step` (vpg.py:244-293).

        Several ranks: `minibatch_size` is the GLOBAL size.  Shards may differ in length (envs that
        terminate early), but every rank must issue the same collectives, so all ranks run the same
        number S of steps per epoch (from the longest shard) and rank r contributes ceil(B_r / S) of
        its rows to each; the gradient of a step is the mean over the union of the ranks' rows."""
        shape, flat = module_owner.mlp_shape, module_owner.flat
        m, v, step = opt.state()
        B = f.B
        mbs = opt.minibatch_size or B
        world = self._world()
        counts = None
        if world > 1:
            import torch.distributed as dist
            grad = torch.empty(shape.P, dtype=torch.float32, device=flat.device)
            bmax = torch.tensor([B], dtype=torch.int64, device=flat.device)
            dist.all_reduce(bmax, op=dist.ReduceOp.MAX, group=self._pg)
            bmax = int(bmax.item())
            per_rank = max(1, (opt.minibatch_size // world) if opt.minibatch_size else bmax)
            S = -(-bmax // per_rank)
            mbs = -(-B // S)
            local = [max(0, min(mbs, B - s * mbs)) for s in range(S)]
            counts = torch.tensor(local, dtype=torch.int64, device=flat.device)
            dist.all_reduce(counts, group=self._pg)
            counts = counts.tolist()
            starts = [s * mbs for s in range(S)]
        else:
            starts = list(range(0, B, mbs))
        host_ids = None
        if self._shuffle == 'numpy' and opt.minibatch_size is not None:
            # BatchDataset: ids shuffled at construction and AFTER every epoch, the last one
            # included (minibatch_dataset.py:14-35) -- same draws from numpy's global RNG
            host_ids = np.arange(B)
            np.random.shuffle(host_ids)
        kw = self._policy_kwargs() if policy else dict(loss_kind=ops.LOSS_VF)
        data = dict(act=f.act, adv=f.adv, old_logp=f.old_logp) if policy else dict(ret=f.ret)
        for _ in range(opt.max_optimization_epochs):
            idx = self._epoch_ids(B, opt, host_ids, step)
            for si, start in enumerate(starts):
                mb = max(0, min(mbs, B - start))
                if mb > 0:
                    nb = ops.mlp_minibatch_grad(
                        shape, flat, f.obs, idx=None if idx is None else idx[start:start + mb],
                        start=start if idx is None else 0, mb=mb, step_counter=step, **data, **kw)
                if world > 1:
                    if mb > 0:
                        ops.mlp_reduce(shape, nb, 1.0, grad_out=grad)
                    else:  # this shard is exhausted: it still joins the collective and counts the step
                        grad.zero_()
                        ops.counter_add(step, 1)
                    dist.all_reduce(grad, group=self._pg)
                    ops.adam_step(flat, grad, 1.0 / counts[si], m, v, step, opt.lr, opt.beta1,
                                  opt.beta2, opt.eps)
                else:
                    ops.mlp_reduce(shape, nb, 1.0 / mb, params=flat, adam_m=m, adam_v=v,
                                   step_counter=step, lr=opt.lr, beta1=opt.beta1, beta2=opt.beta2,
                                   eps=opt.eps)
            if host_ids is not None:
                np.random.shuffle(host_ids)

    def _fused_applies(self):
        """True when one persistent launch runs every minibatch step of both networks: the tcgen05 loop
        (hidden 64, the built (obs, out) pairs; also multi-GPU) or the general CUDA-core loop (any built
        shape with one hidden size, one GPU) -- b2g_mlp_train_loop_kind."""
        if self._update_mode == 'stepwise':
            return False
        adam = self._policy_optimizer.cg is None and self._vf_optimizer.cg is None
        kind = ops.train_loop_kind([self.policy.mlp_shape, self._value_function.mlp_shape], self._world()) if adam else 0
        if ops.mlp_path() == 1:
            kind = 0  # B2G_MLP_PATH=ffma: the step-wise CUDA-core kernels are what the caller asked for
        if not kind and self._update_mode == 'fused':
            raise NotImplementedError("update_mode='fused': no fused loop for these networks (tcgen05: hidden 64 with "
                                      "(obs, out) in {(3,1), (3,2), (4,1)}; general: one hidden size in {32, 64, 128}, "
                                      "obs <= 8, out <= 4, one GPU)")
        return bool(kind)

    def _loop_net(self, opt, module_owner, f, policy):
        """One network of the fused loop; draws the reference's BatchDataset shuffles here
        (numpy's global RNG, in the reference's order) when shuffle='numpy'."""
        m, v, step = opt.state()
        B, E = f.B, opt.max_optimization_epochs
        net = dict(shape=module_owner.mlp_shape, params=module_owner.flat, adam_m=m, adam_v=v,
                   step_counter=step, lr=opt.lr, beta1=opt.beta1, beta2=opt.beta2, eps=opt.eps,
                   obs=f.obs, minibatch=opt.minibatch_size, epochs=E)
        if policy:
            net.update(act=f.act, adv=f.adv, old_logp=f.old_logp, **self._policy_kwargs())
        else:
            net.update(ret=f.ret, loss_kind=ops.LOSS_VF)
        if opt.minibatch_size is None:
            net['shuffle'] = ops.SHUFFLE_NONE
        elif self._shuffle == 'numpy':
            ids = np.empty((E, B), dtype=np.int32)
            host_ids = np.arange(B)
            np.random.shuffle(host_ids)  # BatchDataset.__init__ (minibatch_dataset.py:14-24)
            for e in range(E):
                ids[e] = host_ids
                np.random.shuffle(host_ids)  # after every epoch, the last included (:35)
            net.update(shuffle=ops.SHUFFLE_IDS, idx=torch.from_numpy(ids).to(f.obs.device))
        else:
            net.update(shuffle=ops.SHUFFLE_DEVICE, shuffle_seed=self._shuffle_seed)
        return net

    def _train(self, f):
        """vpg.py:230-248: all policy minibatches, then all value-function minibatches (the two
        loops touch disjoint state, so the fused path runs them side by side)."""
        if self._fused_applies():
            world, comm = self._world(), None
            if world > 1:
                # every rank must run the same number of steps: the minibatch size is global, the
                # shards must be equally long (fixed-length episodes); otherwise step-wise + NCCL
                import torch.distributed as dist
                # The decision must be the same on every rank (a rank that skips this all_reduce while
                # another issues it deadlocks the job): the shortcut is taken only for envs that never
                # terminate -- a property of the env class, identical on all ranks, under which the
                # batch size is fixed by the configuration -- and never from what one rank's own
                # batch happened to look like (CartPole / PointEnv shards can be full-length on one
                # rank and ragged on another).
                static = bool(getattr(getattr(self._sampler, '_env', None), 'never_terminates', False))
                if static and self._dist_B == f.B:
                    same = True  # the size every rank agreed on before
                else:
                    b = torch.tensor([f.B, -f.B], dtype=torch.int64, device=f.obs.device)
                    dist.all_reduce(b, op=dist.ReduceOp.MAX, group=self._pg)
                    same = int(b[0]) == -int(b[1])
                    self._dist_B = f.B if (same and static) else None
                if not same:
                    if self._update_mode == 'fused':
                        raise NotImplementedError("update_mode='fused' needs equally long shards on every rank")
                    world = 0
                else:
                    if self._comm is None:
                        self._comm = _dist.PeerComm(f.obs.device, self._pg)
                        self._use_peer = self._use_peer if self._use_peer is not None else True
                    comm = self._comm
            if world:
                nets = [self._loop_net(self._policy_optimizer, self.policy, f, True),
                        self._loop_net(self._vf_optimizer, self._value_function, f, False)]
                if world > 1:
                    for n in nets:
                        if n['minibatch']:
                            n['minibatch'] = max(1, n['minibatch'] // world)  # minibatch_size is the global size
                ops.mlp_train_loop(nets, comm)
                return
        self._run_minibatches(self._policy_optimizer, self.policy, f, policy=True)
        self._run_minibatches(self._vf_optimizer, self._value_function, f, policy=False)

    def _performance(self, f, out):
        """log_performance (_functions.py:233-275) on the device (b2g_episode_stats): out[0:7] =
        [mean discounted return, mean/std/max/min undiscounted return, termination rate,
        success rate (nan when the env reports no 'success')].  Returns the undiscounted returns."""
        per, _ = ops.episode_stats(f.rew, self._discount, step_types=f.step_types, success=f.success,
                                   lengths=None if f.fixed_len else f.lengths, fixed_len=f.fixed_len,
                                   out=out[0:7])
        return per[0]

    def _enqueue_iteration(self, eps):
        """Everything of `_train_once` that runs on the device, without any host
        synchronisation (CUDA-graph capturable when the batch shape is static).
        Returns (stats [48] float64 device tensor, B, n_eps)."""
        f = self._prepare(eps)
        stats = torch.zeros(56, dtype=torch.float64, device=f.obs.device)
        pad_kl = self._pad_value_mode == 'reference' and f.fixed_len is None and self._log_stats
        if pad_kl:
            # the reference logs KL over the zero-PADDED observations too (vpg.py:172-173,
            # 181-184): remember the old policy's mean at the all-zero observation
            zero_obs = torch.zeros(1, f.obs.shape[1], device=f.obs.device)
            old_mean0 = self.policy._module.mean(zero_obs)
        if self._log_stats:
            self._evaluate(f, stats[0:16], vf_stats=getattr(f, 'vf_before', None), before=True)
        self._train(f)
        if self._log_stats:
            self._evaluate(f, stats[16:32])
        self._performance(f, stats[32:39])
        if pad_kl:
            dummy = torch.zeros(1, device=f.obs.device)
            nb = ops.mlp_eval(self.policy.mlp_shape, self.policy.flat, zero_obs,
                              act=torch.zeros(1, f.act.shape[1], device=f.obs.device), adv=dummy,
                              old_logp=dummy, old_mean=old_mean0, old_log_std=f.old_log_std,
                              **self._policy_kwargs())
            ops.mlp_reduce(self.policy.mlp_shape, nb, 1.0, stats_out=stats[40:48])
            stats[39:40].fill_(float(f.n_eps * int(self.max_episode_length) - f.B))  # padded entries
        if self._world() > 1:
            comm = self._peer_comm()
            if comm is not None:
                # the logging scalars of all ranks, combined on the device before the single D2H read:
                # stats[0:40] become the job-wide values, [48] / [49] the job-wide samples / episodes,
                # [50] marks them as reduced (see _log_iteration)
                comm.reduce_f64(comm.PEER_LOGS, stats, stats, scalar0=f.B, scalar1=f.n_eps)
        return stats, f.B, f.n_eps

    def _train_once(self, itr, eps):
        """One policy-optimisation iteration (reference vpg.py:136-206).

        Returns:
            numpy.float64: mean undiscounted return of the batch.
        """
        stats, B, n_eps = self._enqueue_iteration(eps)
        return self._log_iteration(itr, stats.cpu().numpy(), B, n_eps)  # the single D2H read

    def _log_iteration(self, itr, s, B, n_eps):
        """The dowel records of vpg.py:186-199 and log_performance (_functions.py:262-273)."""
        B = float(B)
        world = self._world()
        if world > 1:
            if len(s) > 50 and s[50] == 1.0:  # already combined on the device (_enqueue_iteration)
                B, n_eps = float(s[48]), int(round(s[49]))
            else:
                s, B, n_eps = self._reduce_logs(s, B, n_eps, world)
        pol_before, vf_before = s[0] / B, s[8] / B
        pol_after, vf_after = s[16] / B, s[24] / B
        with tabular.prefix(self.policy.name):
            tabular.record('/LossBefore', pol_before)
            tabular.record('/LossAfter', pol_after)
            tabular.record('/dLoss', pol_before - pol_after)
            tabular.record('/KLBefore', s[1] / B)
            # s[39] padded entries with KL s[41] each (pad_value_mode='reference' only)
            tabular.record('/KL', (s[17] + s[39] * s[41]) / (B + s[39]))
            tabular.record('/Entropy', s[18] / B)
        with tabular.prefix(self._value_function.name):
            tabular.record('/LossBefore', vf_before)
            tabular.record('/LossAfter', vf_after)
            tabular.record('/dLoss', vf_before - vf_after)
        with tabular.prefix('Evaluation/'):
            tabular.record('Iteration', itr)
            tabular.record('NumEpisodes', n_eps)
            tabular.record('AverageDiscountedReturn', s[32])
            tabular.record('AverageReturn', s[33])
            tabular.record('StdReturn', s[34])
            tabular.record('MaxReturn', s[35])
            tabular.record('MinReturn', s[36])
            tabular.record('TerminationRate', s[37])
            if not np.isnan(s[38]):
                tabular.record('SuccessRate', s[38])
        self.last_stats = tabular.as_dict
        self._episode_reward_mean.append(s[33])
        return np.float64(s[33])

    def _reduce_logs(self, s, B, n_eps, world):
        """Logging scalars over ranks: what the reference would log for the union of the shards
        (_dist.reduce_logs: sums, episode-weighted means, the global std, max / min)."""
        t = torch.tensor(s, dtype=torch.float64, device=self.policy.flat.device)
        red, Bt, nt = _dist.reduce_logs(t, B, n_eps, self._pg)
        # s[40:48] (KL at the zero-padded observation) depends on the parameters only: same on every rank
        return np.concatenate([red.cpu().numpy(), s[40:]]), float(Bt.item()), int(round(float(nt.item())))

    def train(self, trainer):
        """Reference vpg.py:208-228."""
        last_return = None
        for _ in trainer.step_epochs():
            for _ in range(self._n_samples):
                eps = trainer.obtain_episodes(trainer.step_itr)
                last_return = self._train_once(trainer.step_itr, eps)
                trainer.step_itr += 1
        return last_return
