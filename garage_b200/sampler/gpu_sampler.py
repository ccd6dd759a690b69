"""GPU sampler: `Sampler.obtain_samples` (reference sampler/sampler.py:73-93) with
the semantics of `LocalSampler` + `VecWorker` (sampler/local_sampler.py:134-166,
sampler/vec_worker.py:107-219), executed by ONE persistent rollout kernel per
chunk of time steps and assembled into an EpisodeBatch by the pack kernels.

Semantics kept from the reference:
  * every `obtain_samples` pushes the agent update and therefore restarts all
    environments from `reset()` (vec_worker.py:72-73,107-137); partial episodes
    of the previous call are discarded;
  * the batch holds whole episodes only, in completion order (ties in env
    order), and ends with the first time step at which the number of completed
    samples reaches `num_samples` (all episodes finishing in that step included);
  * `total_env_steps` accumulates `sum(lengths)`.
"""
import abc
import math

import numpy as np
import psutil
import torch

from garage_b200 import ops
from garage_b200._dtypes import DeviceEpisodeBatch
from garage_b200.experiment import deterministic as _seeding


class _LazyLogStd:
    """Clamped log-std of the rollout policy, snapshotted on the device (async) and
    read back only if somebody asks for agent_infos['log_std']."""

    def __init__(self, module):
        self._snap = module.flat[0:1].clone()
        self._lo, self._hi = module.min_log_std, module.max_log_std
        self.device_value = self._snap.clamp(min=self._lo, max=self._hi)

    def __float__(self):
        return min(max(float(self._snap.item()), self._lo), self._hi)


class Sampler(abc.ABC):
    """Reference sampler/sampler.py:7-100."""

    @abc.abstractmethod
    def obtain_samples(self, itr, num_samples, agent_update, env_update=None):
        pass

    @abc.abstractmethod
    def shutdown_worker(self):
        pass


class VecWorker:
    """Marker accepted as `worker_class=` for source compatibility: the GPU sampler
    IS a vectorised worker; `worker_args={'n_envs': k}` sets envs per worker."""
    DEFAULT_N_ENVS = 8


class GPUSampler(Sampler):
    """Drop-in for `LocalSampler(agents, envs, max_episode_length=, n_workers=,
    worker_class=VecWorker, worker_args={'n_envs': k})`; the total number of
    environment copies is `n_envs` if given, else `n_workers * worker_args['n_envs']`.

    Args:
        agents (GaussianMLPPolicy or list): the policy (its flat CUDA parameter vector is used); a list
            (one per worker, length `n_workers`) gives every worker its own, as in the reference.
        envs (BatchedEnv or list): environment description (PointEnv / PendulumEnv / CartPoleEnv); a list
            gives every worker its own (multi-task sampling): each worker is then a group of
            `worker_args['n_envs']` device envs and `obtain_samples` visits the groups round-robin.
        max_episode_length (int): required, like the reference (local_sampler.py:63-65).
        n_envs (int): environment copies on this rank.
        seed (int): Philox key (defaults to `deterministic.get_seed()` or 0).
        deterministic (bool): act with the mean of the action distribution (evaluation rollouts,
            np/_functions.py rollout(deterministic=True)); log-probs of such a batch are undefined.
        env_id_offset (int): global id of this rank's first env (rank * n_envs), so
            results do not depend on how envs are sharded over GPUs.
    """

    def __init__(self, agents, envs, *, worker_factory=None, max_episode_length=None,
                 is_tf_worker=False, seed=None, n_workers=None, worker_class=VecWorker,
                 worker_args=None, n_envs=None, env_id_offset=0, keep_agent_infos=True,
                 deterministic=False):
        if worker_factory is None and max_episode_length is None:
            raise TypeError('Must construct a sampler from WorkerFactory or parameters (at least '
                            'max_episode_length)')
        if is_tf_worker:
            raise NotImplementedError('TF workers are out of scope of garage_b200')
        self._groups = None
        if isinstance(agents, list) or isinstance(envs, list):
            # one agent / env PER WORKER (local_sampler.py:50-86, worker_factory.py:68-95)
            lists = [x for x in (agents, envs) if isinstance(x, list)]
            if n_workers is None:
                n_workers = len(lists[0])
            agents_l, envs_l = self._per_worker(agents, n_workers), self._per_worker(envs, n_workers)
            if all(x is agents_l[0] for x in agents_l) and all(x is envs_l[0] for x in envs_l):
                agents, envs = agents_l[0], envs_l[0]  # the same objects everywhere: one vectorised worker
            else:
                if n_envs is not None and n_envs % n_workers:
                    raise ValueError('n_envs must be a multiple of n_workers when agents / envs are lists')
                per = n_envs // n_workers if n_envs is not None else (worker_args or {}).get(
                    'n_envs', VecWorker.DEFAULT_N_ENVS)
                self._groups = [
                    GPUSampler(ag, en, max_episode_length=max_episode_length, seed=seed, n_envs=per,
                               env_id_offset=int(env_id_offset) + i * per, keep_agent_infos=keep_agent_infos,
                               deterministic=deterministic)
                    for i, (ag, en) in enumerate(zip(agents_l, envs_l))]
                n_envs = per * n_workers
                agents, envs = agents_l[0], envs_l[0]
        if n_envs is None:
            per = (worker_args or {}).get('n_envs', VecWorker.DEFAULT_N_ENVS)
            n_envs = per * (n_workers if n_workers is not None else psutil.cpu_count(logical=False))
        if n_envs <= 0:
            raise ValueError('n_envs must be positive')
        self._n_workers = n_workers
        self._policy = agents
        self._env = envs
        self._max_episode_length = max_episode_length
        self._n_envs = int(n_envs)
        if seed is None:
            seed = _seeding.get_seed()
        self._seed = 0 if seed is None else int(seed)
        self._env_id_offset = int(env_id_offset)
        self._keep_agent_infos = keep_agent_infos
        self._deterministic = bool(deterministic)
        self._t_dev = None  # device-side global step counter (Philox counter word)
        self._es = None
        self._gs = None  # GridWorld device state (GridWorldEnv + ScriptedPolicy path)
        self._static_pack = None  # (shape key, row statistics, row map) of a never-terminating env
        self._buf = None
        self.total_env_steps = 0

    @staticmethod
    def _per_worker(objs, n_workers):
        """`WorkerFactory.prepare_worker_messages` (worker_factory.py:68-95): a list must have one entry per
        worker, anything else is given to every worker."""
        if isinstance(objs, list):
            if len(objs) != n_workers:
                raise ValueError('Length of list doesn\'t match number of workers')
            return list(objs)
        return [objs] * n_workers

    @classmethod
    def from_worker_factory(cls, worker_factory, agents, envs):
        return cls(agents, envs, max_episode_length=worker_factory.max_episode_length,
                   seed=getattr(worker_factory, 'seed', None), n_workers=worker_factory.n_workers,
                   worker_args=getattr(worker_factory, 'worker_args', None))

    # ------------------------------------------------------------------ helpers --
    @property
    def n_envs(self):
        return self._n_envs

    def _device(self):
        if hasattr(self._policy, 'flat'):
            return self._policy.flat.device
        return torch.device('cuda', torch.cuda.current_device())

    def _max_len(self):
        m = self._max_episode_length
        if m is None or (isinstance(m, float) and math.isinf(m)):
            return 0
        return int(m)

    def _max_log_std(self):
        # b2g_rollout: max_log_std = -inf clamps the std to exactly 0, i.e. action = mean
        return -math.inf if self._deterministic else self._policy._module.max_log_std

    def _ensure_state(self):
        if self._es is None or self._es.env_id != self._env.env_id:
            self._es = ops.EnvState(self._env.env_id, self._n_envs, self._device())
        if self._t_dev is None:
            self._t_dev = torch.zeros(1, dtype=torch.int64, device=self._device())
        carry = self.__dict__.pop('_carry', None)
        if carry is not None and carry['env_id'] == self._es.env_id and carry['n'] == self._es.n:
            # state restored from a snapshot: the random streams continue where they stopped
            self._t_dev.copy_(carry['t'])
            self._es.episode.copy_(carry['episode'])
            if carry['norm_stats'] is not None:
                self._es.norm_stats = carry['norm_stats'].to(self._device())

    def _ensure_buffers(self, rows):
        if self._buf is None or self._buf.T < rows or self._buf.obs.device != self._device():
            old = self._buf
            self._buf = ops.RolloutBuffers(self._env.env_id, rows, self._n_envs, self._device(),
                                           want_mean=self._keep_agent_infos,
                                           want_success=self._env.has_success)
            return old
        return None

    def _update(self, agent_update, env_update):
        if agent_update is not None:
            if isinstance(agent_update, dict):
                self._policy.set_param_values(agent_update)
            elif agent_update is not self._policy:
                self._policy.set_param_values(agent_update.get_param_values())
        if env_update is not None:
            if not hasattr(env_update, 'env_id'):
                raise TypeError('env_update must be a garage_b200 BatchedEnv; got {}'.format(type(env_update)))
            self._env = env_update
            self._es = None
            self._buf = None
            self._static_pack = None

    # -------------------------------------------------------------------- API --
    def obtain_samples(self, itr, num_samples, agent_update, env_update=None):
        """Collect at least `num_samples` transitions (reference
        local_sampler.py:134-166).  Returns a DeviceEpisodeBatch."""
        if self._groups is not None:
            return self._obtain_grouped(itr, num_samples, agent_update, env_update)
        if getattr(self._env if env_update is None else env_update, 'env_id', None) == 'gridworld':
            return self._obtain_gridworld(num_samples, agent_update, env_update)
        self._update(agent_update, env_update)
        env, N = self._env, self._n_envs
        L = self._max_len()
        if L == 0 and getattr(env, 'never_terminates', False):
            raise ValueError('an env that never terminates needs a finite max_episode_length')
        self._ensure_state()
        prm = env.params_array()
        pol = self._policy
        # normalize(env): the wrapper's arithmetic runs inside the kernels; its running estimates
        # stay in self._es.norm_stats across calls, as each worker's env copy keeps them
        norm = env.norm_config() if hasattr(env, 'norm_config') else None
        # update_agent -> every env restarts from reset() (vec_worker.py:107-137)
        ops.env_reset(self._es, prm, self._seed, self._env_id_offset, norm=norm)
        static = bool(getattr(env, 'never_terminates', False)) and L > 0
        if static:
            rows = int(math.ceil(num_samples / float(N * L))) * L
        else:
            rows = int(math.ceil(num_samples / float(N)))
            if L:
                rows = int(math.ceil(rows / float(L))) * L
        rows = max(rows, 1)
        self._ensure_buffers(rows)
        done_rows = 0
        cached = key = None
        while True:
            ops.rollout(self._es, prm, L, pol.flat, pol.mlp_shape.H, pol._module.min_log_std,
                        self._max_log_std(), self._seed, self._env_id_offset, 0,
                        self._buf, done_rows, rows - done_rows, t_counter_dev=self._t_dev, norm=norm)
            ops.counter_add(self._t_dev, rows - done_rows)
            done_rows = rows
            if static:
                # never-terminating env: every episode is exactly L steps, so the episode structure
                # (row statistics, packed row of every step) depends on the shape alone -- planned
                # once per shape and reused; only the field data is packed every call
                t_star, n_eps, n_steps = rows - 1, N * (rows // L), N * rows
                key = (env.env_id, rows, N, L, int(num_samples), str(self._device()))
                cached = self._static_pack if self._static_pack and self._static_pack[0] == key else None
                ws = cached[1] if cached else self._plan(rows, num_samples, sync=False)[0]
                break
            ws, (t_star, n_eps, n_steps, reached) = self._plan(rows, num_samples, sync=True)
            if reached:
                break
            # not enough completed samples yet: keep the envs running (local_sampler.py:158-166)
            more = max(L if L else 64, rows // 2)
            old = self._buf
            self._buf = None
            self._ensure_buffers(rows + more)
            for name in ('obs', 'act', 'rew', 'logp', 'step_type', 'mean', 'success', 'last_obs', 'ep_len'):
                src = getattr(old, name)
                if src is not None:
                    getattr(self._buf, name)[:rows].copy_(src[:rows])
            rows += more
        batch = self._pack(ws, rows, t_star, n_eps, n_steps, static,
                           row_map=cached[2] if static and cached else None)
        if static and not cached:
            self._static_pack = (key, ws, batch.row_map)
        self.total_env_steps += n_steps
        return batch

    def _plan(self, rows, num_samples, sync):
        b = self._buf
        return ops.pack_plan(b.step_type[:rows], b.ep_len[:rows], rows, num_samples, sync=sync)

    def _pack(self, ws, rows, t_star, n_eps, n_steps, static, row_map=None):
        b, env, pol = self._buf, self._env, self._policy
        lengths, ep_off, last_obs, ep_env, ep_end = ops.pack_index(
            b.step_type, b.ep_len, b.last_obs, ws, rows, t_star, n_eps)
        fields = [b.obs, b.act, b.rew, b.logp, b.step_type]
        if b.mean is not None:
            fields.append(b.mean)
        if b.success is not None:
            fields.append(b.success)
        if row_map is None:
            row_map = ops.pack_rowmap(b.step_type, b.ep_len, ep_off, t_star)
        packed = ops.pack_fields(b.step_type, b.ep_len, ep_off, t_star, n_steps, fields, row_map=row_map)
        dev = DeviceEpisodeBatch._Dev()
        dev.observations, dev.actions, dev.rewards, dev.logp, dev.step_types = packed[:5]
        k = 5
        dev.mean = None
        if b.mean is not None:
            dev.mean = packed[k]
            k += 1
        dev.success = packed[k] if b.success is not None else None
        dev.last_observations, dev.lengths = last_obs, lengths
        dev.ep_env, dev.ep_end = ep_env, ep_end
        return DeviceEpisodeBatch(
            env_spec=env.spec, dev=dev, n_eps=n_eps, n_steps=n_steps,
            log_std=_LazyLogStd(pol._module), tm=b, ep_off=ep_off, t_star=t_star, row_map=row_map,
            episode_info_consts=env.episode_info(), env_info_consts=env.env_info_consts(),
            lengths_host=np.full(n_eps, self._max_len(), dtype=np.int64) if static else None)

    def _obtain_gridworld(self, num_samples, agent_update, env_update):
        """GridWorldEnv + ScriptedPolicy: the VecWorker loop as one integer kernel
        (b2g_gridworld_rollout), packed by the same pack kernels as every other batch."""
        if env_update is not None:
            self._env, self._gs = env_update, None
        if agent_update is not None and hasattr(agent_update, 'action_table'):
            self._policy = agent_update
        env, pol, N = self._env, self._policy, self._n_envs
        if not hasattr(pol, 'action_table'):
            raise NotImplementedError('GridWorldEnv has discrete actions: the device path runs it with a '
                                      'ScriptedPolicy (garage_b200.np.policies)')
        dev = self._device()
        gs = self._gs
        if gs is None or gs.n != N:
            gs = self._gs = env.grid_state(N, dev)
        gs.reset()  # update_agent -> every env restarts (vec_worker.py:107-137)
        table = torch.as_tensor(pol.action_table(gs.cells.numel())).to(dev)
        L = self._max_len()
        rows = int(math.ceil(num_samples / float(N)))
        if L:
            rows = int(math.ceil(rows / float(L))) * L
        rows = max(rows, 1)
        buf = ops.GridBuffers(rows, N, dev)
        done_rows = 0
        while True:
            ops.gridworld_rollout(gs, L, table, buf, done_rows, rows - done_rows)
            done_rows = rows
            ws, (t_star, n_eps, n_steps, reached) = ops.pack_plan(buf.step_type[:rows], buf.ep_len[:rows], rows,
                                                                  num_samples, sync=True)
            if reached:
                break
            if n_eps == 0 and rows >= 1 << 20:
                raise RuntimeError('no episode completes: the scripted walk never ends and there is no '
                                   'max_episode_length')
            more = max(L if L else 64, rows // 2)
            old, buf = buf, ops.GridBuffers(rows + more, N, dev)
            for name in ('obs', 'act', 'rew', 'step_type', 'last_obs', 'ep_len'):
                getattr(buf, name)[:rows].copy_(getattr(old, name)[:rows])
            rows += more
        lengths, ep_off, last_obs, ep_env, ep_end = ops.pack_index(
            buf.step_type, buf.ep_len, buf.last_obs.view(torch.float32), ws, rows, t_star, n_eps)
        row_map = ops.pack_rowmap(buf.step_type, buf.ep_len, ep_off, t_star)
        packed = ops.pack_fields(buf.step_type, buf.ep_len, ep_off, t_star, n_steps,
                                 [buf.obs, buf.act, buf.rew, buf.step_type], row_map=row_map)
        d = DeviceEpisodeBatch._Dev()
        d.observations, d.actions, d.rewards, d.step_types = packed
        d.logp = d.mean = d.success = None
        d.last_observations, d.lengths = last_obs.view(torch.int32), lengths
        d.ep_env, d.ep_end = ep_env, ep_end
        self.total_env_steps += n_steps
        return DeviceEpisodeBatch(env_spec=env.spec, dev=d, n_eps=n_eps, n_steps=n_steps, log_std=0.0,
                                  tm=buf, ep_off=ep_off, t_star=t_star, row_map=row_map)

    def _obtain_grouped(self, itr, num_samples, agent_update, env_update):
        """Per-worker agents / envs: the reference's loop (local_sampler.py:156-166) visits the workers
        round-robin, takes whatever episodes one `rollout()` of a worker completes and stops as soon as the
        total reaches `num_samples`.  A worker never contributes more than `num_samples` steps, so every group
        collects that many on the device, and the visiting order is replayed on the host over the groups'
        completion rows (one `VecWorker.rollout()` = the rows up to the next one in which an episode ends).
        Returns a host `EpisodeBatch` (episode / env infos differ per worker): a compatibility path, not the
        fast one."""
        from garage_b200._dtypes import EpisodeBatch
        n = len(self._groups)
        a_ups, e_ups = self._per_worker(agent_update, n), self._per_worker(env_update, n)
        rounds = []
        for g, au, eu in zip(self._groups, a_ups, e_ups):
            before = g.total_env_steps
            b = g.obtain_samples(itr, num_samples, au, eu)
            g.total_env_steps = before
            ends = b.dev.ep_end.cpu().numpy()
            eps = b.to_host().split()
            per_row = []
            for e, row in zip(eps, ends):  # completion order: a rollout() is a run of equal rows
                if not per_row or per_row[-1][0] != row:
                    per_row.append((row, []))
                per_row[-1][1].append(e)
            rounds.append([r[1] for r in per_row])
        out, total, r = [], 0, 0
        while True:
            for g, rr in zip(self._groups, rounds):
                for e in rr[r]:
                    out.append(e)
                    total += int(e.lengths.sum())
                    g.total_env_steps += int(e.lengths.sum())
                if total >= num_samples:
                    self.total_env_steps += total
                    return EpisodeBatch.concatenate(*out)
            r += 1

    def obtain_exact_episodes(self, n_eps_per_worker, agent_update, env_update=None):
        """Exactly `n_eps_per_worker` whole episodes from EVERY env copy, env-major order (reference
        local_sampler.py:168-200 with one env per worker): every env runs from `reset()` for
        n_eps_per_worker * max_episode_length steps, all episodes completed by then are packed and
        the first n_eps_per_worker of each env are kept."""
        k = int(n_eps_per_worker)
        if k <= 0:
            raise ValueError('n_eps_per_worker must be positive')
        if self._groups is not None:  # worker-major, like the reference's loop over workers
            from garage_b200._dtypes import EpisodeBatch
            n = len(self._groups)
            parts = [g.obtain_exact_episodes(k, au, eu).to_host() for g, au, eu in
                     zip(self._groups, self._per_worker(agent_update, n), self._per_worker(env_update, n))]
            self.total_env_steps += sum(int(p.lengths.sum()) for p in parts)
            return EpisodeBatch.concatenate(*parts)
        self._update(agent_update, env_update)
        env, N, L = self._env, self._n_envs, self._max_len()
        if L == 0:
            raise ValueError('obtain_exact_episodes needs a finite max_episode_length')
        self._ensure_state()
        prm, pol = env.params_array(), self._policy
        norm = env.norm_config() if hasattr(env, 'norm_config') else None
        ops.env_reset(self._es, prm, self._seed, self._env_id_offset, norm=norm)
        rows = k * L
        self._ensure_buffers(rows)
        ops.rollout(self._es, prm, L, pol.flat, pol.mlp_shape.H, pol._module.min_log_std,
                    self._max_log_std(), self._seed, self._env_id_offset, 0, self._buf, 0, rows,
                    t_counter_dev=self._t_dev, norm=norm)
        ops.counter_add(self._t_dev, rows)
        ws, (t_star, n_eps, n_steps, _) = self._plan(rows, N * rows + 1, sync=True)
        full = self._pack(ws, rows, t_star, n_eps, n_steps, static=False)
        batch = self._select_first(full, k)
        self.total_env_steps += batch.n_steps
        return batch

    def _select_first(self, full, k):
        """The first k episodes of every env out of a packed batch, reordered env-major.
        Index arithmetic on the device (not on the training path: evaluation batches)."""
        d, N = full.dev, self._n_envs
        n_all = full.n_episodes
        ep_env = d.ep_env.to(torch.int64)
        # episodes are in completion order, so a stable sort by env keeps each env's own order
        order = torch.sort(ep_env, stable=True).indices
        first_of_env = torch.searchsorted(ep_env[order], torch.arange(N, device=ep_env.device))
        rank = torch.arange(n_all, device=ep_env.device) - first_of_env[ep_env[order]]
        keep = order[rank < k]  # env-major, completion order inside an env
        if keep.numel() != N * k:
            raise RuntimeError('an env completed fewer than {} episodes'.format(k))
        off_all = torch.cumsum(d.lengths, 0) - d.lengths
        lengths = d.lengths[keep]
        new_off = torch.cumsum(lengths, 0) - lengths
        n_steps = int(lengths.sum().item())
        ep_of_row = torch.repeat_interleave(torch.arange(keep.numel(), device=keep.device), lengths,
                                            output_size=n_steps)
        rows = off_all[keep][ep_of_row] + (torch.arange(n_steps, device=keep.device) - new_off[ep_of_row])
        dev = DeviceEpisodeBatch._Dev()
        for name in ('observations', 'actions', 'rewards', 'logp', 'step_types', 'mean', 'success'):
            t = getattr(d, name)
            setattr(dev, name, None if t is None else t[rows].contiguous())
        dev.last_observations, dev.lengths = d.last_observations[keep].contiguous(), lengths
        dev.ep_env, dev.ep_end = d.ep_env[keep], d.ep_end[keep]
        return DeviceEpisodeBatch(
            env_spec=full.env_spec, dev=dev, n_eps=keep.numel(), n_steps=n_steps, log_std=full._log_std,
            episode_info_consts=full._episode_info_consts, env_info_consts=full._env_info_consts)

    def shutdown_worker(self):
        """Nothing to shut down: there are no worker processes."""
        for g in self._groups or ():
            g.shutdown_worker()
        self._es = None
        self._buf = None
        self._static_pack = None

    def __getstate__(self):
        """cloudpickle-able: device buffers are rebuilt lazily; what identifies the position in the
        random streams (Philox step counter, per-env episode counters) and the running estimates
        of normalize(env) travel as host tensors, so a resumed run continues the same sequence."""
        state = self.__dict__.copy()
        es = self._es
        if es is not None and self._t_dev is not None:
            state['_carry'] = dict(env_id=es.env_id, n=es.n, t=self._t_dev.cpu(), episode=es.episode.cpu(),
                                   norm_stats=None if es.norm_stats is None else es.norm_stats.cpu())
        state['_es'] = None
        state['_buf'] = None
        state['_gs'] = None
        state['_static_pack'] = None
        state['_t_dev'] = None
        return state


LocalSampler = GPUSampler
