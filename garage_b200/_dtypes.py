"""Batch data types of the on-policy path: `StepType` and `EpisodeBatch`.

API mirror of garage `_dtypes.py` (StepType :15-68, EpisodeBatch :454-977) so
that code written against the reference keeps working; the storage model is
new: `DeviceEpisodeBatch` keeps everything on the GPU (time-major rollout
buffers + episode-major packed views made by libb2g's pack kernels) and only
materialises NumPy arrays when an attribute is read.
"""
import enum

import numpy as np
import torch


class StepType(enum.IntEnum):
    """Status of a time step inside an episode (reference `_dtypes.py:15-42`)."""
    FIRST = 0
    MID = 1
    TERMINAL = 2
    TIMEOUT = 3

    @classmethod
    def get_step_type(cls, step_cnt, max_episode_length, done):
        """Reference `_dtypes.py:44-68`: timeout wins over done; a one-step
        terminal episode is TERMINAL, not FIRST."""
        if max_episode_length is not None and step_cnt >= max_episode_length:
            return cls.TIMEOUT
        if done:
            return cls.TERMINAL
        if step_cnt == 1:
            return cls.FIRST
        if step_cnt < 1:
            raise ValueError('Expect step_cnt to be >= 1, but got {} instead. Did you forget '
                             'to call `reset()`?'.format(step_cnt))
        return cls.MID


_STEP_TYPES = np.array(list(StepType), dtype=object)


def step_types_from_codes(codes):
    """uint8 codes -> the reference's object array of StepType members."""
    return _STEP_TYPES[np.asarray(codes, dtype=np.intp)].astype(StepType)


def _soft_contains(space, element):
    if space.contains(element):
        return True
    if hasattr(space, 'flat_dim'):
        return space.flat_dim == np.prod(np.shape(element))
    return False


def _slice_nested(d, start, stop):
    return {k: (_slice_nested(v, start, stop) if isinstance(v, dict) else v[start:stop])
            for k, v in d.items()}


def _concat_nested(dicts):
    first = dicts[0]
    return {k: (_concat_nested([d[k] for d in dicts]) if isinstance(first[k], dict) else
                np.concatenate([d[k] for d in dicts])) for k in first}


def pad_batch_array(array, lengths, max_length=None):
    """Packed [sum(T), ...] -> padded [N, max_length, ...] (reference
    np/_functions.py:375-406, vectorised with one fancy-index scatter)."""
    lengths = np.asarray(lengths, dtype=np.int64)
    assert array.shape[0] == int(lengths.sum())
    longest = int(lengths.max()) if len(lengths) else 0
    if max_length is None:
        max_length = longest
    elif max_length < longest:
        import warnings
        warnings.warn('Creating a padded array with longer length than requested')
        max_length = longest
    out = np.zeros((len(lengths), max_length) + array.shape[1:], dtype=array.dtype)
    if array.shape[0]:
        rows = np.repeat(np.arange(len(lengths)), lengths)
        starts = np.repeat(np.cumsum(lengths) - lengths, lengths)
        cols = np.arange(array.shape[0]) - starts
        out[rows, cols] = array
    return out


class EpisodeBatch:
    """A batch of whole episodes, episode-major and packed (reference
    `_dtypes.py:454-977`): per-step fields have shape `(sum(lengths), ...)`,
    per-episode fields `(N, ...)`.  Raises ValueError on malformed input like
    the reference (`:528-589`, `check_timestep_batch :1001-1083`)."""

    def __init__(self, env_spec, episode_infos, observations, last_observations, actions,
                 rewards, env_infos, agent_infos, step_types, lengths):
        lengths = np.asarray(lengths) if not isinstance(lengths, np.ndarray) else lengths
        if lengths.ndim != 1:
            raise ValueError('lengths has shape {} but must be a tensor of shape (N,)'.format(lengths.shape))
        if lengths.dtype.kind not in 'ui':
            raise ValueError('lengths has dtype {}, but must have an integer dtype'.format(lengths.dtype))
        n_eps = len(lengths)
        n_steps = int(lengths.sum())
        for key, val in episode_infos.items():
            if not isinstance(val, np.ndarray):
                raise ValueError('Entry {!r} in episode_infos is of type {!r} but must be of type '
                                 '{!r}'.format(key, type(val), np.ndarray))
            if val.shape[0] != n_eps:
                raise ValueError('Entry {!r} in episode_infos has batch size {}, but must have batch '
                                 'size {} to match the number of episodes'.format(key, val.shape[0], n_eps))
        if not isinstance(last_observations, np.ndarray):
            raise ValueError('last_observations is not of type {!r}'.format(np.ndarray))
        if last_observations.shape[0] != n_eps:
            raise ValueError('last_observations has batch size {} but must have batch size {} to '
                             'match the number of episodes'.format(last_observations.shape[0], n_eps))
        for name, val in (('observations', observations), ('actions', actions),
                          ('rewards', rewards), ('step_types', step_types)):
            if not isinstance(val, np.ndarray):
                raise ValueError('{} is not of type {!r}'.format(name, np.ndarray))
            if val.shape[0] != n_steps:
                raise ValueError('{} has batch size {}, but must have batch size {} to match '
                                 'sum(lengths)'.format(name, val.shape[0], n_steps))
        if n_steps:
            if not _soft_contains(env_spec.observation_space, observations[0]):
                raise ValueError('Each observation has shape {} but must match the observation_space '
                                 '{}'.format(observations[0].shape, env_spec.observation_space))
            if not _soft_contains(env_spec.observation_space, last_observations[0]):
                raise ValueError('last_observations must match the observation_space')
            if not _soft_contains(env_spec.action_space, actions[0]):
                raise ValueError('Each action has shape {} but must match the action_space {}'.format(
                    actions[0].shape, env_spec.action_space))
        if rewards.shape != (n_steps, ):
            raise ValueError('rewards has shape {} but must have batch size {}'.format(rewards.shape, n_steps))
        if step_types.shape != (n_steps, ):
            raise ValueError('step_types has shape {} but must have batch size {}'.format(step_types.shape, n_steps))
        if step_types.dtype != StepType:
            raise ValueError('step_types has dtype {} but must have dtype StepType'.format(step_types.dtype))
        for field, d in (('agent_infos', agent_infos), ('env_infos', env_infos)):
            for key, val in d.items():
                if not isinstance(val, (np.ndarray, dict)):
                    raise ValueError('Entry {!r} in {} is of type {} but must be {!r} or dict'.format(
                        key, field, type(val), np.ndarray))
                if hasattr(val, 'shape') and val.shape[0] != n_steps:
                    raise ValueError('Entry {!r} in {} has batch size {} but must have batch size '
                                     '{}'.format(key, field, val.shape[0], n_steps))
        self.env_spec = env_spec
        self.episode_infos_by_episode = episode_infos
        self.observations = observations
        self.last_observations = last_observations
        self.actions = actions
        self.rewards = rewards
        self.env_infos = env_infos
        self.agent_infos = agent_infos
        self.step_types = step_types
        self.lengths = lengths

    # -- construction helpers -------------------------------------------------
    @classmethod
    def concatenate(cls, *batches):
        """Reference `_dtypes.py:591-632`."""
        b0 = batches[0]
        return cls(env_spec=b0.env_spec,
                   episode_infos=_concat_nested([b.episode_infos_by_episode for b in batches]),
                   observations=np.concatenate([b.observations for b in batches]),
                   last_observations=np.concatenate([b.last_observations for b in batches]),
                   actions=np.concatenate([b.actions for b in batches]),
                   rewards=np.concatenate([b.rewards for b in batches]),
                   env_infos=_concat_nested([b.env_infos for b in batches]),
                   agent_infos=_concat_nested([b.agent_infos for b in batches]),
                   step_types=np.concatenate([b.step_types for b in batches]),
                   lengths=np.concatenate([b.lengths for b in batches]))

    @classmethod
    def from_list(cls, env_spec, paths):
        """EpisodeBatch from a list of per-episode dicts (reference `:730-808`).  Keys per path:
        observations (T or T+1 rows), optional next_observations, actions, rewards, step_types
        (or legacy `dones`), env_infos, agent_infos, episode_infos.  With T+1 observation rows the
        last row is the episode's final observation; otherwise next_observations[-1] is, and
        failing that the last observation is duplicated (the reference's lenient fallback)."""
        lengths = np.asarray([len(p['rewards']) for p in paths])
        if all(len(p['observations']) == n + 1 for p, n in zip(paths, lengths)):
            last_observations = np.asarray([p['observations'][-1] for p in paths])
            observations = np.concatenate([p['observations'][:-1] for p in paths])
        else:
            observations = np.concatenate([p['observations'] for p in paths])
            key = 'next_observations' if paths[0].get('next_observations') is not None else 'observations'
            last_observations = np.asarray([p[key][-1] for p in paths])

        def stacked(name):
            return np.concatenate([np.asarray(p[name]) for p in paths])

        if 'step_types' in paths[0]:
            step_types = stacked('step_types')
        elif 'dones' in paths[0]:
            step_types = np.array([StepType.TERMINAL if d else StepType.MID for d in stacked('dones')],
                                  dtype=StepType)
        else:
            raise KeyError('paths need step_types (or dones)')
        infos = [p.get('episode_infos', {}) for p in paths]
        episode_infos = {k: np.asarray([i[k] for i in infos]) for k in (infos[0] if infos else {})}
        return cls(env_spec=env_spec, episode_infos=episode_infos, observations=observations,
                   last_observations=last_observations, actions=stacked('actions'), rewards=stacked('rewards'),
                   env_infos=_concat_nested([p['env_infos'] for p in paths]),
                   agent_infos=_concat_nested([p['agent_infos'] for p in paths]),
                   step_types=step_types, lengths=lengths)

    def _episode_ranges(self):
        stops = np.cumsum(self.lengths)
        return zip(stops - self.lengths, stops)

    def split(self):
        """One single-episode EpisodeBatch per episode (reference `:648-674`)."""
        out = []
        for i, (start, stop) in enumerate(self._episode_ranges()):
            out.append(EpisodeBatch(
                env_spec=self.env_spec,
                episode_infos=_slice_nested(self.episode_infos_by_episode, i, i + 1),
                observations=self.observations[start:stop],
                last_observations=np.asarray([self.last_observations[i]]),
                actions=self.actions[start:stop], rewards=self.rewards[start:stop],
                env_infos=_slice_nested(self.env_infos, start, stop),
                agent_infos=_slice_nested(self.agent_infos, start, stop),
                step_types=self.step_types[start:stop], lengths=np.asarray([self.lengths[i]])))
        return out

    def to_list(self):
        """List of per-episode dicts (reference `:676-737`)."""
        out = []
        nxt = self.next_observations
        for i, (start, stop) in enumerate(self._episode_ranges()):
            out.append(dict(
                episode_infos={k: v[i:i + 1] for k, v in self.episode_infos.items()},
                observations=self.observations[start:stop], next_observations=nxt[start:stop],
                actions=self.actions[start:stop], rewards=self.rewards[start:stop],
                env_infos=_slice_nested(self.env_infos, start, stop),
                agent_infos=_slice_nested(self.agent_infos, start, stop),
                step_types=self.step_types[start:stop]))
        return out

    # -- derived views ----------------------------------------------------------
    @property
    def next_observations(self):
        """Reference `:810-827`, vectorised: shift by one inside each episode and
        put the episode's last observation at its final step."""
        nxt = np.empty_like(self.observations)
        nxt[:-1] = self.observations[1:]
        ends = np.cumsum(self.lengths) - 1
        nxt[ends] = self.last_observations
        return nxt

    @property
    def episode_infos(self):
        """Per-step view of the per-episode infos (reference `:829-850`)."""
        return {k: np.repeat(v, self.lengths, axis=0) for k, v in self.episode_infos_by_episode.items()}

    def _pad(self, arr):
        return pad_batch_array(arr, self.lengths, self.env_spec.max_episode_length)

    @property
    def padded_observations(self):
        return self._pad(self.observations)

    @property
    def padded_actions(self):
        return self._pad(self.actions)

    @property
    def padded_rewards(self):
        return self._pad(self.rewards)

    @property
    def padded_step_types(self):
        return self._pad(self.step_types)

    @property
    def padded_next_observations(self):
        return self._pad(self.next_observations)

    @property
    def padded_agent_infos(self):
        return {k: self._pad(v) for k, v in self.agent_infos.items()}

    @property
    def padded_env_infos(self):
        return {k: self._pad(v) for k, v in self.env_infos.items() if not isinstance(v, dict)}

    @property
    def valids(self):
        """[N, P] array of 1 where a step exists, 0 in the padding (reference `:960-977`)."""
        return self._pad(np.ones(int(self.lengths.sum())))

    @property
    def observations_list(self):
        return [self.observations[a:b] for a, b in self._episode_ranges()]

    @property
    def actions_list(self):
        return [self.actions[a:b] for a, b in self._episode_ranges()]


class DeviceEpisodeBatch(EpisodeBatch):
    """EpisodeBatch whose storage is GPU memory.

    Produced by `garage_b200.sampler.GPUSampler`; consumed directly (device
    pointers) by `garage_b200.torch.algos.VPG/PPO._train_once`.  Every
    reference attribute (`observations`, `rewards`, `step_types`, `padded_*`,
    `split()`, ...) works and is materialised on first access with one D2H
    copy, in the reference's dtypes (observations/actions float32, rewards
    float64, step_types object array of StepType, lengths int64).

    Device-side fields (torch CUDA tensors), episode-major packed:
      dev.observations [B,O], dev.actions [B,A], dev.rewards [B] f32,
      dev.logp [B] (log-prob under the rollout policy), dev.step_types [B] u8,
      dev.mean [B,A], dev.success [B] u8 or None, dev.last_observations [N,O],
      dev.lengths [N] int64;
    plus the time-major rollout buffers (`tm`) and the pack index needed to
    pack further per-step fields (advantages, returns) later.
    """

    class _Dev:
        pass

    def __init__(self, env_spec, dev, n_eps, n_steps, log_std, tm=None, ep_off=None, t_star=None,
                 episode_info_consts=None, env_info_consts=None, lengths_host=None, row_map=None):
        # no validation pass: the fields come from the pack kernels, whose
        # invariants the parity tests check bit-exactly
        self.env_spec = env_spec
        self.dev = dev
        self.n_episodes = int(n_eps)
        self.n_steps = int(n_steps)
        self._log_std = log_std  # float or lazy device snapshot
        self.tm = tm
        self.ep_off = ep_off
        self.t_star = t_star
        self.row_map = row_map  # int32 [t_star+1, N]: packed row of every rollout step (ops.pack_rowmap)
        self._episode_info_consts = dict(episode_info_consts or {})
        self._env_info_consts = dict(env_info_consts or {})
        self._cache = {}
        if lengths_host is not None:
            self._cache['lengths'] = np.asarray(lengths_host, dtype=np.int64)

    @property
    def log_std(self):
        if not isinstance(self._log_std, float):
            self._log_std = float(self._log_std)
        return self._log_std

    def _get(self, name, make):
        if name not in self._cache:
            self._cache[name] = make()
        return self._cache[name]

    @property
    def lengths(self):
        return self._get('lengths', lambda: self._host('lengths'))

    def _host_field(self, name):
        """Discrete spaces (GridWorld): the device holds int32 [n, 1]; the reference's batches hold
        int64 (n,) (vec_worker.py:206-219 over python ints)."""
        arr = self._host(name)
        return arr.reshape(-1).astype(np.int64) if arr.dtype == np.int32 else arr

    @property
    def observations(self):
        return self._get('observations', lambda: self._host_field('observations'))

    @property
    def last_observations(self):
        return self._get('last_observations', lambda: self._host_field('last_observations'))

    @property
    def actions(self):
        return self._get('actions', lambda: self._host_field('actions'))

    @property
    def rewards(self):
        # python floats in the reference -> float64 (Q9)
        return self._get('rewards', lambda: self._host('rewards').astype(np.float64, copy=False))

    @property
    def step_types(self):
        return self._get('step_types', lambda: step_types_from_codes(self._host('step_types')))

    @property
    def agent_infos(self):
        def make():
            if self.dev.mean is None:  # scripted / agent-info-free rollouts
                return {}
            mean = self._host('mean')
            return dict(mean=mean, log_std=np.full_like(mean, np.float32(self.log_std)))
        return self._get('agent_infos', make)

    @property
    def env_infos(self):
        def make():
            out = {}
            if self.dev.success is not None:
                out['success'] = self.dev.success.cpu().numpy().astype(bool)
            for k, v in self._env_info_consts.items():
                out[k] = ({kk: np.broadcast_to(np.asarray(vv), (self.n_steps, ) + np.shape(vv)).copy()
                           for kk, vv in v.items()} if isinstance(v, dict) else
                          np.broadcast_to(np.asarray(v), (self.n_steps, ) + np.shape(v)).copy())
            return out
        return self._get('env_infos', make)

    @property
    def episode_infos_by_episode(self):
        return self._get('episode_infos_by_episode', lambda: {
            k: np.broadcast_to(np.asarray(v), (self.n_episodes, ) + np.shape(v)).copy()
            for k, v in self._episode_info_consts.items()})

    _pinned_pool = {}

    def prefetch_host(self, fields=('observations', 'actions', 'rewards', 'step_types', 'mean',
                                    'last_observations', 'lengths')):
        """Start asynchronous D2H copies of the packed fields into pinned host buffers on a
        side stream (they overlap with whatever is enqueued next, e.g. `_train_once`).
        The NumPy attributes then wait for the copy instead of doing a pageable `.cpu()`.
        The pinned buffers are pooled: the arrays are valid until the next `prefetch_host`
        of a batch of the same shape.  Returns the number of bytes copied."""
        cur = torch.cuda.current_stream(self.dev.observations.device)
        side = DeviceEpisodeBatch._pinned_pool.setdefault(('stream', cur.device), torch.cuda.Stream(cur.device))
        side.wait_stream(cur)
        self._pinned = {}
        nbytes = 0
        with torch.cuda.stream(side):
            for name in fields:
                t = getattr(self.dev, name, None)
                if t is None:
                    continue
                if name == 'rewards' and t.dtype == torch.float32:
                    # the reference's rewards are float64 (Q9): widen on the device, so the host
                    # attribute is a view of the pinned buffer instead of an astype() pass
                    t = t.double()
                key = (name, tuple(t.shape), t.dtype)
                buf = DeviceEpisodeBatch._pinned_pool.get(key)
                if buf is None:
                    buf = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
                    DeviceEpisodeBatch._pinned_pool[key] = buf
                buf.copy_(t, non_blocking=True)
                t.record_stream(side)
                self._pinned[name] = buf
                nbytes += t.numel() * t.element_size()
            self._pinned_event = torch.cuda.Event()
            self._pinned_event.record(side)
        return nbytes

    def _host(self, name):
        """NumPy view of a prefetched field, or a synchronous copy."""
        pinned = getattr(self, '_pinned', None)
        if pinned and name in pinned:
            if self._pinned_event is not None:  # None: copies replayed by a CUDA graph, the caller synchronised
                self._pinned_event.synchronize()
            return pinned[name].numpy()
        return getattr(self.dev, name).cpu().numpy()

    def to_host(self):
        """A plain, validated `EpisodeBatch` copy (all fields on the host)."""
        return EpisodeBatch(env_spec=self.env_spec, episode_infos=self.episode_infos_by_episode,
                            observations=self.observations, last_observations=self.last_observations,
                            actions=self.actions, rewards=self.rewards, env_infos=self.env_infos,
                            agent_infos=self.agent_infos, step_types=self.step_types,
                            lengths=self.lengths)

    def __getstate__(self):  # cloudpickle: ship the host copy, never device handles
        self.to_host()
        return dict(env_spec=self.env_spec, cache=self._cache, n_eps=self.n_episodes,
                    n_steps=self.n_steps, log_std=self.log_std,
                    ei=self._episode_info_consts, vi=self._env_info_consts)

    def __setstate__(self, s):
        self.env_spec, self._cache = s['env_spec'], s['cache']
        self.n_episodes, self.n_steps, self._log_std = s['n_eps'], s['n_steps'], s['log_std']
        self._episode_info_consts, self._env_info_consts = s['ei'], s['vi']
        self.dev = self.tm = self.ep_off = self.t_star = None
