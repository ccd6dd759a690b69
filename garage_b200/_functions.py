"""Drop-ins for `garage._functions` on the on-policy path: `log_performance`,
`log_multitask_performance` and `obtain_evaluation_episodes` (reference
_functions.py:143-275).  The per-episode reductions (undiscounted / discounted returns,
termination, success) run in `b2g_episode_stats` on the packed episode-major arrays -- the
device copy of a `DeviceEpisodeBatch` directly, a host `EpisodeBatch` after one upload of its
rewards, step types and lengths -- instead of the reference's per-episode Python loop."""
import collections

import numpy as np
import torch

from garage_b200 import ops
from garage_b200._dtypes import DeviceEpisodeBatch, EpisodeBatch, StepType
from garage_b200.tabular import tabular


def _device():
    return torch.device('cuda', torch.cuda.current_device())


def _episode_stats(batch, discount):
    """(per_episode [4, n_eps] float64 numpy, summary [7] float64 numpy) of a batch."""
    if isinstance(batch, DeviceEpisodeBatch) and batch.dev is not None:
        d = batch.dev
        per, out = ops.episode_stats(d.rewards, discount, step_types=d.step_types, success=d.success,
                                     lengths=d.lengths)
    else:
        dev = _device()
        lengths = torch.as_tensor(np.asarray(batch.lengths, dtype=np.int64)).to(dev)
        rewards = torch.as_tensor(np.ascontiguousarray(batch.rewards, dtype=np.float64)).to(dev)
        codes = np.fromiter((int(s) for s in batch.step_types), dtype=np.uint8, count=len(batch.step_types))
        success = None
        if 'success' in batch.env_infos:
            success = torch.as_tensor(np.ascontiguousarray(batch.env_infos['success']).astype(np.uint8)).to(dev)
        per, out = ops.episode_stats(rewards, discount, step_types=torch.as_tensor(codes).to(dev),
                                     success=success, lengths=lengths)
    both = torch.cat([per.reshape(-1), out]).cpu().numpy()  # one D2H read
    return both[:-7].reshape(4, -1), both[-7:]


def log_performance(itr, batch, discount, prefix='Evaluation'):
    """Records the reference's keys (_functions.py:262-273) under `prefix/` and returns the
    undiscounted returns, one per episode (numpy float64 array)."""
    if len(batch.lengths) == 0:
        raise ValueError('log_performance needs at least one episode')
    per, s = _episode_stats(batch, discount)
    with tabular.prefix(prefix + '/'):
        tabular.record('Iteration', itr)
        tabular.record('NumEpisodes', per.shape[1])
        tabular.record('AverageDiscountedReturn', s[0])
        tabular.record('AverageReturn', s[1])
        tabular.record('StdReturn', s[2])
        tabular.record('MaxReturn', s[3])
        tabular.record('MinReturn', s[4])
        tabular.record('TerminationRate', s[5])
        if not np.isnan(s[6]):
            tabular.record('SuccessRate', s[6])
    return per[0].copy()


_NAN_KEYS = ('AverageDiscountedReturn', 'AverageReturn', 'StdReturn', 'MaxReturn', 'MinReturn',
             'TerminationRate', 'SuccessRate')


def log_multitask_performance(itr, batch, discount, name_map=None):
    """Per-task `log_performance` plus the 'Average' block (_functions.py:170-230): episodes
    are grouped by `env_infos['task_name']`, else `env_infos['task_id']` through `name_map`."""
    groups = collections.defaultdict(list)
    for eps in batch.split():
        name = '__unnamed_task__'
        if 'task_name' in eps.env_infos:
            name = eps.env_infos['task_name'][0]
        elif 'task_id' in eps.env_infos:
            task_id = eps.env_infos['task_id'][0]
            name = (name_map or {}).get(task_id, 'Task #{}'.format(task_id))
        groups[name].append(eps)
    for name in (groups.keys() if name_map is None else name_map.values()):
        if name in groups:
            log_performance(itr, EpisodeBatch.concatenate(*groups[name]), discount, prefix=name)
        else:
            with tabular.prefix(name + '/'):
                tabular.record('Iteration', itr)
                tabular.record('NumEpisodes', 0)
                for key in _NAN_KEYS:
                    tabular.record(key, np.nan)
    return log_performance(itr, batch, discount=discount, prefix='Average')


def obtain_evaluation_episodes(policy, env, max_episode_length=1000, num_eps=100, deterministic=True):
    """`num_eps` evaluation episodes, each from a fresh `reset()` (_functions.py:143-167).  The
    reference runs them one after the other; here `num_eps` device environments run side by side
    and each contributes its first episode.  `deterministic=True` acts with the mean of the
    action distribution (np/_functions.py rollout(): `agent_info['mean']`)."""
    from garage_b200.sampler import GPUSampler
    sampler = GPUSampler(policy, env, max_episode_length=max_episode_length, n_envs=int(num_eps),
                         deterministic=deterministic)
    return sampler.obtain_exact_episodes(1, policy)
