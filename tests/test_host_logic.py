"""Host-side pieces of the GridWorld / evaluation rows that need no GPU: Discrete space,
ScriptedPolicy, GridWorldEnv construction (maps, start state, cell codes), GymEnv id shim."""
import numpy as np
import pytest

from garage_b200.envs import GridWorldEnv, GymEnv
from garage_b200.envs.grid_world_env import MAPS
from garage_b200.np.policies import ScriptedPolicy
from garage_b200.spaces import Box, Discrete


def test_discrete_space():
    d = Discrete(5)
    assert d.flat_dim == 5 and d.n == 5 and d == Discrete(5) and d != Discrete(4) and d != Box(0, 1, (5, ))
    assert d.contains(3) and d.contains(np.int64(0)) and not d.contains(5) and not d.contains(-1)
    assert not d.contains(1.0) and not d.contains(np.array([1]))
    np.testing.assert_array_equal(d.flatten(2), [0, 0, 1, 0, 0])
    assert d.unflatten(d.flatten(4)) == 4
    np.testing.assert_array_equal(d.unflatten_n(d.flatten_n([0, 3, 1])), [0, 3, 1])
    assert 0 <= d.sample() < 5 and hash(d) == hash(Discrete(5))


def test_scripted_policy():
    p = ScriptedPolicy(scripted_actions=[2, 2, 1, 0], agent_env_infos=None)
    assert p.get_action(2) == (1, {}) and p.get_actions([0, 3]) == ([2, 0], {})
    assert p.get_param_values() == [] and p.action_table(4).dtype == np.int32
    q = ScriptedPolicy({0: 3, 1: 1}, agent_env_infos={0: {'k': 1}, 1: {'k': 2}})
    assert q.get_action(1) == (1, {'k': 2})
    np.testing.assert_array_equal(q.action_table(2), [3, 1])


@pytest.mark.parametrize('desc', sorted(MAPS))
def test_gridworld_construction(desc):
    env = GridWorldEnv(desc=desc, max_episode_length=7)
    rows = MAPS[desc]
    n = len(rows) * len(rows[0])
    assert env.spec.observation_space == Discrete(n) and env.spec.action_space == Discrete(4)
    assert env.spec.max_episode_length == 7 and env.render_modes == []
    flat = ''.join(rows)
    assert flat[env._start_state] == 'S'
    codes = {'S': 0, 'F': 0, 'W': 1, 'H': 2, 'G': 3}
    np.testing.assert_array_equal(env._cells, [codes[c] for c in flat])


def test_gridworld_custom_desc_and_errors():
    env = GridWorldEnv(desc=['S.x', '.oG'])
    np.testing.assert_array_equal(env._cells, [0, 0, 1, 0, 2, 3])
    with pytest.raises(NotImplementedError):
        GridWorldEnv(desc=['SQ', 'FG'])
    with pytest.raises(ValueError):
        GridWorldEnv(desc=['FF', 'FG'])
    with pytest.raises(KeyError):
        GridWorldEnv(desc='5x5')


def test_gym_env_shim():
    assert GymEnv('Pendulum-v0').spec.max_episode_length == 200
    assert GymEnv('CartPole-v0').spec.max_episode_length == 200
    assert GymEnv('CartPole-v1', max_episode_length=77).spec.max_episode_length == 77
    with pytest.raises(NotImplementedError):
        GymEnv('InvertedDoublePendulum-v2')
    with pytest.raises(NotImplementedError):
        GymEnv('Pendulum-v0', is_image=True)
    with pytest.raises(ValueError):
        GymEnv(object())


def _batch():
    from garage_b200 import EnvSpec, EpisodeBatch, StepType
    lengths = np.array([3, 1, 2])
    n = int(lengths.sum())
    rng = np.random.default_rng(0)
    st = np.array([StepType.FIRST, StepType.MID, StepType.TERMINAL, StepType.TIMEOUT, StepType.FIRST,
                   StepType.TIMEOUT], dtype=StepType)
    spec = EnvSpec(Box(-np.ones(3), np.ones(3)), Box(-np.ones(2), np.ones(2)), max_episode_length=3)
    return EpisodeBatch(
        spec, episode_infos={'goal': rng.normal(size=(3, 2))},
        observations=rng.uniform(-1, 1, size=(n, 3)).astype(np.float32),
        last_observations=rng.uniform(-1, 1, size=(3, 3)).astype(np.float32),
        actions=rng.uniform(-1, 1, size=(n, 2)).astype(np.float32), rewards=rng.normal(size=n),
        env_infos={'success': rng.random(n) < 0.5, 'task': {'id': np.arange(n)}},
        agent_infos={'mean': rng.normal(size=(n, 2)).astype(np.float32)}, step_types=st, lengths=lengths)


def test_episode_batch_to_list_from_list_round_trip():
    """Reference tests/garage/test_dtypes.py: to_list() -> from_list() reproduces the batch, with
    next_observations carrying each episode's final observation."""
    from garage_b200 import EpisodeBatch, StepType
    b = _batch()
    paths = b.to_list()
    assert [len(p['rewards']) for p in paths] == [3, 1, 2]
    np.testing.assert_array_equal(paths[0]['next_observations'][-1], b.last_observations[0])
    np.testing.assert_array_equal(paths[0]['next_observations'][:-1], b.observations[1:3])
    back = EpisodeBatch.from_list(b.env_spec, paths)
    for name in ('observations', 'last_observations', 'actions', 'rewards', 'lengths'):
        np.testing.assert_array_equal(getattr(back, name), getattr(b, name), err_msg=name)
    assert list(back.step_types) == list(b.step_types) and back.step_types.dtype == StepType
    np.testing.assert_array_equal(back.env_infos['task']['id'], np.arange(6))
    np.testing.assert_array_equal(back.agent_infos['mean'], b.agent_infos['mean'])
    assert back.episode_infos_by_episode['goal'].shape[0] == 3
    # T + 1 observation rows: the extra row is the final observation
    plus = [dict(p, observations=np.concatenate([p['observations'], p['next_observations'][-1:]])) for p in paths]
    for p in plus:
        p.pop('next_observations')
    again = EpisodeBatch.from_list(b.env_spec, plus)
    np.testing.assert_array_equal(again.observations, b.observations)
    np.testing.assert_array_equal(again.last_observations, b.last_observations)
    # legacy `dones`, no next_observations: last observation duplicated, TERMINAL/MID step types
    legacy = [{k: v for k, v in p.items() if k not in ('step_types', 'next_observations')} for p in paths]
    for p, n in zip(legacy, [3, 1, 2]):
        p['dones'] = np.array([False] * (n - 1) + [True])
    old = EpisodeBatch.from_list(b.env_spec, legacy)
    np.testing.assert_array_equal(old.last_observations[0], b.observations[2])
    assert [int(s) for s in old.step_types] == [1, 1, 2, 2, 1, 2]
    # split() / concatenate() round trip
    cat = EpisodeBatch.concatenate(*b.split())
    np.testing.assert_array_equal(cat.observations, b.observations)
    np.testing.assert_array_equal(cat.lengths, b.lengths)


def test_sampler_list_arguments_need_one_entry_per_worker():
    """worker_factory.py:68-95: a list of agents / envs must have length n_workers (checked without a GPU)."""
    from garage_b200.sampler import LocalSampler
    with pytest.raises(ValueError, match="Length of list doesn't match number of workers"):
        LocalSampler(agents=object(), envs=[object(), object(), object()], max_episode_length=10, n_workers=2)
    with pytest.raises(ValueError, match="Length of list doesn't match number of workers"):
        LocalSampler(agents=[object()], envs=object(), max_episode_length=10, n_workers=2)
