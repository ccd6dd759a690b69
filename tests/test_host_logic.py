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
