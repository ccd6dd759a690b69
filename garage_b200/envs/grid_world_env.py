"""GridWorldEnv on the device (reference envs/grid_world_env.py:8-250): same maps, constructor,
spaces (Discrete observation = cell index, Discrete(4) action), `reset()` / `step()` results and
errors; the transition runs in `b2g_gridworld_step`, whole VecWorker rollouts with a
ScriptedPolicy in `b2g_gridworld_rollout` (through `garage_b200.sampler.LocalSampler`)."""
import numpy as np
import torch

from garage_b200 import ops
from garage_b200._dtypes import StepType
from garage_b200._environment import Environment, EnvSpec, EnvStep, finite_len
from garage_b200.spaces import Discrete

MAPS = {
    'chain': ['GFFFFFFFFFFFFFSFFFFFFFFFFFFFG'],
    '4x4_safe': ['SFFF', 'FWFW', 'FFFW', 'WFFG'],
    '4x4': ['SFFF', 'FHFH', 'FFFH', 'HFFG'],
    '8x8': ['SFFFFFFF', 'FFFFFFFF', 'FFFHFFFF', 'FFFFFHFF', 'FFFHFFFF', 'FHHFFFHF', 'FHFFHFHF', 'FFFHFFFG'],
}
_CODES = {'S': 0, 'F': 0, '.': 0, 'W': 1, 'x': 1, 'H': 2, 'o': 2, 'G': 3}


class GridWorldEnv(Environment):
    """| 'S' start | 'F' or '.' free | 'W' or 'x' wall | 'H' or 'o' hole (ends the episode) | 'G' goal"""

    env_id = 'gridworld'
    has_success = False
    never_terminates = False

    def __init__(self, desc='4x4', max_episode_length=None):
        rows = MAPS[desc] if isinstance(desc, str) else desc
        grid = np.array([list(r) for r in rows])
        if grid.ndim != 2:
            raise ValueError('desc must be a rectangular grid')
        unknown = set(grid.reshape(-1)) - set(_CODES)
        if unknown:
            raise NotImplementedError('unknown cell types {}'.format(sorted(unknown)))
        starts = np.argwhere(grid == 'S')
        if len(starts) != 1:
            raise ValueError('the grid needs exactly one start cell')
        self._desc = grid
        self._n_row, self._n_col = grid.shape
        self._start_state = int(starts[0][0] * self._n_col + starts[0][1])
        self._cells = np.vectorize(_CODES.get)(grid).astype(np.uint8).reshape(-1)
        self._max_episode_length = max_episode_length
        self._action_space = Discrete(4)
        self._observation_space = Discrete(self._n_row * self._n_col)
        self._spec = EnvSpec(action_space=self._action_space, observation_space=self._observation_space,
                             max_episode_length=max_episode_length)
        self._gs = None

    action_space = property(lambda self: self._action_space)
    observation_space = property(lambda self: self._observation_space)
    spec = property(lambda self: self._spec)
    render_modes = property(lambda self: [])
    max_len = property(lambda self: finite_len(self._max_episode_length))

    def grid_state(self, n, device=None):
        """Fresh device state for n copies of this map (used by the sampler)."""
        device = device or torch.device('cuda', torch.cuda.current_device())
        return ops.GridState(self._cells, self._n_row, self._n_col, self._start_state, n, device)

    def episode_info(self):
        return {}

    def env_info_consts(self):
        return {}

    def reset(self):
        if self._gs is None:
            self._gs = self.grid_state(1)
        self._gs.reset()
        return self._start_state, dict()

    def step(self, action):
        if self._gs is None:
            raise RuntimeError('reset() must be called before step()!')
        act = torch.full((1, ), int(action), dtype=torch.int32, device=self._gs.device)
        obs, rew, st = ops.gridworld_step(self._gs, self.max_len, act)
        return EnvStep(env_spec=self.spec, action=action, reward=float(rew.item()), observation=int(obs.item()),
                       env_info={}, step_type=StepType(int(st.item())))

    def render(self, mode):
        pass

    def visualize(self):
        pass

    def close(self):
        self._gs = None

    def __getstate__(self):
        d = self.__dict__.copy()
        d['_gs'] = None
        return d
