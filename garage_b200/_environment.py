"""Environment API of the on-policy path (mirror of garage `_environment.py`:
EnvSpec :21-41, EnvStep :84-144, Environment :146-356)."""
import abc
import math
from dataclasses import dataclass

import numpy as np

from garage_b200._dtypes import StepType


class EnvSpec:
    """observation_space, action_space, max_episode_length (reference `:21-41`)."""

    def __init__(self, observation_space, action_space, max_episode_length=None):
        self.observation_space = observation_space
        self.action_space = action_space
        self.max_episode_length = max_episode_length

    def __eq__(self, other):
        return (isinstance(other, EnvSpec) and self.observation_space == other.observation_space and
                self.action_space == other.action_space and
                self.max_episode_length == other.max_episode_length)

    def __repr__(self):
        return 'EnvSpec(obs={}, act={}, max_episode_length={})'.format(
            self.observation_space, self.action_space, self.max_episode_length)


@dataclass(frozen=True)
class EnvStep:
    """One environment step (reference `:84-144`)."""
    env_spec: EnvSpec
    action: np.ndarray
    reward: float
    observation: np.ndarray
    env_info: dict
    step_type: StepType

    @property
    def first(self):
        return self.step_type is StepType.FIRST

    @property
    def mid(self):
        return self.step_type is StepType.MID

    @property
    def terminal(self):
        return self.step_type is StepType.TERMINAL

    @property
    def timeout(self):
        return self.step_type is StepType.TIMEOUT

    @property
    def last(self):
        return self.step_type is StepType.TERMINAL or self.step_type is StepType.TIMEOUT


class Environment(abc.ABC):
    """reset()/step() contract of the reference (`:146-356`)."""

    @property
    @abc.abstractmethod
    def action_space(self):
        pass

    @property
    @abc.abstractmethod
    def observation_space(self):
        pass

    @property
    @abc.abstractmethod
    def spec(self):
        pass

    @property
    def render_modes(self):
        return []

    @abc.abstractmethod
    def reset(self):
        pass

    @abc.abstractmethod
    def step(self, action):
        pass

    def render(self, mode):
        raise NotImplementedError('rendering is unavailable on this box (no graphics stack)')

    def visualize(self):
        pass

    def close(self):
        pass


def finite_len(max_episode_length):
    """None / inf -> 0 (the C ABI's "no limit")."""
    if max_episode_length is None or (isinstance(max_episode_length, float) and math.isinf(max_episode_length)):
        return 0
    return int(max_episode_length)
