"""garage_b200 -- B200-native drop-in for garage's on-policy hot path
(rollout -> EpisodeBatch -> GAE/returns -> PPO/VPG minibatch update).

Everything numerical runs in hand-written sm_100a kernels behind the C ABI of
`libb2g.so` (include/b2g.h); there is no CPU fallback."""
from garage_b200._dtypes import DeviceEpisodeBatch, EpisodeBatch, StepType
from garage_b200._environment import Environment, EnvSpec, EnvStep
from garage_b200.experiment import wrap_experiment
from garage_b200.experiment.snapshotter import SnapshotConfig
from garage_b200.trainer import Trainer
from garage_b200._functions import (log_multitask_performance, log_performance,
                                    obtain_evaluation_episodes)

__all__ = ['EpisodeBatch', 'DeviceEpisodeBatch', 'StepType', 'EnvSpec', 'EnvStep', 'Environment',
           'Trainer', 'wrap_experiment', 'SnapshotConfig',
           'log_performance', 'log_multitask_performance', 'obtain_evaluation_episodes']
