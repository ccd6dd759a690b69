from garage_b200.experiment import deterministic
from garage_b200.experiment.experiment import ExperimentContext, ExperimentTemplate, wrap_experiment
from garage_b200.experiment.snapshotter import NotAFileError, SnapshotConfig, Snapshotter

__all__ = ['deterministic', 'wrap_experiment', 'ExperimentContext', 'ExperimentTemplate',
           'Snapshotter', 'SnapshotConfig', 'NotAFileError']
