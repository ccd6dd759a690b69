"""`wrap_experiment` for launcher scripts (reference experiment/experiment.py:98-448).

Same calling conventions: bare decorator or `@wrap_experiment(snapshot_mode='all', ...)`, the
wrapped function takes `ctxt` first and keyword arguments after it, an optional single dict
argument overrides the decorator options for one call.  Per call it creates the log directory
(`data/local/<prefix>/<name>[_k=v...]`, `_1`, `_2`, ... appended when it exists unless
`use_existing_dir`), writes `variant.json`, attaches a `progress.csv` writer to the tabular
stand-in (dowel and TensorBoard are absent here) and hands the function an `ExperimentContext`
that `Trainer(ctxt)` turns into a `Snapshotter`.  Launch-repo archiving and git metadata are
launcher services outside the accelerated path and are not reproduced (`metadata.json` only
records the launcher's path)."""
import collections
import enum
import functools
import inspect
import json
import os
import sys

import numpy as np

from garage_b200.tabular import tabular

_OPTION_KEYS = ('name', 'function', 'prefix', 'name_parameters', 'log_dir', 'archive_launch_repo',
                'snapshot_gap', 'snapshot_mode', 'use_existing_dir', 'x_axis', 'signature')


class ExperimentContext:
    """What the wrapped function receives (experiment.py:98-120); same fields as SnapshotConfig."""

    def __init__(self, *, snapshot_dir, snapshot_mode, snapshot_gap):
        self.snapshot_dir = snapshot_dir
        self.snapshot_mode = snapshot_mode
        self.snapshot_gap = snapshot_gap


def _experiment_signature(function):
    """Signature without the leading `ctxt`; the rest keyword-only (experiment.py:55-95)."""
    params = list(inspect.signature(function).parameters.values())
    if not params:
        raise ValueError("Experiment functions should have a first parameter named 'ctxt', "
                         'but {!r} has no parameters'.format(function))
    if params[0].name != 'ctxt':
        raise ValueError("Experiment functions should have a first parameter named 'ctxt' "
                         'instead of {!r}'.format(params[0].name))
    return inspect.Signature([p.replace(kind=inspect.Parameter.KEYWORD_ONLY) for p in params[1:]],
                             return_annotation=inspect.signature(function).return_annotation)


def _claim_log_dir(log_dir):
    """mkdir log_dir, or the first free log_dir_1, log_dir_2, ... (experiment.py:28-52)."""
    suffix = 0
    while True:
        candidate = log_dir if suffix == 0 else '{}_{}'.format(log_dir, suffix)
        try:
            os.makedirs(candidate)
            return candidate
        except FileExistsError:
            suffix += 1


def _jsonable(o, depth=0):
    """Best-effort plain-data view of experiment arguments (stands in for LogEncoder,
    experiment.py:562-640)."""
    if isinstance(o, (str, bool, int, float)) or o is None:
        return o
    if isinstance(o, enum.Enum):
        return '{}.{}'.format(type(o).__name__, o.name)
    if isinstance(o, np.generic):
        return o.item()
    if isinstance(o, np.ndarray):
        return o.tolist()
    if depth > 4:
        return repr(o)
    if isinstance(o, dict):
        return {str(k): _jsonable(v, depth + 1) for k, v in o.items()}
    if isinstance(o, (list, tuple, set)):
        return [_jsonable(v, depth + 1) for v in o]
    if isinstance(o, type) or callable(o):
        return '{}.{}'.format(getattr(o, '__module__', ''), getattr(o, '__qualname__', repr(o)))
    return repr(o)


def dump_json(filename, data):
    os.makedirs(os.path.dirname(os.path.abspath(filename)), exist_ok=True)
    with open(filename, 'w') as fh:
        json.dump(_jsonable(data), fh, indent=2, sort_keys=True)


class ExperimentTemplate:
    """Created by `wrap_experiment` only (experiment.py:123-373)."""

    def __init__(self, *, function, log_dir, name, prefix, snapshot_mode, snapshot_gap,
                 archive_launch_repo, name_parameters, use_existing_dir, x_axis):
        self.function = function
        self.log_dir = log_dir
        self.name = name
        self.prefix = prefix
        self.snapshot_mode = snapshot_mode
        self.snapshot_gap = snapshot_gap
        self.archive_launch_repo = archive_launch_repo
        self.name_parameters = name_parameters
        self.use_existing_dir = use_existing_dir
        self.x_axis = x_axis
        if function is not None:
            self._adopt(function)

    def _adopt(self, function):
        functools.update_wrapper(self, function)
        self.function = function
        self.__signature__ = _experiment_signature(function)

    @staticmethod
    def _augment_name(options, name, params):
        mode = options['name_parameters']
        if mode not in (None, 'passed', 'all'):
            raise ValueError('wrap_experiment.name_parameters should be set to one of None, '
                             '"passed", or "all"')
        shown = collections.OrderedDict()
        if mode is not None:
            for p in options['signature'].parameters.values():
                if p.name in params:
                    shown[p.name] = params[p.name]
                elif mode == 'all':
                    shown[p.name] = p.default
        tail = '_'.join('{}={}'.format(k, v) for k, v in shown.items())
        return '{}_{}'.format(name, tail) if tail else name

    def _get_options(self, *args):
        options = {k: getattr(self, k, None) for k in _OPTION_KEYS if k != 'signature'}
        options['signature'] = self.__signature__
        if args:
            if len(args) != 1 or not isinstance(args[0], dict):
                raise ValueError('garage.experiment currently only supports keyword arguments')
            for k in args[0]:
                if k not in options:
                    raise ValueError('Unknown key {} in wrap_experiment options'.format(k))
            options.update(args[0])
        return options

    @classmethod
    def _make_context(cls, options, **kwargs):
        name = options['name'] or options['function'].__name__
        name = cls._augment_name(options, name, kwargs)
        log_dir = options['log_dir']
        if log_dir is None:
            log_dir = os.path.join(os.getcwd(), 'data', 'local', options['prefix'], name)
        if options['use_existing_dir']:
            os.makedirs(log_dir, exist_ok=True)
        else:
            log_dir = _claim_log_dir(log_dir)
        dump_json(os.path.join(log_dir, 'variant.json'), kwargs)
        dump_json(os.path.join(log_dir, 'metadata.json'),
                  {'launcher': os.path.abspath(sys.argv[0]) if sys.argv and sys.argv[0] else None})
        tabular.add_csv_output(os.path.join(log_dir, 'progress.csv'))
        return ExperimentContext(snapshot_dir=log_dir, snapshot_mode=options['snapshot_mode'],
                                 snapshot_gap=options['snapshot_gap'])

    def __call__(self, *args, **kwargs):
        if self.function is None:
            if len(args) != 1 or kwargs or not callable(args[0]):
                raise ValueError('Please apply the result of wrap_experiment() to a single function')
            self._adopt(args[0])
            return self
        ctxt = self._make_context(self._get_options(*args), **kwargs)
        try:
            return self.function(ctxt, **kwargs)
        finally:
            tabular.remove_all_outputs()


def wrap_experiment(function=None, *, log_dir=None, prefix='experiment', name=None,
                    snapshot_mode='last', snapshot_gap=1, archive_launch_repo=True,
                    name_parameters=None, use_existing_dir=False, x_axis='TotalEnvSteps'):
    """See the module docstring; argument meaning as the reference (experiment.py:376-448)."""
    return ExperimentTemplate(function=function, log_dir=log_dir, prefix=prefix, name=name,
                              snapshot_mode=snapshot_mode, snapshot_gap=snapshot_gap,
                              archive_launch_repo=archive_launch_repo,
                              name_parameters=name_parameters, use_existing_dir=use_existing_dir,
                              x_axis=x_axis)
