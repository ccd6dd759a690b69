"""`SnapshotConfig` / `Snapshotter` with the reference's file naming and modes
(experiment/snapshotter.py:9-170): 'all' -> itr_N.pkl every iteration, 'last' -> params.pkl
overwritten, 'gap' -> itr_N.pkl every `snapshot_gap`, 'gap_overwrite' -> params.pkl every
`snapshot_gap`, 'gap_and_last' -> both, 'none'.  Snapshots are cloudpickle files; the
garage_b200 classes drop their device buffers in `__getstate__` and rebuild them on the current
CUDA device when loaded, so a snapshot written on one GPU resumes on any other."""
import collections
import errno
import os

import cloudpickle

SnapshotConfig = collections.namedtuple('SnapshotConfig',
                                        ['snapshot_dir', 'snapshot_mode', 'snapshot_gap'])

_MODES = ('all', 'last', 'gap', 'gap_overwrite', 'gap_and_last', 'none')


class NotAFileError(Exception):
    """The snapshot path exists but is not a regular file (snapshotter.py:173)."""


class Snapshotter:

    def __init__(self, snapshot_dir=None, snapshot_mode='last', snapshot_gap=1):
        if snapshot_dir is None:
            snapshot_dir = os.path.join(os.getcwd(), 'data/local/experiment')
        if snapshot_mode == 'gap_overwrite' and snapshot_gap <= 1:
            raise ValueError('snapshot_gap must be > 1 when using snapshot_mode="gap_overwrite". '
                             'Use snapshot_mode="last" to snapshot after every iteration.')
        if snapshot_mode == 'last' and snapshot_gap != 1:
            raise ValueError('snapshot_gap should be set to 1 if using snapshot_mode="last". '
                             'Did you mean to use snapshot_mode="gap"?')
        self._snapshot_dir = snapshot_dir
        self._snapshot_mode = snapshot_mode
        self._snapshot_gap = snapshot_gap
        os.makedirs(snapshot_dir, exist_ok=True)

    snapshot_dir = property(lambda self: self._snapshot_dir)
    snapshot_mode = property(lambda self: self._snapshot_mode)
    snapshot_gap = property(lambda self: self._snapshot_gap)

    def _targets(self, itr):
        """File names one call writes, in the reference's order."""
        mode, gap = self._snapshot_mode, self._snapshot_gap
        if mode not in _MODES:
            raise ValueError('Invalid snapshot mode {}'.format(mode))
        numbered, last = 'itr_%d.pkl' % itr, 'params.pkl'
        on_gap = itr % gap == 0 if mode.startswith('gap') else False
        return {
            'all': [numbered],
            'last': [last],
            'gap': [numbered] if on_gap else [],
            'gap_overwrite': [last] if on_gap else [],
            'gap_and_last': [last] + ([numbered] if on_gap else []),
            'none': [],
        }[mode]

    def save_itr_params(self, itr, params):
        names = self._targets(itr)
        if not names:
            return
        blob = cloudpickle.dumps(params)
        for name in names:
            with open(os.path.join(self._snapshot_dir, name), 'wb') as fh:
                fh.write(blob)

    def load(self, load_dir, itr='last'):
        return load_snapshot(load_dir, itr)


def load_snapshot(load_dir, itr='last'):
    """`Snapshotter.load` (snapshotter.py:128-170) without needing an instance.
    `itr`: an integer (or digit string) -> itr_N.pkl; 'last' / 'first' -> params.pkl if it
    exists, else the last / first *.pkl by name."""
    if isinstance(itr, int) or itr.isdigit():
        path = os.path.join(load_dir, 'itr_{}.pkl'.format(itr))
    else:
        if itr not in ('last', 'first'):
            raise ValueError("itr should be an integer or 'last' or 'first'")
        path = os.path.join(load_dir, 'params.pkl')
        if not os.path.isfile(path):
            files = sorted(f for f in os.listdir(load_dir) if f.endswith('.pkl'))
            if not files:
                raise FileNotFoundError(errno.ENOENT, os.strerror(errno.ENOENT),
                                        '*.pkl file in', load_dir)
            path = os.path.join(load_dir, files[0 if itr == 'first' else -1])
    if not os.path.isfile(path):
        raise NotAFileError('File not existing: ', path)
    with open(path, 'rb') as fh:
        return cloudpickle.load(fh)
