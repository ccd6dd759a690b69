"""Minimal stand-in for `dowel.tabular` / `dowel.logger` (third-party, absent
here): same `record(key, value)` / `prefix(...)` calls as the reference's
`_train_once` (torch/algos/vpg.py:186-199) and `log_performance`
(_functions.py:262-273); values are kept in a dict per iteration."""
import contextlib


class Tabular:

    def __init__(self):
        self._data = {}
        self._prefix = ''

    def record(self, key, val):
        self._data[self._prefix + str(key)] = val

    @contextlib.contextmanager
    def prefix(self, prefix):
        old = self._prefix
        self._prefix = old + str(prefix)
        try:
            yield
        finally:
            self._prefix = old

    @property
    def as_dict(self):
        return dict(self._data)

    def clear(self):
        self._data.clear()


tabular = Tabular()
