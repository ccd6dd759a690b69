"""Minimal stand-in for `dowel.tabular` / `dowel.logger` (third-party, absent
here): same `record(key, value)` / `prefix(...)` calls as the reference's
`_train_once` (torch/algos/vpg.py:186-199) and `log_performance`
(_functions.py:262-273); values are kept in a dict per iteration and, when an
output is attached (`wrap_experiment` attaches `progress.csv`, the file dowel's
CsvOutput would write), appended as one CSV row per `dump()`."""
import contextlib
import csv
import os


class _CsvOutput:
    """One row per dump; the header is the key set of the first row, later new keys widen it
    (the file is rewritten once when that happens)."""

    def __init__(self, path):
        self.path = path
        self.fields = []
        self.rows = 0
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        open(path, 'w').close()

    def write(self, record):
        new = [k for k in record if k not in self.fields]
        if new and self.rows:
            with open(self.path, newline='') as fh:
                old = list(csv.DictReader(fh))
            self.fields += new
            with open(self.path, 'w', newline='') as fh:
                w = csv.DictWriter(fh, fieldnames=self.fields, restval='')
                w.writeheader()
                w.writerows(old)
        elif new:
            self.fields += new
            with open(self.path, 'w', newline='') as fh:
                csv.DictWriter(fh, fieldnames=self.fields).writeheader()
        with open(self.path, 'a', newline='') as fh:
            csv.DictWriter(fh, fieldnames=self.fields, restval='').writerow(record)
        self.rows += 1


class Tabular:

    def __init__(self):
        self._data = {}
        self._prefix = ''
        self._outputs = []

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

    # ---- outputs (dowel: logger.add_output(CsvOutput(...)), logger.log(tabular), dump_all) ----
    def add_csv_output(self, path):
        self._outputs.append(_CsvOutput(path))

    def remove_all_outputs(self):
        self._outputs = []

    def dump(self):
        """Write the current record to every attached output (the trainer calls this once per
        epoch, where the reference calls `logger.log(tabular); logger.dump_all(itr)`)."""
        rec = {k: (float(v) if hasattr(v, '__float__') else v) for k, v in self._data.items()}
        for out in self._outputs:
            out.write(rec)


tabular = Tabular()
