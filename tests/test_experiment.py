"""Snapshotter / wrap_experiment / tabular CSV: the cases of the reference's
tests/garage/experiment/test_snapshotter.py and test_experiment.py that do not need git,
dowel or TensorBoard, run against garage_b200's implementations (host logic, CPU only)."""
import csv
import os
import pickle

import pytest

from garage_b200 import wrap_experiment
from garage_b200.experiment import NotAFileError, SnapshotConfig, Snapshotter
from garage_b200.tabular import Tabular

DATA = [{'testparam': 1}, {'testparam': 4}]


@pytest.mark.parametrize('mode, files', [
    ('all', {'itr_1.pkl': 0, 'itr_2.pkl': 1}),
    ('last', {'params.pkl': 1}),
    ('gap', {'itr_2.pkl': 1}),
    ('gap_and_last', {'itr_2.pkl': 1, 'params.pkl': 1}),
    ('none', {}),
])
def test_snapshot_modes(tmp_path, mode, files):
    s = Snapshotter(str(tmp_path), mode, 1)
    assert (s.snapshot_dir, s.snapshot_mode, s.snapshot_gap) == (str(tmp_path), mode, 1)
    s.save_itr_params(1, DATA[0])
    s.save_itr_params(2, DATA[1])
    for name, which in files.items():
        with open(tmp_path / name, 'rb') as fh:
            assert pickle.load(fh) == DATA[which]
    if not files:
        assert list(tmp_path.iterdir()) == []


def test_snapshot_gap_modes(tmp_path):
    s = Snapshotter(str(tmp_path / 'a'), 'gap_overwrite', 2)
    s.save_itr_params(1, DATA[0])
    assert not (tmp_path / 'a' / 'params.pkl').exists()
    s.save_itr_params(2, DATA[1])
    assert s.load(str(tmp_path / 'a')) == DATA[1]
    g = Snapshotter(str(tmp_path / 'b'), 'gap', 3)
    for itr in range(7):
        g.save_itr_params(itr, {'itr': itr})
    assert sorted(os.listdir(tmp_path / 'b')) == ['itr_0.pkl', 'itr_3.pkl', 'itr_6.pkl']
    assert g.load(str(tmp_path / 'b'), 'first') == {'itr': 0}
    assert g.load(str(tmp_path / 'b'), 'last') == {'itr': 6}
    assert g.load(str(tmp_path / 'b'), 3) == {'itr': 3} == g.load(str(tmp_path / 'b'), '3')


def test_snapshot_errors(tmp_path):
    with pytest.raises(ValueError):
        Snapshotter(str(tmp_path), 'invalid').save_itr_params(2, {})
    with pytest.raises(ValueError):
        Snapshotter(str(tmp_path), 'last', 2)
    with pytest.raises(ValueError):
        Snapshotter(str(tmp_path), 'gap_overwrite', 1)
    s = Snapshotter(str(tmp_path))
    with pytest.raises(ValueError):
        s.load(str(tmp_path), 'middle')
    with pytest.raises(FileNotFoundError):
        s.load(str(tmp_path))
    with pytest.raises(NotAFileError):
        s.load(str(tmp_path), 5)
    assert SnapshotConfig('d', 'last', 1).snapshot_mode == 'last'


def test_wrap_experiment_log_dirs(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    root = tmp_path / 'data' / 'local' / 'experiment'
    seen = []

    @wrap_experiment
    def exp(ctxt=None, seed=1):
        seen.append((ctxt.snapshot_dir, ctxt.snapshot_mode, ctxt.snapshot_gap, seed))
        return seed * 2

    assert exp() == 2 and exp(seed=3) == 6
    assert sorted(p.name for p in root.iterdir()) == ['exp', 'exp_1']
    assert seen[0][1:] == ('last', 1, 1) and os.path.samefile(seen[1][0], root / 'exp_1')
    for d in ('exp', 'exp_1'):
        assert {'variant.json', 'metadata.json', 'progress.csv'} <= set(os.listdir(root / d))
    assert '"seed": 3' in (root / 'exp_1' / 'variant.json').read_text()
    assert exp.__name__ == 'exp'


def test_wrap_experiment_name_parameters(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)

    @wrap_experiment(prefix='p_all', name_parameters='all')
    def exp(ctxt=None, seed=1, env='test-env'):
        return ctxt.snapshot_dir

    assert os.path.basename(exp()) == 'exp_seed=1_env=test-env'
    assert os.path.basename(exp(env='test-env-v2', seed=2)) == 'exp_seed=2_env=test-env-v2'
    assert os.path.basename(exp(env='test-env-v2')) == 'exp_seed=1_env=test-env-v2'

    @wrap_experiment(prefix='p_passed', name_parameters='passed', snapshot_mode='gap', snapshot_gap=5)
    def exp2(ctxt=None, seed=1, env='test-env'):
        assert (ctxt.snapshot_mode, ctxt.snapshot_gap) == ('gap', 5)
        return ctxt.snapshot_dir

    assert os.path.basename(exp2()) == 'exp2'
    assert os.path.basename(exp2(env='e', seed=2)) == 'exp2_seed=2_env=e'
    assert os.path.basename(exp2(env='e')) == 'exp2_env=e'
    with pytest.raises(ValueError):
        wrap_experiment(name_parameters='some')(lambda ctxt=None: None)()


def test_wrap_experiment_options_dict_and_existing_dir(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    logdir = os.path.join('data', 'local', 'dyn', 'given')

    @wrap_experiment
    def exp(ctxt=None, seed=1):
        return ctxt.snapshot_dir

    assert os.path.samefile(exp(dict(prefix='dyn', log_dir=logdir)), tmp_path / logdir)
    assert os.path.samefile(exp(dict(prefix='dyn', log_dir=logdir)), str(tmp_path / logdir) + '_1')
    assert os.path.samefile(exp(dict(log_dir=logdir, use_existing_dir=True)), tmp_path / logdir)
    assert len(os.listdir(tmp_path / 'data' / 'local' / 'dyn')) == 2
    with pytest.raises(ValueError):
        exp(dict(logdir=logdir))
    with pytest.raises(ValueError):
        exp(1)


def test_wrap_experiment_signature_errors():
    with pytest.raises(ValueError, match='ctxt'):
        @wrap_experiment
        def bad(seed=1):
            pass
    with pytest.raises(ValueError, match='ctxt'):
        @wrap_experiment
        def empty():
            pass
    with pytest.raises(ValueError):
        wrap_experiment()(1)


def test_tabular_csv_rows(tmp_path):
    t = Tabular()
    t.add_csv_output(str(tmp_path / 'progress.csv'))
    t.record('a', 1)
    with t.prefix('P/'):
        t.record('b', 2.5)
    t.dump()
    t.clear()
    t.record('a', 2)
    t.record('c', 7)  # a key that appears later widens the header
    t.dump()
    with open(tmp_path / 'progress.csv', newline='') as fh:
        rows = list(csv.DictReader(fh))
    assert rows == [{'a': '1.0', 'P/b': '2.5', 'c': ''}, {'a': '2.0', 'P/b': '', 'c': '7.0'}]
