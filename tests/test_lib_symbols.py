"""The C-ABI library builds, loads and exports every symbol include/b2g.h declares
(no compute calls: this runs without a GPU)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, 'include', 'b2g.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(b2g_[a-z0-9_]+)\s*\(', text)))


def test_library_builds_and_exports_every_declared_symbol():
    from garage_b200.csrc import build
    path = build.build()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)
    names = _declared()
    assert len(names) >= 8
    for n in names:
        assert hasattr(lib, n), 'symbol {} declared in b2g.h but not exported'.format(n)


def test_ctypes_table_matches_header():
    from garage_b200 import _lib
    assert sorted(_lib.SIGNATURES) == _declared()
    lib = _lib.load()
    assert lib.b2g_version() == 100


def test_no_oracle_import_in_product():
    """The product must never route through the oracle (or any CPU fallback)."""
    pkg = os.path.join(ROOT, 'garage_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith('.py'):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle\b', src, flags=re.M), f
