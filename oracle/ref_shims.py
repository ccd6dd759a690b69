"""TEST INFRASTRUCTURE ONLY -- import shims that let the UNMODIFIED reference
(`/root/reference/src/garage`, v2021.03.0) be imported in the build container.

Only `oracle/gen_golden.py` (run here, by hand) and the `not gpu` tests that
cross-check the restatement use this; nothing under `garage_b200/` may import
it.  `/root/reference` does not exist on the GPU box, so `install()` returns
False there and callers skip.

What is stubbed (the reference imports them at package-import time, none is on
the arithmetic of the hot path):
  dowel, gym, ray, tensorflow(+_probability), cma, skimage, mujoco_py,
  dm_control, metaworld, click-free.  `dowel.tabular` is a real collector so
  the keys `_train_once` records can be read back.
`akro` gets a ~60 line real implementation of `Box` / `Discrete`
(flat_dim / flatten / flatten_n / contains / sample) because PointEnv and the
policies call it.
"""
import contextlib
import importlib.abc
import importlib.machinery
import os
import sys
import types

import numpy as np

REFERENCE_SRC = '/root/reference/src'
# staged copy of the unmodified package (oracle/make_ref.py; git-ignored, travels with gpurun)
STAGED_SRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), '_ref')

_STUB_ROOTS = ('dowel', 'gym', 'ray', 'tensorflow', 'tensorflow_probability',
               'cma', 'skimage', 'mujoco_py', 'dm_control', 'metaworld',
               'dm_env', 'pybullet', 'pybullet_envs', 'cv2', 'PIL',
               'matplotlib', 'setproctitle', 'glfw')


class _Anything:
    """Permissive placeholder: callable, subscriptable, attribute-chainable."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, name):
        if name.startswith('__'):
            raise AttributeError(name)
        return _Anything()

    def __getitem__(self, k):
        return _Anything()

    def __iter__(self):
        return iter(())

    def __mro_entries__(self, bases):
        return (object, )


class _StubModule(types.ModuleType):
    __path__ = []

    def __getattr__(self, name):
        if name.startswith('__'):
            raise AttributeError(name)
        # classes so that `class Foo(gym.Wrapper)` works
        cls = type(name, (), {'__init__': lambda self, *a, **k: None,
                              '__getattr__': lambda self, n: _Anything()})
        setattr(self, name, cls)
        return cls


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):

    def find_spec(self, fullname, path, target=None):
        if fullname.split('.')[0] in _STUB_ROOTS:
            return importlib.machinery.ModuleSpec(fullname, self,
                                                  is_package=True)
        return None

    def create_module(self, spec):
        return _StubModule(spec.name)

    def exec_module(self, module):
        pass


class Tabular:
    """Stand-in for dowel.tabular: records into a dict that tests can read."""

    def __init__(self):
        self.data = {}
        self._prefix = ''

    def record(self, key, val):
        self.data[self._prefix + str(key)] = val

    @contextlib.contextmanager
    def prefix(self, p):
        old = self._prefix
        self._prefix = old + str(p)
        try:
            yield
        finally:
            self._prefix = old

    def clear(self):
        self.data.clear()


def _make_akro():
    akro = types.ModuleType('akro')

    class Space:
        pass

    class Box(Space):

        def __init__(self, low, high, shape=None, dtype=np.float32):
            if shape is None:
                low = np.asarray(low)
                shape = low.shape
            self.shape = tuple(shape)
            self.dtype = np.dtype(dtype)
            self.low = np.broadcast_to(np.asarray(low, dtype=dtype),
                                       self.shape).copy()
            self.high = np.broadcast_to(np.asarray(high, dtype=dtype),
                                        self.shape).copy()

        @property
        def flat_dim(self):
            return int(np.prod(self.shape))

        @property
        def bounds(self):
            return self.low, self.high

        def flatten(self, x):
            return np.asarray(x).flatten()

        def unflatten(self, x):
            return np.asarray(x).reshape(self.shape)

        def flatten_n(self, xs):
            xs = np.asarray(xs)
            return xs.reshape((xs.shape[0], -1))

        def unflatten_n(self, xs):
            xs = np.asarray(xs)
            return xs.reshape((xs.shape[0], ) + self.shape)

        def contains(self, x):
            x = np.asarray(x)
            return (x.shape == self.shape and np.all(x >= self.low)
                    and np.all(x <= self.high))

        def sample(self):
            lo = np.where(np.isfinite(self.low), self.low, -1.0)
            hi = np.where(np.isfinite(self.high), self.high, 1.0)
            return np.random.uniform(lo, hi).astype(self.dtype)

        def __eq__(self, other):
            return (isinstance(other, Box) and self.shape == other.shape
                    and np.allclose(self.low, other.low)
                    and np.allclose(self.high, other.high))

        def __repr__(self):
            return 'Box{}'.format(self.shape)

    class Discrete(Space):

        def __init__(self, n):
            self.n = n
            self.shape = ()
            self.dtype = np.dtype(np.int64)

        @property
        def flat_dim(self):
            return self.n

        def flatten(self, x):
            out = np.zeros(self.n)
            out[x] = 1
            return out

        def flatten_n(self, xs):
            out = np.zeros((len(xs), self.n))
            out[np.arange(len(xs)), xs] = 1
            return out

        def contains(self, x):
            return 0 <= int(x) < self.n

        def sample(self):
            return np.random.randint(self.n)

        def __eq__(self, other):
            return isinstance(other, Discrete) and self.n == other.n

    akro.Space = Space
    akro.Box = Box
    akro.Discrete = Discrete
    for name in ('Dict', 'Tuple', 'Image'):
        setattr(akro, name, type(name, (Space, ), {}))
    akro.from_gym = lambda space, **kw: space
    return akro


_installed = False
tabular = Tabular()


def source_dir():
    """Directory holding the reference's `garage` package: the reference tree itself in the
    build container, the staged copy on the GPU box; None when neither exists."""
    for d in (REFERENCE_SRC, STAGED_SRC):
        if os.path.isdir(os.path.join(d, 'garage')):
            return d
    return None


def available():
    return source_dir() is not None


def install():
    """Make `import garage` resolve to the reference.  False if absent."""
    global _installed
    if not available():
        return False
    if _installed:
        return True
    sys.meta_path.insert(0, _StubFinder())
    import dowel  # noqa: F401  (the stub)
    sys.modules['dowel'].tabular = tabular
    sys.modules['dowel'].logger = _Anything()
    sys.modules['akro'] = _make_akro()
    src = source_dir()
    if src not in sys.path:
        sys.path.insert(0, src)
    _installed = True
    return True
