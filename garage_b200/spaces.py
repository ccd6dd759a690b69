"""Minimal Box space (the subset of `akro.Box` the on-policy path touches:
`flat_dim`, `flatten(_n)`, `contains`, `low/high/shape`, `sample`).  akro is a
third-party dependency of the reference and is not installed here."""
import numpy as np


class Space:
    pass


class Box(Space):

    def __init__(self, low, high, shape=None, dtype=np.float32):
        if shape is None:
            shape = np.shape(low)
        self.shape = tuple(int(s) for s in shape)
        self.dtype = np.dtype(dtype)
        self.low = np.broadcast_to(np.asarray(low, dtype=self.dtype), self.shape).copy()
        self.high = np.broadcast_to(np.asarray(high, dtype=self.dtype), self.shape).copy()

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
        return bool(x.shape == self.shape and np.all(x >= self.low) and np.all(x <= self.high))

    def sample(self):
        lo = np.where(np.isfinite(self.low), self.low, -1.0)
        hi = np.where(np.isfinite(self.high), self.high, 1.0)
        return np.random.uniform(lo, hi).astype(self.dtype)

    def __eq__(self, other):
        return (isinstance(other, Box) and self.shape == other.shape and
                np.array_equal(self.low, other.low) and np.array_equal(self.high, other.high))

    def __hash__(self):
        return hash((self.shape, self.low.tobytes(), self.high.tobytes()))

    def __repr__(self):
        return 'Box{}'.format(self.shape)


class Discrete(Space):
    """The subset of `akro.Discrete` the GridWorld path touches: `n`, `flat_dim` (one-hot width),
    `flatten(_n)` / `unflatten(_n)` (one-hot), `contains`, `sample`."""

    def __init__(self, n):
        self.n = int(n)
        self.shape = ()
        self.dtype = np.dtype(np.int64)

    @property
    def flat_dim(self):
        return self.n

    def flatten(self, x):
        out = np.zeros(self.n)
        out[int(x)] = 1
        return out

    def unflatten(self, x):
        return int(np.nonzero(np.asarray(x))[0][0])

    def flatten_n(self, xs):
        xs = np.asarray(xs, dtype=np.int64).reshape(-1)
        out = np.zeros((len(xs), self.n))
        out[np.arange(len(xs)), xs] = 1
        return out

    def unflatten_n(self, xs):
        return np.nonzero(np.asarray(xs))[1]

    def contains(self, x):
        x = np.asarray(x)
        return bool(x.shape == () and x.dtype.kind in 'iu' and 0 <= int(x) < self.n)

    def sample(self):
        return int(np.random.randint(self.n))

    def __eq__(self, other):
        return isinstance(other, Discrete) and self.n == other.n

    def __hash__(self):
        return hash(('Discrete', self.n))

    def __repr__(self):
        return 'Discrete({})'.format(self.n)
