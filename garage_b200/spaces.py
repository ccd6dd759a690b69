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
