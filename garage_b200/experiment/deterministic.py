"""Seeding (reference experiment/deterministic.py:12-56): seeds python, numpy and
torch; the same integer keys garage_b200's Philox streams."""
import random

import numpy as np
import torch

_seed = None


def set_seed(seed):
    global _seed
    seed %= 4294967294
    _seed = seed
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)


def get_seed():
    return _seed
