"""B200-native drop-ins for the `garage.torch` names on the on-policy hot path."""
from garage_b200.torch._functions import compute_advantages, filter_valids

__all__ = ['compute_advantages', 'filter_valids']
