"""B200-native drop-ins for the `garage.np` names on the on-policy hot path."""
from garage_b200.np._functions import discount_cumsum

__all__ = ['discount_cumsum']
