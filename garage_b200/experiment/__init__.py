from garage_b200.experiment import deterministic

__all__ = ['deterministic']
