"""Samplers of the on-policy hot path."""
from garage_b200.sampler.gpu_sampler import GPUSampler, LocalSampler, Sampler, VecWorker

__all__ = ['Sampler', 'GPUSampler', 'LocalSampler', 'VecWorker']
