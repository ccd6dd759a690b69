from garage_b200.torch.optimizers.conjugate_gradient_optimizer import ConjugateGradientOptimizer
from garage_b200.torch.optimizers.optimizer_wrapper import OptimizerWrapper

__all__ = ['OptimizerWrapper', 'ConjugateGradientOptimizer']
