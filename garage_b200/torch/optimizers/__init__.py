from garage_b200.torch.optimizers.optimizer_wrapper import OptimizerWrapper

__all__ = ['OptimizerWrapper']
