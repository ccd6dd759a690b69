from garage_b200.torch.value_functions.gaussian_mlp_value_function import (GaussianMLPValueFunction,
                                                                            ValueFunction)

__all__ = ['ValueFunction', 'GaussianMLPValueFunction']
