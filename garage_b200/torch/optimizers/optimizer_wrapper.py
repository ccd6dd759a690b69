"""OptimizerWrapper drop-in (reference torch/optimizers/optimizer_wrapper.py:6-63 +
np/optimizers/minibatch_dataset.py:4-35 + make_optimizer _functions.py:25-65).

Same constructor; the optimiser is `torch.optim.Adam` (the one on the PPO/VPG path:
ppo.py:65-76, vpg.py:97-105) -- its state (exp_avg, exp_avg_sq, step) lives in flat CUDA
vectors next to the module's flat parameter vector and the update is libb2g's fused
reduce+Adam kernel -- or `ConjugateGradientOptimizer` (TRPO's policy optimiser,
trpo.py:62-66), whose instance is kept in `cg`."""
import numpy as np
import torch

_ADAM_DEFAULTS = dict(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False)


class OptimizerWrapper:

    def __init__(self, optimizer, module, max_optimization_epochs=1, minibatch_size=None):
        if isinstance(optimizer, tuple):
            opt_type, opt_args = optimizer
        else:
            opt_type, opt_args = optimizer, {}
        self._module = module
        self._max_optimization_epochs = max_optimization_epochs
        self._minibatch_size = minibatch_size
        self._state = None
        self.cg = None
        from garage_b200.torch.optimizers.conjugate_gradient_optimizer import ConjugateGradientOptimizer
        if opt_type is ConjugateGradientOptimizer:
            self.cg = ConjugateGradientOptimizer(None, **opt_args)
            self.lr = self.beta1 = self.beta2 = self.eps = None
            return
        if opt_type is not torch.optim.Adam:
            raise NotImplementedError('garage_b200 implements torch.optim.Adam and ConjugateGradientOptimizer '
                                      '(got {})'.format(opt_type))
        args = dict(_ADAM_DEFAULTS)
        unknown = set(opt_args) - set(args)
        if unknown:
            raise NotImplementedError('unsupported Adam arguments: {}'.format(sorted(unknown)))
        args.update(opt_args)
        if args['weight_decay'] != 0 or args['amsgrad']:
            raise NotImplementedError('weight_decay / amsgrad are not implemented in the Adam kernel')
        self.lr = float(args['lr'])
        self.beta1, self.beta2 = (float(b) for b in args['betas'])
        self.eps = float(args['eps'])

    @property
    def module(self):
        return self._module

    @property
    def max_optimization_epochs(self):
        return self._max_optimization_epochs

    @property
    def minibatch_size(self):
        return self._minibatch_size

    def state(self):
        """(exp_avg, exp_avg_sq, step) flat CUDA tensors, created lazily."""
        flat = self._module.flat
        if self._state is None or self._state[0].device != flat.device:
            self._state = (torch.zeros_like(flat), torch.zeros_like(flat),
                           torch.zeros(1, dtype=torch.int64, device=flat.device))
        return self._state

    def get_minibatch(self, *inputs):
        """Yield minibatches exactly like the reference (`optimizer_wrapper.py:31-49`
        over BatchDataset, minibatch_dataset.py:4-35): ids shuffled with numpy's global
        RNG at construction and after every epoch."""
        n = inputs[0].shape[0]
        if self._minibatch_size is None:
            for _ in range(self._max_optimization_epochs):
                yield list(inputs)
            return
        ids = np.arange(n)
        np.random.shuffle(ids)
        for _ in range(self._max_optimization_epochs):
            for start in range(0, n, self._minibatch_size):
                sel = ids[start:start + self._minibatch_size]
                yield [d[sel] for d in inputs]
            np.random.shuffle(ids)

    def zero_grad(self):
        """Gradients are produced and consumed inside the fused kernels."""

    def step(self, **closure):
        raise NotImplementedError('the parameter update is fused into VPG/PPO._train; there is no '
                                  'separate autograd step in garage_b200')

    def __getstate__(self):
        s = self.__dict__.copy()
        if self._state is not None:
            s['_state'] = tuple(t.cpu() for t in self._state)
        return s

    def __setstate__(self, s):
        self.__dict__.update(s)
        if self._state is not None:
            dev = torch.device('cuda', torch.cuda.current_device())
            self._state = tuple(t.to(dev) for t in self._state)
