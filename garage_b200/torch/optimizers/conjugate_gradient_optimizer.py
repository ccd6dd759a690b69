"""ConjugateGradientOptimizer drop-in (reference
torch/optimizers/conjugate_gradient_optimizer.py:113-260): same constructor arguments, same `state`
dictionary, same algorithm -- conjugate gradient on Hessian-vector products of the constraint, the
step size that meets the constraint to second order, backtracking line search -- with the arithmetic
in libb2g:

  Hessian-vector product   b2g_mlp_fvp (forward-mode pass + backward pass through the MLP kernels; exact
                           for the KL of the Gaussian MLP policy when the step starts at the old policy)
  CG vector algebra        b2g_cg_init / b2g_cg_step / b2g_cg_sanitize / b2g_cg_descent (scalars stay on
                           the device: one solve is enqueued without a host synchronisation)
  line search              b2g_line_search_point + one forward pass (loss and KL together) per point

Used through `OptimizerWrapper((ConjugateGradientOptimizer, dict(max_constraint_value=0.01)), policy)`
exactly like the reference (trpo.py:62-66); `TRPO._train_policy` calls `step`.
"""
import warnings

import numpy as np
import torch

from garage_b200 import ops


class ConjugateGradientOptimizer:
    """Performs constrained optimization via backtracking line search (reference docstring :113-136).

    Args:
        params: the policy's parameters (kept for signature compatibility; the flat parameter vector
            of the policy module is what the kernels update).
        max_constraint_value (float): Maximum constraint value.
        cg_iters (int): The number of CG iterations used to calculate A^-1 g.
        max_backtracks (int): Max number of iterations for backtrack linesearch.
        backtrack_ratio (float): backtrack ratio for backtracking line search.
        hvp_reg_coeff (float): A small value so that A -> A + reg*I.
        accept_violation (bool): whether to accept the descent step if it violates the line search
            condition after exhausting all backtracking budgets.
    """

    def __init__(self, params, max_constraint_value, cg_iters=10, max_backtracks=15, backtrack_ratio=0.8,
                 hvp_reg_coeff=1e-5, accept_violation=False):
        self._params = params
        self._max_constraint_value = max_constraint_value
        self._cg_iters = cg_iters
        self._max_backtracks = max_backtracks
        self._backtrack_ratio = backtrack_ratio
        self._hvp_reg_coeff = hvp_reg_coeff
        self._accept_violation = accept_violation
        self.last_info = {}

    @property
    def state(self):
        """dict: The hyper-parameters of the optimizer (reference :180-190)."""
        return {
            'max_constraint_value': self._max_constraint_value,
            'cg_iters': self._cg_iters,
            'max_backtracks': self._max_backtracks,
            'backtrack_ratio': self._backtrack_ratio,
            'hvp_reg_coeff': self._hvp_reg_coeff,
            'accept_violation': self._accept_violation,
        }

    @state.setter
    def state(self, state):
        self._max_constraint_value = state.get('max_constraint_value', 0.01)
        self._cg_iters = state.get('cg_iters', 10)
        self._max_backtracks = state.get('max_backtracks', 15)
        self._backtrack_ratio = state.get('backtrack_ratio', 0.8)
        self._hvp_reg_coeff = state.get('hvp_reg_coeff', 1e-5)
        self._accept_violation = state.get('accept_violation', False)

    def __getstate__(self):
        return {'state': self.state}

    def __setstate__(self, state):
        if 'hvp_reg_coeff' not in state['state']:
            warnings.warn('Resuming ConjugateGradientOptimizer with lost state. '
                          'This behavior is fixed if pickling from garage>=2020.02.0.')
        self._params = None
        self.last_info = {}
        self.state = state['state']

    def zero_grad(self):
        """Gradients are produced and consumed inside the kernels."""

    def step(self, f_loss, f_constraint, *, flat, grad, f_Ax):
        """Take an optimization step (reference :141-178).

        Args:
            f_loss (callable): () -> float, the loss at the current parameters.
            f_constraint (callable): () -> float, the constraint value at the current parameters.
            flat (torch.Tensor): the flat float32 parameter vector to update in place.
            grad (torch.Tensor): gradient of the loss at `flat` (what `p.grad` holds in the reference).
            f_Ax (callable): (vec, out) -> out, the Hessian-vector product of the constraint.
        """
        dev, n = flat.device, flat.numel()
        f32 = dict(dtype=torch.float32, device=dev)
        x, r, p, z = (torch.empty(n, **f32) for _ in range(4))
        scal = torch.zeros(8, dtype=torch.float64, device=dev)
        # step direction: _conjugate_gradient(f_Ax, flat_loss_grads, cg_iters)
        ops.cg_init(grad, x, r, p, scal)
        for _ in range(self._cg_iters):
            f_Ax(p, z)
            ops.cg_step(z, x, r, p, scal)
        ops.cg_sanitize(x)                       # step_dir[step_dir.ne(step_dir)] = 0.
        f_Ax(x, z)
        descent = torch.empty(n, **f32)
        ops.cg_descent(x, z, self._max_constraint_value, descent, scal)
        self._backtracking_line_search(flat, descent, f_loss, f_constraint)
        s = scal.cpu().numpy()
        self.last_info.update(cg_iterations=int(s[4]), step_size=float(s[2]), xAx=float(s[3]))

    def _backtracking_line_search(self, flat, descent, f_loss, f_constraint):
        """Reference :218-260, with (loss, constraint) read back once per point."""
        prev = flat.clone()
        ratio_list = self._backtrack_ratio**np.arange(self._max_backtracks)
        loss_before = f_loss()
        loss = constraint_val = float('nan')
        used = None
        for k, ratio in enumerate(ratio_list):
            ops.line_search_point(prev, descent, ratio, flat)
            loss, constraint_val = f_loss(), f_constraint()
            used = k
            if loss < loss_before and constraint_val <= self._max_constraint_value:
                break
        rejected = False
        if ((np.isnan(loss) or np.isnan(constraint_val) or loss >= loss_before
             or constraint_val >= self._max_constraint_value) and not self._accept_violation):
            rejected = True
            self._log_violation(loss, constraint_val, loss_before)
            flat.copy_(prev)
        self.last_info.update(backtracks=used, loss_before=loss_before, loss=loss, constraint=constraint_val,
                              rejected=rejected)

    def _log_violation(self, loss, constraint_val, loss_before):
        msgs = ['Line search condition violated. Rejecting the step!']
        if np.isnan(loss):
            msgs.append('Violated because loss is NaN')
        if np.isnan(constraint_val):
            msgs.append('Violated because constraint is NaN')
        if loss >= loss_before:
            msgs.append('Violated because loss not improving')
        if constraint_val >= self._max_constraint_value:
            msgs.append('Violated because constraint is violated')
        self.last_info['messages'] = msgs
