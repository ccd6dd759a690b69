"""Trust Region Policy Optimization -- drop-in for `garage.torch.algos.TRPO`
(reference torch/algos/trpo.py:9-144): VPG whose policy step is the reference's
ConjugateGradientOptimizer (torch/optimizers/conjugate_gradient_optimizer.py) on the
unclipped likelihood-ratio surrogate (trpo.py:90-117), with the arithmetic in libb2g:

  loss + gradient           b2g_mlp_minibatch_grad (loss kind PPO with an infinite clip range:
                            min(r A, clip(r) A) = r A exactly) + b2g_mlp_reduce
  Hessian-vector products   b2g_mlp_fvp (Fisher-vector product: exact for the KL constraint of
                            vpg.py:381-406 because the step starts at the old policy)
  CG / step size            b2g_cg_*                 (conjugate_gradient_optimizer.py:69-110, 141-178)
  line search               b2g_line_search_point + b2g_mlp_eval (loss and KL in one pass, :218-260)
  value function            the Adam minibatch path of VPG (fused loop when it applies)
"""
import math

import torch

from garage_b200 import ops
from garage_b200.torch.algos.vpg import VPG
from garage_b200.torch.optimizers import ConjugateGradientOptimizer, OptimizerWrapper


class TRPO(VPG):
    """See the reference docstring (trpo.py:10-45) for the arguments."""
    _loss_kind = ops.LOSS_PPO
    _clip = math.inf  # surrogate = likelihood_ratio * advantages, no clipping (trpo.py:109-117)

    def __init__(self, env_spec, policy, value_function, sampler, policy_optimizer=None,
                 vf_optimizer=None, num_train_per_epoch=1, discount=0.99, gae_lambda=0.98,
                 center_adv=True, positive_adv=False, policy_ent_coeff=0.0,
                 use_softplus_entropy=False, stop_entropy_gradient=False,
                 entropy_method='no_entropy', **kwargs):
        # defaults of the reference, trpo.py:62-72
        if policy_optimizer is None:
            policy_optimizer = OptimizerWrapper(
                (ConjugateGradientOptimizer, dict(max_constraint_value=0.01)), policy)
        if vf_optimizer is None:
            vf_optimizer = OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), value_function,
                                            max_optimization_epochs=10, minibatch_size=64)
        if policy_optimizer.cg is None:
            raise NotImplementedError('garage_b200 TRPO needs a ConjugateGradientOptimizer policy optimiser '
                                      '(with Adam it is VPG on the unclipped surrogate: use PPO/VPG)')
        if policy_optimizer.minibatch_size is not None or policy_optimizer.max_optimization_epochs != 1:
            # The Fisher-vector product equals the Hessian of KL(old || new) only at new == old, i.e. for the
            # first step after the rollout; the reference's defaults (one full-batch step) are that case.
            raise NotImplementedError('TRPO policy optimiser: one full-batch step per iteration '
                                      '(max_optimization_epochs=1, minibatch_size=None, the reference default)')
        super().__init__(env_spec=env_spec, policy=policy, value_function=value_function,
                         sampler=sampler, policy_optimizer=policy_optimizer, vf_optimizer=vf_optimizer,
                         num_train_per_epoch=num_train_per_epoch, discount=discount,
                         gae_lambda=gae_lambda, center_adv=center_adv, positive_adv=positive_adv,
                         policy_ent_coeff=policy_ent_coeff, use_softplus_entropy=use_softplus_entropy,
                         stop_entropy_gradient=stop_entropy_gradient, entropy_method=entropy_method,
                         **kwargs)

    # ------------------------------------------------------------------ train --
    def _train(self, f):
        """vpg.py:230-248 with trpo.py:119-144 as the policy step."""
        self._train_policy_cg(f)
        vf, opt = self._value_function, self._vf_optimizer
        if self._update_mode != 'stepwise' and ops.mlp_path() != 1 and ops.train_loop_kind([vf.mlp_shape], self._world()):
            ops.mlp_train_loop([self._loop_net(opt, vf, f, False)])
        else:
            self._run_minibatches(opt, vf, f, policy=False)

    def _all_reduce(self, t):
        import torch.distributed as dist
        dist.all_reduce(t, group=self._pg)
        return t

    def _train_policy_cg(self, f):
        """`TRPO._train_policy` (trpo.py:119-144): gradient of the loss, then
        `ConjugateGradientOptimizer.step(f_loss, f_constraint)`."""
        pol, cg = self.policy, self._policy_optimizer.cg
        shape, flat = pol.mlp_shape, pol.flat
        dev = flat.device
        world = self._world()
        B = f.B
        if world > 1:  # the batch of the job: per-rank sums are weighted by 1 / B_global
            B = int(self._all_reduce(torch.tensor([f.B], dtype=torch.int64, device=dev)).item())
        kw = self._policy_kwargs()
        data = dict(act=f.act, adv=f.adv, old_logp=f.old_logp)
        nb = ops.mlp_minibatch_grad(shape, flat, f.obs, **data, **kw)
        grad = torch.empty(shape.P, dtype=torch.float32, device=dev)
        ops.mlp_reduce(shape, nb, 1.0 / B, grad_out=grad)
        if world > 1:
            self._all_reduce(grad)
        scratch = torch.empty(f.B * shape.A, dtype=torch.float32, device=dev)
        mod = pol._module
        stats = torch.empty(8, dtype=torch.float64, device=dev)
        last = {}

        def f_Ax(vec, out):
            if world == 1:
                return ops.mlp_fvp(shape, flat, vec, f.obs, min_log_std=mod.min_log_std, max_log_std=mod.max_log_std,
                                   reg_coeff=cg._hvp_reg_coeff, out=out, scratch=scratch)
            ops.mlp_fvp(shape, flat, vec, f.obs, min_log_std=mod.min_log_std, max_log_std=mod.max_log_std,
                        reg_coeff=0.0, out=out, scratch=scratch)
            out.mul_(f.B / B)
            self._all_reduce(out)
            return out.add_(vec, alpha=cg._hvp_reg_coeff)

        def f_loss():
            # one forward pass gives the loss and the KL of the current parameters; f_constraint() is always
            # called right after f_loss() at the same parameters (conjugate_gradient_optimizer.py:238-239)
            n = ops.mlp_eval(shape, flat, f.obs, old_mean=f.old_mean, old_log_std=f.old_log_std, **data, **kw)
            ops.mlp_reduce(shape, n, 1.0, stats_out=stats)
            s = stats[0:2].clone()
            if world > 1:
                self._all_reduce(s)
            s = s.cpu().numpy()
            last['kl'] = float(s[1]) / B
            return float(s[0]) / B

        def f_constraint():
            return last['kl']

        cg.step(f_loss, f_constraint, flat=flat, grad=grad, f_Ax=f_Ax)
