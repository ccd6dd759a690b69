"""Proximal Policy Optimization -- drop-in for `garage.torch.algos.PPO`
(reference torch/algos/ppo.py:9-132): VPG with the clipped surrogate, whose forward
AND backward run inside libb2g's fused minibatch kernel (b2g_mlp_minibatch_grad,
loss_kind = PPO).  `old_ll` is the rollout log-prob: exact, because the reference's
`_old_policy` is the policy the batch was sampled with (vpg.py:107, 201; quirk Q8)."""
import torch

from garage_b200 import ops
from garage_b200.torch.algos.vpg import VPG
from garage_b200.torch.optimizers import OptimizerWrapper


class PPO(VPG):
    _loss_kind = ops.LOSS_PPO

    def __init__(self, env_spec, policy, value_function, sampler, policy_optimizer=None,
                 vf_optimizer=None, lr_clip_range=2e-1, num_train_per_epoch=1, discount=0.99,
                 gae_lambda=0.97, center_adv=True, positive_adv=False, policy_ent_coeff=0.0,
                 use_softplus_entropy=False, stop_entropy_gradient=False,
                 entropy_method='no_entropy', **kwargs):
        # defaults of the reference, ppo.py:65-76
        if policy_optimizer is None:
            policy_optimizer = OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), policy,
                                                max_optimization_epochs=10, minibatch_size=64)
        if vf_optimizer is None:
            vf_optimizer = OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), value_function,
                                            max_optimization_epochs=10, minibatch_size=64)
        super().__init__(env_spec=env_spec, policy=policy, value_function=value_function,
                         sampler=sampler, policy_optimizer=policy_optimizer,
                         vf_optimizer=vf_optimizer, num_train_per_epoch=num_train_per_epoch,
                         discount=discount, gae_lambda=gae_lambda, center_adv=center_adv,
                         positive_adv=positive_adv, policy_ent_coeff=policy_ent_coeff,
                         use_softplus_entropy=use_softplus_entropy,
                         stop_entropy_gradient=stop_entropy_gradient,
                         entropy_method=entropy_method, **kwargs)
        self._lr_clip_range = lr_clip_range
        self._clip = float(lr_clip_range)
