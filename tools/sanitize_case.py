"""Small-shape invocations of every kernel family for compute-sanitizer (memcheck /
racecheck per kernel).  One line per case: `CASE <name> ok` or `CASE <name> FAILED <error>`.

    compute-sanitizer --tool memcheck  python tools/sanitize_case.py
    compute-sanitizer --tool racecheck python tools/sanitize_case.py
"""
import math
import os
import sys
import traceback

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from garage_b200 import ops  # noqa: E402

dev = torch.device('cuda:0')


def algo_case(env_cls, hidden, cls_name='PPO', n_envs=32, T=20, mb=256, epochs=2, mode='auto'):
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch import algos
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    torch.manual_seed(0)
    env = env_cls(max_episode_length=T)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=hidden)
    vf = GaussianMLPValueFunction(env.spec, hidden_sizes=hidden)
    sampler = GPUSampler(policy, env, max_episode_length=T, n_envs=n_envs, seed=1)
    kw = {}
    if cls_name != 'TRPO':
        mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), m, max_optimization_epochs=epochs, minibatch_size=mb)
        kw = dict(policy_optimizer=mk(policy), vf_optimizer=mk(vf), update_mode=mode)
    algo = getattr(algos, cls_name)(env.spec, policy, vf, sampler, **kw)
    for it in range(2):
        eps = sampler.obtain_samples(it, n_envs * T, policy)
        r = algo._train_once(it, eps)
    assert np.isfinite(r)


def gae_case(T, N):
    g = torch.Generator(device=dev).manual_seed(0)
    rew, val = torch.randn(T, N, device=dev, generator=g), torch.randn(T, N, device=dev, generator=g)
    st = torch.ones(T, N, dtype=torch.uint8, device=dev)
    st[torch.rand(T, N, device=dev, generator=g) < 0.05] = 2
    st[T - 1] = 3
    a, r = ops.gae_time_major(rew, val, st, discount=0.99, gae_lambda=0.95)
    b = ops.gae_episode_major(val.T.contiguous(), rew.T.contiguous(), discount=0.99, gae_lambda=0.95)
    ops.discount_cumsum_f64(rew.T.contiguous().double(), 0.99)
    assert bool(torch.isfinite(a).all()) and bool(torch.isfinite(b).all())


def fvp_case():
    from tests.test_mlp_gpu import make_case
    for (O_, H, A) in [(6, 128, 3), (3, 32, 2)]:
        flat, obs, *_ = make_case(O_, H, A, 300, seed=1)
        sh = ops.MlpShape(O_, H, A, dev)
        t = lambda a: torch.from_numpy(a).to(dev)
        v = torch.randn(sh.P, device=dev)
        ops.mlp_fvp(sh, t(flat), v, t(obs), min_log_std=math.log(1e-6))


def gridworld_case():
    from garage_b200.envs import GridWorldEnv
    from garage_b200.np.policies import ScriptedPolicy
    from garage_b200.sampler import GPUSampler
    env = GridWorldEnv(desc='4x4', max_episode_length=16)
    pol = ScriptedPolicy(scripted_actions=[2, 2, 1, 0, 3, 1, 1, 1, 2, 2, 1, 1, 1, 2, 2, 1])
    s = GPUSampler(pol, env, max_episode_length=16, n_envs=8)
    s.obtain_samples(0, 100, pol)


def main():
    from garage_b200.envs import CartPoleEnv, PendulumEnv, PointEnv
    cases = [
        ('rollout+pack+gae+tcgen05 loop (Pendulum 64x64, PPO fused)', lambda: algo_case(PendulumEnv, (64, 64))),
        ('tcgen05 step-wise kernels (Pendulum 64x64, PPO stepwise)', lambda: algo_case(PendulumEnv, (64, 64), mode='stepwise')),
        ('ragged episodes + tcgen05 loop (CartPole 64x64, VPG)', lambda: algo_case(CartPoleEnv, (64, 64), 'VPG')),
        ('general loop, one CTA per network (PointEnv 32x32, minibatch 64)', lambda: algo_case(PointEnv, (32, 32), mb=64)),
        ('general loop, several CTAs (PointEnv 32x32, minibatch 512)', lambda: algo_case(PointEnv, (32, 32), mb=512)),
        ('FFMA step-wise kernels (PointEnv 32x32)', lambda: algo_case(PointEnv, (32, 32), mode='stepwise')),
        ('TRPO: fvp + cg + line search (Pendulum 64x64)', lambda: algo_case(PendulumEnv, (64, 64), 'TRPO')),
        ('general kernels: fvp hidden 128 / 32', fvp_case),
        ('gae time-major sequential + split, episode-major, discount_cumsum', lambda: (gae_case(50, 4096), gae_case(200, 64), gae_case(33, 40000))),
        ('gridworld rollout', gridworld_case),
    ]
    only = sys.argv[1:]
    for name, fn in cases:
        if only and not any(o in name for o in only):
            continue
        try:
            fn()
            torch.cuda.synchronize()
            print('CASE', name, 'ok', flush=True)
        except Exception as e:  # keep going: one report per kernel family
            print('CASE', name, 'FAILED', repr(e)[:300], flush=True)
            traceback.print_exc()


if __name__ == '__main__':
    main()
