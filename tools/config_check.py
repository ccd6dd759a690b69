"""BASELINE.json configs [0] and [2] at full size through the public API (eager iterations, CUDA
events): do they run, which update path do they take, and how fast.
    python tools/config_check.py
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def run(name, env, algo_cls, hidden, n_envs, T, batch, mb, epochs, iters=6):
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    torch.manual_seed(0)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=hidden)
    vf = GaussianMLPValueFunction(env.spec, hidden_sizes=hidden)
    sampler = GPUSampler(policy, env, max_episode_length=T, n_envs=n_envs, seed=1)
    mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), m, max_optimization_epochs=epochs,
                                    minibatch_size=mb)
    algo = algo_cls(env.spec, policy, vf, sampler, policy_optimizer=mk(policy), vf_optimizer=mk(vf), discount=0.99)
    steps = 0
    for it in range(iters + 2):
        if it == 2:
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            steps = 0
        eps = sampler.obtain_samples(it, batch, policy)
        algo._train_once(it, eps)
        steps += eps.n_steps
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print('{}: {} env-steps in {:.1f} ms/iteration -> {:.2f} M env-steps/s; fused loop: {}; last batch {} episodes / {} steps; '
          'AverageReturn {:.2f}'.format(name, steps, 1e3 * dt / iters, steps / dt / 1e6, algo._fused_applies(),
                                        eps.n_episodes, eps.n_steps, algo.last_stats['Evaluation/AverageReturn']), flush=True)


def main():
    from garage_b200.envs import CartPoleEnv, PointEnv
    from garage_b200.torch.algos import PPO, VPG
    run('configs[0] PPO PointEnv (32,32), batch 1000 (1 env copy like LocalSampler)', PointEnv(max_episode_length=100),
        PPO, (32, 32), 1, 100, 1000, 64, 10, iters=3)
    run('configs[0] at 4,096 envs', PointEnv(max_episode_length=100), PPO, (32, 32), 4096, 100, 4096 * 100, 32768, 10)
    run('configs[2] VPG CartPole 16,384 envs x 200 steps, full-batch update', CartPoleEnv(max_episode_length=200), VPG,
        (64, 64), 16384, 200, 16384 * 200, None, 1)
    run('configs[2] with 32,768-row minibatches, 10 epochs', CartPoleEnv(max_episode_length=200), VPG,
        (64, 64), 16384, 200, 16384 * 200, 32768, 10)


if __name__ == '__main__':
    main()
