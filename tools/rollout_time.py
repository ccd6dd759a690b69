"""Rollout kernel alone: time per launch for each built-in env (4,096 envs x 200 steps, hidden 64)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    from garage_b200.envs import CartPoleEnv, PendulumEnv, PointEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.policies import GaussianMLPPolicy
    for name, env in (('pendulum', PendulumEnv(max_episode_length=200)), ('point', PointEnv(max_episode_length=200)),
                      ('cartpole', CartPoleEnv(max_episode_length=200))):
        policy = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
        sampler = GPUSampler(policy, env, max_episode_length=200, n_envs=4096, seed=1)
        for _ in range(2):
            sampler.obtain_samples(0, 4096 * 200, policy)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            sampler.obtain_samples(0, 4096 * 200, policy)
        e1.record()
        torch.cuda.synchronize()
        print('{}: obtain_samples {:.1f} us (rollout + packing)'.format(name, 1e3 * e0.elapsed_time(e1) / 5), flush=True)


if __name__ == '__main__':
    main()
