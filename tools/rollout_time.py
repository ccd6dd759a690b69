"""Rollout kernel alone: time per launch for each built-in env (4,096 envs x 200 steps, hidden 64;
HID / NENVS in the environment change the hidden size and the number of envs)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402


def main():
    hid, n = int(os.environ.get('HID', '64')), int(os.environ.get('NENVS', '4096'))
    from garage_b200.envs import CartPoleEnv, PendulumEnv, PointEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.policies import GaussianMLPPolicy
    for name, env in (('pendulum', PendulumEnv(max_episode_length=200)), ('point', PointEnv(max_episode_length=200)),
                      ('cartpole', CartPoleEnv(max_episode_length=200))):
        policy = GaussianMLPPolicy(env.spec, hidden_sizes=(hid, hid))
        sampler = GPUSampler(policy, env, max_episode_length=200, n_envs=n, seed=1)
        for _ in range(2):
            sampler.obtain_samples(0, n * 200, policy)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(5):
            sampler.obtain_samples(0, n * 200, policy)
        e1.record()
        torch.cuda.synchronize()
        us = 1e3 * e0.elapsed_time(e1) / 5
        print('{} hidden {} envs {}: obtain_samples {:.1f} us (rollout + packing), {:.2f} G env-steps/s'.format(
            name, hid, n, us, n * 200 / us / 1e3), flush=True)


if __name__ == '__main__':
    main()
