"""Two eager PPO iterations of the bench workload (BASELINE configs[1]); the second one is the
one ncu captures (`-s <launches of iteration 1>`).  No timing is reported from here."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--iters', type=int, default=2)
    ap.add_argument('--epochs', type=int, default=10)
    ap.add_argument('--n-envs', type=int, default=4096)
    ap.add_argument('--minibatch', type=int, default=37888)
    p = ap.parse_args()
    a = argparse.Namespace(n_envs=p.n_envs, horizon=200, hidden=64, epochs=p.epochs, minibatch=p.minibatch)
    algo, policy, vf, sampler = bench.build_algo(a, 0, 1, torch.device('cuda:0'))
    B = a.n_envs * a.horizon
    for i in range(p.iters):
        n = bench.count_launches(lambda: algo._train_once(i, sampler.obtain_samples(i, B, policy)))
        torch.cuda.synchronize()
        print('iteration', i, 'launches', n, 'return', algo.last_stats['Evaluation/AverageReturn'])


if __name__ == '__main__':
    main()
