"""Per-stage device times of one PPO iteration at the bench configuration, on 1..N GPUs
(run directly for one GPU or under torchrun): where does a multi-GPU iteration spend its time?
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29513 tools/dist_time.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    rank, local, world = (int(os.environ.get(k, d)) for k, d in (('RANK', 0), ('LOCAL_RANK', 0), ('WORLD_SIZE', 1)))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    from garage_b200.envs import PendulumEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.algos import PPO
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    n_envs, T = 4096, 200
    torch.manual_seed(1)
    env = PendulumEnv(max_episode_length=T)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
    vf = GaussianMLPValueFunction(env.spec, hidden_sizes=(64, 64))
    sampler = GPUSampler(policy, env, max_episode_length=T, n_envs=n_envs, seed=1234, env_id_offset=rank * n_envs)
    mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), m, max_optimization_epochs=10,
                                    minibatch_size=32768 * world)
    algo = PPO(env.spec, policy, vf, sampler, policy_optimizer=mk(policy), vf_optimizer=mk(vf), discount=0.99,
               gae_lambda=0.95)
    names = ['rollout', 'prepare', 'eval0', 'train', 'eval1', 'perf']
    acc = [0.0] * len(names)
    iters = 8
    for it in range(iters + 2):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(names) + 1)]
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev[0].record()
        eps = sampler.obtain_samples(it, n_envs * T, policy)
        ev[1].record()
        f = algo._prepare(eps)
        ev[2].record()
        stats = torch.zeros(48, dtype=torch.float64, device='cuda')
        algo._evaluate(f, stats[0:16])
        ev[3].record()
        algo._train(f)
        ev[4].record()
        algo._evaluate(f, stats[16:32])
        ev[5].record()
        algo._performance(f, stats[32:39])
        ev[6].record()
        torch.cuda.synchronize()
        if it >= 2:
            for i in range(len(names)):
                acc[i] += ev[i].elapsed_time(ev[i + 1]) / iters
    if rank == 0:
        print('world', world, 'fused', algo._fused_applies(), 'comm', algo._comm is not None,
              ' '.join('{} {:.3f} ms'.format(n, t) for n, t in zip(names, acc)), 'total {:.3f} ms'.format(sum(acc)), flush=True)
    if world > 1:
        dist.barrier()
        sys.stdout.flush()
        os._exit(0)


if __name__ == '__main__':
    main()
