"""Multi-GPU parity of the fused loop's in-kernel all-reduce (run under torchrun, one rank per GPU):
fused (peer-memory all-reduce inside the kernel) vs step-wise (NCCL all-reduce per minibatch) must
leave the same parameters on every rank, and all ranks must hold identical parameters.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tools/dist_check.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def run(mode, rank, world, n_envs=64, epochs=3, mb=1024, iters=2, use_peer=None):
    from garage_b200.envs import PendulumEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.algos import PPO
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    torch.manual_seed(1)
    env = PendulumEnv(max_episode_length=40)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
    vf = GaussianMLPValueFunction(env.spec, hidden_sizes=(64, 64))
    sampler = GPUSampler(policy, env, max_episode_length=40, n_envs=n_envs, seed=1234,
                         env_id_offset=rank * n_envs)
    mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), m, max_optimization_epochs=epochs,
                                    minibatch_size=mb)
    algo = PPO(env.spec, policy, vf, sampler, policy_optimizer=mk(policy), vf_optimizer=mk(vf), discount=0.99,
               gae_lambda=0.95, update_mode=mode)
    algo._use_peer = use_peer  # None: peer-memory exchange of moments / logs; False: NCCL + host arithmetic
    for it in range(iters):
        eps = sampler.obtain_samples(it, n_envs * 40, policy)
        algo._train_once(it, eps)
    torch.cuda.synchronize()
    if rank == 0:
        print(mode, 'steps', int(algo._policy_optimizer.state()[2].item()), int(algo._vf_optimizer.state()[2].item()),
              'fused applies', algo._fused_applies(), 'comm', algo._comm is not None, '|policy|', float(policy.flat.norm()), '|vf|', float(vf.flat.norm()), flush=True)
    return policy.flat.clone(), vf.flat.clone(), dict(algo.last_stats)


def run_ragged(rank, world, n_envs=48, epochs=2, mb=700, iters=3):
    """CartPole terminates early, so the ranks' shards differ in length from iteration to iteration
    (and a shard may happen to be full-length on one rank only): update_mode='auto' must take the
    same path on every rank, issue the same collectives and leave identical parameters."""
    from garage_b200.envs import CartPoleEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.algos import PPO
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    torch.manual_seed(2)
    env = CartPoleEnv(max_episode_length=30)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
    vf = GaussianMLPValueFunction(env.spec, hidden_sizes=(64, 64))
    sampler = GPUSampler(policy, env, max_episode_length=30, n_envs=n_envs, seed=77, env_id_offset=rank * n_envs)
    mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), m, max_optimization_epochs=epochs,
                                    minibatch_size=mb)
    algo = PPO(env.spec, policy, vf, sampler, policy_optimizer=mk(policy), vf_optimizer=mk(vf), discount=0.99,
               gae_lambda=0.95)
    sizes = []
    for it in range(iters):
        eps = sampler.obtain_samples(it, n_envs * 30, policy)
        sizes.append(int(eps.n_steps))
        algo._train_once(it, eps)
    torch.cuda.synchronize()
    return policy.flat.clone(), vf.flat.clone(), sizes, int(algo._policy_optimizer.state()[2].item())


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


def main():
    rank, local, world = int(os.environ['RANK']), int(os.environ['LOCAL_RANK']), int(os.environ['WORLD_SIZE'])
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    pf, vf_, lp = run('fused', rank, world)
    ps, vs, ln = run('stepwise', rank, world, use_peer=False)
    log_err = max(abs(float(lp[k]) - float(ln[k])) / max(1.0, abs(float(ln[k]))) for k in ln if ln[k] == ln[k])
    errs = torch.tensor([rel(pf, ps), rel(vf_, vs), log_err], device='cuda')
    # identical bits on every rank
    ref = pf.clone()
    dist.broadcast(ref, 0)
    same = torch.tensor([float(torch.equal(ref, pf))], device='cuda')
    dist.all_reduce(errs, op=dist.ReduceOp.MAX)
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    # ragged shards (CartPole): same collectives on every rank, identical parameters afterwards
    rp, rv, sizes, steps = run_ragged(rank, world)
    allsz = [None] * world
    dist.all_gather_object(allsz, (sizes, steps))
    refp = rp.clone()
    dist.broadcast(refp, 0)
    same_r = torch.tensor([float(torch.equal(refp, rp))], device='cuda')
    dist.all_reduce(same_r, op=dist.ReduceOp.MIN)
    ragged_ok = bool(same_r.item()) and len({s_[1] for s_ in allsz}) == 1 and bool(torch.isfinite(rp).all())
    if rank == 0:
        print('ragged CartPole shards per rank (env steps per iteration, optimiser steps):', allsz,
              '-> identical parameters on all ranks:', bool(same_r.item()))
    if rank == 0:
        print('world', world, 'fused vs stepwise rel err policy {:.3e} vf {:.3e}; logged scalars peer-memory vs NCCL {:.3e}; '
              'ranks bit-identical: {}'.format(float(errs[0]), float(errs[1]), float(errs[2]), bool(same.item())))
        ok = float(errs.max()) < 1e-5 and bool(same.item()) and ragged_ok
        print('DIST_CHECK', 'OK' if ok else 'FAIL')
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
