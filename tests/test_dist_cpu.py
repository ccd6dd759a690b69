"""world_size-2 gloo tests (CPU) of the host-side logic of the env-sharded multi-GPU path:
global advantage moments, log reduction, env sharding."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from garage_b200.torch.algos import _dist
    rng = np.random.default_rng(0)
    full = rng.standard_normal(1000) * 3 + 2
    n, off = _dist.shard_envs(1000, rank, world)
    mine = torch.from_numpy(full[off:off + n])
    local = torch.stack([mine.mean(), mine.var(), mine.min()]).to(torch.float64)
    glob = _dist.combine_moments(local, n)
    # per-rank log_performance statistics of UNEQUAL numbers of episodes (terminating envs)
    returns = np.random.default_rng(7).standard_normal(50) * 4 - 3
    mine_r = returns[:20] if rank == 0 else returns[20:]
    s = torch.zeros(48, dtype=torch.float64)
    s[0] = float(rank + 1)          # a sum statistic
    s[32] = 0.5 * mine_r.mean()     # mean discounted return of the shard
    s[33], s[34] = mine_r.mean(), mine_r.std()
    s[35], s[36] = mine_r.max(), mine_r.min()
    s[37] = (mine_r > 0).mean()     # termination rate
    s[38] = float('nan')            # env without 'success'
    red, B, n_eps = _dist.reduce_logs(s, 100 * (rank + 1), len(mine_r))
    q.put((rank, glob.numpy(), red.numpy(), float(B), n, off, float(n_eps)))
    dist.destroy_process_group()


def test_two_rank_moments_logs_and_sharding():
    world = 2
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.default_rng(0)
    full = rng.standard_normal(1000) * 3 + 2
    covered = []
    returns = np.random.default_rng(7).standard_normal(50) * 4 - 3
    for rank, glob, red, B, n, off, n_eps in res:
        assert abs(glob[0] - full.mean()) < 1e-12
        assert abs(glob[1] - full.var(ddof=1)) < 1e-10
        assert glob[2] == full.min()
        assert red[0] == 3.0 and B == 300.0 and n_eps == 50.0   # sums over ranks
        # what log_performance (_functions.py:262-273) reports for the union of the shards
        assert abs(red[33] - returns.mean()) < 1e-12            # episode-weighted, not a mean of means
        assert abs(red[34] - returns.std()) < 1e-12             # the global std
        assert abs(red[32] - 0.5 * (20 * returns[:20].mean() + 30 * returns[20:].mean()) / 50) < 1e-12
        assert red[35] == returns.max() and red[36] == returns.min()
        assert abs(red[37] - (returns > 0).mean()) < 1e-12
        assert np.isnan(red[38])
        covered += list(range(off, off + n))
    assert sorted(covered) == list(range(1000))


def test_shard_envs_remainder():
    from garage_b200.torch.algos._dist import shard_envs
    parts = [shard_envs(10, r, 4) for r in range(4)]
    assert [p[0] for p in parts] == [3, 3, 2, 2]
    assert [p[1] for p in parts] == [0, 3, 6, 8]
    assert shard_envs(65536, 7, 8) == (8192, 57344)
