"""Host-side logic of the multi-GPU (env-sharded, data-parallel) path.

Rank r of G owns envs [r*N/G, (r+1)*N/G) end to end (rollout, GAE, packing: no
communication).  The exchanges are: one gradient all-reduce per minibatch step
(NCCL, in VPG._run_minibatches), the global advantage moments (below), and the
logging scalars once per iteration."""
import torch
import torch.distributed as dist


def world_size(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group)
    return 1


def combine_moments(stats, n, group=None):
    """Global mean / unbiased variance / min from per-rank (mean, var, min) over n local
    values each (Chan et al. parallel variance): one all_gather of 4 doubles.
    `stats` is the double[3] tensor written by b2g_moments; works on any device/backend."""
    world = world_size(group)
    if world == 1:
        return stats
    mine = torch.stack([torch.full((), float(n), dtype=torch.float64, device=stats.device),  # fill kernel: capturable
                        stats[0], stats[1] * (n - 1), stats[2]])
    allv = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allv, mine, group=group)
    a = torch.stack(allv)
    ntot = a[:, 0].sum()
    mean = (a[:, 0] * a[:, 1]).sum() / ntot
    m2 = (a[:, 2] + a[:, 0] * (a[:, 1] - mean)**2).sum()
    return torch.stack([mean, m2 / (ntot - 1), a[:, 3].min()]).contiguous()


def shard_envs(n_envs_total, rank, world):
    """(n_envs on this rank, global id of its first env); the remainder goes to the
    lowest ranks so that every env id in [0, n_envs_total) is owned exactly once."""
    base, extra = divmod(int(n_envs_total), int(world))
    n = base + (1 if rank < extra else 0)
    off = rank * base + min(rank, extra)
    return n, off


def reduce_logs(s, B, n_eps=None, group=None):
    """The logging scalars of one iteration over all ranks, equal to what the reference's
    log_performance (_functions.py:233-275) reports for the union of the shards: sums for the
    loss / KL / entropy statistics and the sample count; the return statistics are combined from
    per-rank (n_eps, mean, population std) -- episode-count weighted means, the GLOBAL standard
    deviation (Chan's pairwise formula, not a mean of per-rank stds), max / min of the extremes.
    `s`: float64 tensor [>=40] laid out like VPG._enqueue_iteration's stats (s[32:39] = mean
    discounted return, mean / std / max / min return, termination rate, success rate).
    Returns (s[:40] reduced, B total, n_eps total) as tensors."""
    world = world_size(group)
    n = float(n_eps) if n_eps is not None else 1.0
    dd = dict(dtype=torch.float64, device=s.device)
    if world == 1:
        return s[:40], torch.full((), float(B), **dd), torch.full((), n, **dd)
    perf = s[32:39]
    mean, std = perf[1], perf[2]
    # [0:32] sums | n_eps | n*disc | n*mean | n*(std^2 + mean^2) | n*term | n*succ | pad (s[39]) | B
    w = torch.stack([torch.full((), n, **dd), n * perf[0], n * mean, n * (std * std + mean * mean),
                     n * perf[5], n * perf[6], s[39], torch.full((), float(B), **dd)])
    t = torch.cat([s[:32], w])
    dist.all_reduce(t, group=group)
    ext = torch.stack([perf[3], -perf[4]])
    dist.all_reduce(ext, op=dist.ReduceOp.MAX, group=group)
    nt = t[32]
    gmean = t[34] / nt
    gstd = torch.sqrt(torch.clamp(t[35] / nt - gmean * gmean, min=0.0))
    out = torch.cat([t[:32], torch.stack([t[33] / nt, gmean, gstd, ext[0], -ext[1], t[36] / nt, t[37] / nt, t[38]])])
    return out, t[39], nt


class PeerComm:
    """Peer-mapped exchange buffers for the in-kernel gradient all-reduce of the fused training
    loop (b2g_mlp_train_loop_dist): every rank allocates one buffer (cudaMalloc inside libb2g),
    the CUDA IPC handles travel through torch.distributed, and each rank maps all peers' buffers.
    One node only (NVLink / NVSwitch peer access); collective constructor."""

    def __init__(self, device, group=None):
        import ctypes

        from garage_b200 import _lib
        lib = _lib.load()
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > _lib.MAX_RANKS:
            raise NotImplementedError('at most {} ranks'.format(_lib.MAX_RANKS))
        self._lib, self._own, self._peers = lib, ctypes.c_void_p(), []
        handle = ctypes.create_string_buffer(64)
        with torch.cuda.device(device):
            _lib.check(lib.b2g_comm_alloc(int(lib.b2g_train_comm_bytes()), ctypes.byref(self._own), handle),
                       'b2g_comm_alloc')
            mine = torch.tensor(list(handle.raw), dtype=torch.uint8, device=device)
            allh = [torch.empty_like(mine) for _ in range(self.world)]
            dist.all_gather(allh, mine, group=group)
            self._c = _lib.TrainComm()
            self._c.world, self._c.rank = self.world, self.rank
            for r, h in enumerate(allh):
                if r == self.rank:
                    self._c.peer_base[r] = self._own.value
                    continue
                p = ctypes.c_void_p()
                _lib.check(lib.b2g_comm_open(bytes(h.cpu().tolist()), ctypes.byref(p)), 'b2g_comm_open')
                self._c.peer_base[r] = p.value
                self._peers.append(p)
        dist.barrier(group=group)  # every buffer is mapped (and still zero) before anyone writes

    def struct(self):
        return self._c

    PEER_GATHER, PEER_MOMENTS, PEER_LOGS = 0, 1, 2

    def reduce_f64(self, op, t_in, out, scalar0=0.0, scalar1=0.0, n=0):
        """b2g_peer_reduce_f64: small all-gather + combine of float64 scalars over the peer-mapped
        buffers (no NCCL, no host synchronisation, CUDA-graph capturable).  Collective: every rank
        calls it in the same order."""
        import ctypes

        from garage_b200 import _lib
        if t_in.dtype != torch.float64 or out.dtype != torch.float64 or not t_in.is_cuda or not out.is_cuda:
            raise TypeError('reduce_f64 needs float64 CUDA tensors')
        if not (t_in.is_contiguous() and out.is_contiguous()):
            raise ValueError('reduce_f64 needs contiguous tensors')
        with torch.cuda.device(t_in.device):
            _lib.call('b2g_peer_reduce_f64', ctypes.byref(self._c), int(op), t_in.data_ptr(), int(n), float(scalar0),
                      float(scalar1), out.data_ptr(), torch.cuda.current_stream(t_in.device).cuda_stream)
        return out

    def close(self):
        from garage_b200 import _lib
        for p in self._peers:
            _lib.check(self._lib.b2g_comm_close(p), 'b2g_comm_close')
        self._peers = []
        if self._own:
            _lib.check(self._lib.b2g_comm_free(self._own), 'b2g_comm_free')
            self._own = None
