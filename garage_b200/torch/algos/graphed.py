"""Whole-iteration CUDA graph: rollout -> pack -> baselines -> GAE -> normalise ->
all minibatch steps -> logging reductions, captured once and replayed.

Valid when the batch shape is static (an env that only ends at the time limit, e.g.
Pendulum) and the shuffle runs on the device: every counter that must advance between
replays (Philox step counter, reset episode counters, Adam step) lives in device memory
and is advanced by the captured kernels themselves."""
import torch


class GraphedIteration:
    """`run(itr)` == `eps = sampler.obtain_samples(...); algo._train_once(itr, eps)`.

    Args:
        algo (VPG | PPO): built with a `GPUSampler` sharing `algo.policy`, `shuffle='device'`.
        batch_size (int): `num_samples` per iteration.
        warmup (int): eager iterations before capture (lazy allocations, module loading).
    """

    def __init__(self, algo, batch_size, warmup=1, host_params=None):
        env = algo._sampler._env
        if not getattr(env, 'never_terminates', False):
            raise ValueError('GraphedIteration needs static shapes: an env that only ends at '
                             'max_episode_length (e.g. PendulumEnv)')
        if algo._shuffle != 'device':
            raise ValueError("GraphedIteration needs shuffle='device'")
        if algo._world() > 1 and not algo._fused_applies():
            raise ValueError('multi-GPU iterations are captured only with the fused training loop '
                             '(the step-wise path all-reduces between kernels)')
        self._algo, self._batch_size = algo, int(batch_size)
        self._graph = None
        self._stats = None
        self._meta = None
        self._warmup = warmup
        self._eager_done = 0
        self.eps = None  # the (static) DeviceEpisodeBatch of the captured iteration
        # host I/O inside the graph (the end-to-end form of an iteration): `host_params` is a dict of PINNED
        # CPU tensors with the policy's state_dict keys -- copied H2D at the start of every replay (the
        # sampler's agent_update), overwritten with the updated parameters at its end; the EpisodeBatch
        # fields and the logging scalars land in pinned buffers too (DeviceEpisodeBatch.prefetch_host)
        self._host_params = host_params
        self._stats_host = None
        self.h2d_bytes = self.d2h_bytes = 0
        if host_params is not None:
            for k, v in host_params.items():
                if v.is_cuda or not v.is_pinned():
                    raise ValueError('host_params[{!r}] must be a pinned CPU tensor'.format(k))

    def _enqueue(self):
        a = self._algo
        hp = self._host_params
        if hp is not None:  # H2D: this iteration's policy parameters (what `agent_update` carries)
            sd = a.policy.state_dict()
            for k, v in hp.items():
                sd[k].copy_(v, non_blocking=True)
            self.h2d_bytes = sum(v.numel() * v.element_size() for v in hp.values())
        eps = a._sampler.obtain_samples(0, self._batch_size, a.policy)
        d2h = eps.prefetch_host() if hp is not None else 0  # D2H: the EpisodeBatch (side stream, pinned)
        stats, B, n_eps = a._enqueue_iteration(eps)
        if hp is not None:
            if self._stats_host is None:
                self._stats_host = torch.empty(stats.shape, dtype=stats.dtype, pin_memory=True)
            self._stats_host.copy_(stats, non_blocking=True)        # D2H: logging scalars
            sd = a.policy.state_dict()
            for k, v in hp.items():                                  # D2H: updated parameters
                v.copy_(sd[k], non_blocking=True)
            cur = torch.cuda.current_stream(stats.device)
            cur.wait_event(eps._pinned_event)                        # join the side stream (graph capture needs it)
            self.d2h_bytes = d2h + stats.numel() * stats.element_size() + self.h2d_bytes
        return eps, stats, B, n_eps

    def enqueue(self):
        """Launch one iteration (graph replay after the warm-up) without synchronising."""
        if self._graph is None and self._eager_done >= self._warmup:
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self.eps, self._stats, B, n_eps = self._enqueue()
            self._graph, self._meta = g, (B, n_eps)
            if self._host_params is not None:
                self.eps._pinned_event = None  # a captured event cannot be waited on: run() synchronises the stream
        if self._graph is not None:
            self._graph.replay()
            return self._stats
        self.eps, stats, B, n_eps = self._enqueue()
        self._stats, self._meta = stats, (B, n_eps)
        self._eager_done += 1
        return stats

    def run(self, itr):
        stats = self.enqueue()
        B, n_eps = self._meta
        self._algo._sampler.total_env_steps += 0  # already counted by obtain_samples when eager
        if self._host_params is not None:
            torch.cuda.current_stream(stats.device).synchronize()  # the one host wait of the iteration
            return self._algo._log_iteration(itr, self._stats_host.numpy(), B, n_eps)
        return self._algo._log_iteration(itr, stats.cpu().numpy(), B, n_eps)
