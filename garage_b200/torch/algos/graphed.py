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

    def __init__(self, algo, batch_size, warmup=1):
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

    def _enqueue(self):
        a = self._algo
        eps = a._sampler.obtain_samples(0, self._batch_size, a.policy)
        stats, B, n_eps = a._enqueue_iteration(eps)
        return eps, stats, B, n_eps

    def enqueue(self):
        """Launch one iteration (graph replay after the warm-up) without synchronising."""
        if self._graph is None and self._eager_done >= self._warmup:
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self.eps, self._stats, B, n_eps = self._enqueue()
            self._graph, self._meta = g, (B, n_eps)
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
        return self._algo._log_iteration(itr, stats.cpu().numpy(), B, n_eps)
