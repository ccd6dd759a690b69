"""Benchmark of the on-policy hot path (BASELINE.json metric: PPO env-steps/sec over a
full rollout+GAE+update iteration; GAE GB/s vs HBM peak).

    python bench.py --gpus N --steps K --warmup W            # this repo (B200 kernels)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path

One "step" = one PPO iteration: obtain_samples (rollout of n_envs x horizon steps,
EpisodeBatch assembly) + _train_once (baselines, GAE/returns, normalisation, E epochs of
minibatch updates for the policy and for the value function, logging passes).
Workload at N=1 = BASELINE.json configs[1]: Pendulum, 64x64 tanh policy + value,
4,096 envs x 200 steps on one B200; for N>1 every rank owns 4,096 envs (weak scaling)
and gradients are all-reduced per minibatch over NCCL.

Prints ONE JSON line on rank 0 (see the keys in `main`).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--n-envs', type=int, default=4096, help='envs per GPU')
    ap.add_argument('--horizon', type=int, default=200)
    ap.add_argument('--hidden', type=int, default=64)
    ap.add_argument('--epochs', type=int, default=10, help='PPO max_optimization_epochs (reference default 10)')
    ap.add_argument('--minibatch', type=int, default=37888,
                    help='minibatch size per GPU; default 148 SMs x 4 tiles x 64 rows: every CTA of the loop '
                         'kernel gets the same number of row tiles (32,768 leaves 80 of 148 CTAs one tile short)')
    ap.add_argument('--no-graph', action='store_true', help='eager launches instead of one CUDA graph')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--cpu-envs', type=int, default=32, help='envs of the bounded CPU sample')
    ap.add_argument('--gae-T', type=int, default=200)
    ap.add_argument('--gae-N', type=int, default=1048576)
    return ap.parse_args()


# ------------------------------------------------------------------ clocks --
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                 '--format=csv,noheader,nounits', '-lms', '25'],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=['nvidia-smi unavailable'])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except (ValueError, IndexError):
                continue
            for name, v in zip(names, r[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(name)
        sm.sort()
        return dict(sm_mhz=(sm[len(sm) // 2] if sm else None), sm_max_mhz=mx, reasons=sorted(reasons),
                    samples=len(sm))


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return d['hbm_gbs'], d.get('bf16_tflops_sustained', d['bf16_tflops']), 'measured'
    return 6650.0, 1590.0, 'fallback'


def workload_config(a, world):
    return dict(workload='PPO Pendulum (gym 0.17.2 dynamics), 64x64 tanh policy+value, '
                         '{} envs x {} steps per GPU'.format(a.n_envs, a.horizon),
                env='Pendulum', n_envs_per_gpu=a.n_envs, horizon=a.horizon,
                batch_per_gpu=a.n_envs * a.horizon, hidden=[a.hidden, a.hidden],
                epochs=a.epochs, minibatch_per_gpu=a.minibatch, discount=0.99, gae_lambda=0.95,
                clip=0.2, lr=2.5e-4, center_adv=True, parallelism='dp{}'.format(world),
                cache='L2 flushed between iterations (a 256 MB fill, inside the timed region); every '
                      'iteration also regenerates all of its buffers (rollout); kernel rooflines: GAE case '
                      'uses inputs > L2, the loop kernel re-reads its L2-resident batch by design')


# ---------------------------------------------------------- reference arm --
def cpu_iteration_sample(a, iters):
    """The oracle port (reference algorithm on the host cores) on a BOUNDED sample of the
    same workload: cpu_envs x horizon steps per iteration, same nets and PPO settings."""
    import torch
    from oracle import port
    mb = min(a.minibatch, a.cpu_envs * a.horizon)
    r = port.time_iteration(env_name='pendulum', n_envs=a.cpu_envs, horizon=a.horizon,
                            hidden=a.hidden, epochs=a.epochs, minibatch_size=mb, iters=iters,
                            gae_lambda=0.95)
    sample = ('oracle/port.py (reference algorithm restated: VecWorker loop + torch-CPU PPO), '
              '{} iteration(s) of {} envs x {} steps, E={}, minibatch {}; rollout {:.2f}s + update {:.2f}s; '
              'rollout single-threaded like LocalSampler, torch intra-op threads {}').format(
                  iters, a.cpu_envs, a.horizon, a.epochs, mb, r['rollout_s'], r['train_s'],
                  torch.get_num_threads())
    return r, sample, torch.get_num_threads()


def run_reference(a):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    for _ in range(a.warmup):
        cpu_iteration_sample(a, 1)
    t0 = time.perf_counter()
    steps = 0
    sample = cores = None
    for _ in range(a.steps):
        r, sample, cores = cpu_iteration_sample(a, 1)
        steps += r['env_steps']
    dt = time.perf_counter() - t0
    value = steps / dt
    line = dict(impl='reference', metric='ppo_env_steps_per_sec', value=value, unit='env-steps/s',
                n_gpus=a.gpus, steps=a.steps, warmup=a.warmup, ms_per_step=1e3 * dt / max(a.steps, 1),
                higher_is_better=True, scaling='weak', vs_baseline=None, dtype='f32',
                data='synthetic', config=workload_config(a, a.gpus),
                cpu_baseline=dict(value=value, unit='env-steps/s', cores=cores, kind='port', sample=sample),
                e2e=dict(value=value, unit='env-steps/s', h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------- b200 arm --
def build_algo(a, rank, world, device):
    import torch
    from garage_b200.envs import PendulumEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.algos import PPO
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    torch.manual_seed(1)  # same initial weights on every rank
    env = PendulumEnv(max_episode_length=a.horizon)
    hs = (a.hidden, a.hidden)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=hs)
    vf = GaussianMLPValueFunction(env.spec, hidden_sizes=hs)
    sampler = GPUSampler(policy, env, max_episode_length=a.horizon, n_envs=a.n_envs, seed=1234,
                         env_id_offset=rank * a.n_envs)
    mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), m,
                                    max_optimization_epochs=a.epochs,
                                    minibatch_size=a.minibatch * world)
    algo = PPO(env.spec, policy, vf, sampler, policy_optimizer=mk(policy), vf_optimizer=mk(vf),
               discount=0.99, gae_lambda=0.95, lr_clip_range=0.2, center_adv=True, shuffle='device')
    return algo, policy, vf, sampler


LAUNCHES_PER_CALL = dict(b2g_moments=3, b2g_pack_plan=2)  # kernels (+memset) per C-ABI call


def count_launches(fn):
    """Run fn() once counting kernels launched through the C ABI."""
    from garage_b200 import _lib
    counter = [0]
    real = _lib.call

    def counting(name, *args):
        counter[0] += LAUNCHES_PER_CALL.get(name, 1)
        return real(name, *args)
    import garage_b200.ops as ops
    ops.call = counting
    try:
        fn()
    finally:
        ops.call = real
    return counter[0]


_flush_buf = None


def flush_l2():
    """Write a buffer larger than the 126 MB L2 (timing rule: no data survives in cache between timed
    iterations).  Enqueued on the current stream, so it sits inside the timed region."""
    import torch
    global _flush_buf
    if _flush_buf is None:
        _flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
    _flush_buf.fill_(1)


def timed(fn, steps, barrier):
    import torch
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        flush_l2()
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    barrier()
    return e0.elapsed_time(e1)


def ncu_traffic(kernel, **shape):
    """dram bytes per launch of `kernel` from the committed `ncu --set full` capture
    (profiles/r1_ncu_traffic.json) when it was taken on this exact shape, else None."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'profiles', 'r1_ncu_traffic.json')) as fh:
            rec = json.load(fh)[kernel]
    except (OSError, KeyError, ValueError):
        return None
    if any(rec.get(k) != v for k, v in shape.items()):
        return None
    return rec['dram_read_bytes'] + rec['dram_write_bytes']


def kernel_rooflines(a, algo, device):
    """Dominant kernel (the fused training loop: every minibatch step of both networks in one
    launch; falls back to the one-minibatch kernel when the loop does not apply) and the GAE
    kernel, each timed alone with CUDA events on the launching stream."""
    import torch
    from garage_b200 import ops
    hbm, tf, kind = peaks()
    pol, vf = algo.policy, algo._value_function
    B, mb = a.n_envs * a.horizon, min(a.minibatch, a.n_envs * a.horizon)
    g = torch.Generator(device=device).manual_seed(0)
    obs = torch.randn(B, 3, device=device, generator=g)
    act = torch.randn(B, 1, device=device, generator=g)
    adv = torch.randn(B, device=device, generator=g)
    olp = torch.randn(B, device=device, generator=g) * 0.1 - 1.0
    ret = torch.randn(B, device=device, generator=g)
    O_, H, A = pol.mlp_shape.O, pol.mlp_shape.H, pol.mlp_shape.A
    flops_row = 3 * 2 * (O_ * H + H * H + H * A)  # 3 x forward FLOPs per row and pass
    kw = algo._policy_kwargs()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if algo._fused_applies():
        def net(owner, policy):
            d = dict(shape=owner.mlp_shape, params=owner.flat.clone(), adam_m=torch.zeros_like(owner.flat),
                     adam_v=torch.zeros_like(owner.flat),
                     step_counter=torch.zeros(1, dtype=torch.int64, device=device), lr=2.5e-4, beta1=0.9,
                     beta2=0.999, eps=1e-8, obs=obs, minibatch=mb, epochs=a.epochs,
                     shuffle=ops.SHUFFLE_DEVICE, shuffle_seed=1)
            d.update(dict(act=act, adv=adv, old_logp=olp, **kw) if policy else dict(ret=ret, loss_kind=ops.LOSS_VF))
            return d
        nets = [net(pol, True), net(vf, False)]
        reps = 5
        for _ in range(3):
            ops.mlp_train_loop(nets)
        e0.record()
        for _ in range(reps):
            ops.mlp_train_loop(nets)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        steps = a.epochs * -(-B // mb)
        flops = 2 * flops_row * B * a.epochs
        ach = flops / (ms * 1e-3) / 1e12
        mlp = dict(bound='tensor', achieved=ach, peak=tf, unit='TFLOP/s', frac=ach / tf,
                   traffic=ncu_traffic('train_loop_kernel', rows=B, minibatch=mb, epochs=a.epochs),
                   traffic_unit='dram bytes per launch (ncu --set full, profiles/r1_ncu_traffic.json); the '
                                'gathered batch (rows x 40 B per epoch) is L2-resident',
                   kernel='train_loop_kernel<O={},A={}> (tcgen05 kind::tf32, 3xTF32 split: the tensor pipe '
                          'executes 3x the algorithmic FLOPs)'.format(O_, A),
                   us_per_launch=ms * 1e3, minibatch_steps_per_launch=steps, us_per_minibatch_step=ms * 1e3 / steps,
                   rows_per_launch=2 * B * a.epochs, peak_kind=kind + ' bf16 sustained',
                   note='algorithmic FLOPs = 3 x 2 x (O*H + H*H + H*A) per row and pass, both networks; '
                        'K = 64 contractions issued as M64/128 x N64..160 x K8 MMAs whose ~55-65 cycle '
                        'instruction floor (tools/ld_probe micro-benchmark), not the FLOP rate, bounds them')
    else:
        idx = ops.permutation(B, 0, device=device)
        reps = 50
        for _ in range(5):
            ops.mlp_minibatch_grad(pol.mlp_shape, pol.flat, obs, idx=idx[:mb], mb=mb, act=act, adv=adv, old_logp=olp, **kw)
        e0.record()
        for r in range(reps):
            s = (r * mb) % max(B - mb + 1, 1)
            ops.mlp_minibatch_grad(pol.mlp_shape, pol.flat, obs, idx=idx[s:s + mb], mb=mb, act=act, adv=adv, old_logp=olp, **kw)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        ach = flops_row * mb / (ms * 1e-3) / 1e12
        mlp = dict(bound='tensor', achieved=ach, peak=tf, unit='TFLOP/s', frac=ach / tf, traffic=None,
                   kernel='mlp minibatch kernel H={},O={},A={} (path {})'.format(H, O_, A, ops.mlp_path()),
                   us_per_launch=ms * 1e3, rows_per_launch=mb, peak_kind=kind + ' bf16 sustained',
                   note='algorithmic FLOPs = 3 x 2 x (O*H + H*H + H*A) per row')
    # GAE: inputs larger than L2
    T, N = a.gae_T, a.gae_N
    rew = torch.randn(T, N, device=device, generator=g)
    val = torch.randn(T, N, device=device, generator=g)
    st = torch.ones(T, N, dtype=torch.uint8, device=device)
    st[T - 1] = 3
    advo, reto = torch.empty_like(rew), torch.empty_like(rew)
    for _ in range(3):
        ops.gae_time_major(rew, val, st, discount=0.99, gae_lambda=0.95, adv_out=advo, ret_out=reto)
    e0.record()
    for _ in range(10):
        ops.gae_time_major(rew, val, st, discount=0.99, gae_lambda=0.95, adv_out=advo, ret_out=reto)
    e1.record()
    torch.cuda.synchronize()
    gms = e0.elapsed_time(e1) / 10
    gbs = 17.0 * T * N / (gms * 1e-3) / 1e9
    gae = dict(bound='hbm', achieved=gbs, peak=hbm, unit='GB/s', frac=gbs / hbm,
               traffic=ncu_traffic('gae_time_major_kernel', T=T, N=N), algorithmic_bytes=17.0 * T * N,
               traffic_unit='dram bytes per launch (ncu --set full, profiles/r1_ncu_traffic.json)',
               kernel='gae_time_major_kernel', T=T, N=N, us_per_launch=gms * 1e3,
               footprint_mb=17.0 * T * N / 1e6, peak_kind=kind,
               note='17 algorithmic B per (t,n): r 4 + V 4 + step_type 1 read, A 4 + R 4 written')
    return mlp, gae


def _leave(dist):
    """End of a multi-rank run: meet the other ranks, then exit without tearing the communicator
    down (a captured graph still references its NCCL kernels; destroying it first can block)."""
    dist.barrier()
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)


def run_b200(a):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py --impl b200 needs a CUDA device (there is no CPU fallback)')
    torch.cuda.set_device(local_rank)
    device = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=device)
    barrier = (lambda: dist.barrier()) if world > 1 else (lambda: None)

    from garage_b200.torch.algos.graphed import GraphedIteration
    algo, policy, vf, sampler = build_algo(a, rank, world, device)
    B = a.n_envs * a.horizon
    use_graph = not a.no_graph  # multi-GPU too: NCCL collectives and the fused loop are capturable

    # ---- device-resident throughput ("value"): inputs (weights, env state) live in HBM
    if use_graph:
        runner = GraphedIteration(algo, B, warmup=2)
        step = lambda i: runner.enqueue()
    else:
        def step(i):
            eps = sampler.obtain_samples(i, B, policy)
            algo._enqueue_iteration(eps)
    launches = count_launches(lambda: step(0))  # eager iteration: counts kernels per step
    for i in range(max(a.warmup, 3)):
        step(i)
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    ms = timed(step, a.steps, barrier)
    clk = clocks.stop() if rank == 0 else None
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = B * world * a.steps / (ms * 1e-3)

    # ---- end to end through the public API with HOST buffers
    host_sd = {k: v.detach().cpu().pin_memory() for k, v in policy.state_dict().items()}
    io = dict(h2d=0, d2h=0)

    def e2e_step(i):
        eps = sampler.obtain_samples(i, B, host_sd)            # H2D: policy parameters
        d2h = eps.prefetch_host()                              # D2H: the EpisodeBatch (async, pinned)
        ret = algo._train_once(i, eps)                         # D2H: logging scalars (sync)
        _ = (eps.observations, eps.actions, eps.rewards, eps.lengths)  # numpy views of the batch
        for k, v in policy.state_dict().items():               # D2H: updated parameters
            host_sd[k].copy_(v, non_blocking=True)
        io['h2d'] = sum(v.numel() * v.element_size() for v in host_sd.values())
        io['d2h'] = d2h + 40 * 8 + io['h2d']
        return ret
    for i in range(max(a.warmup, 3)):
        e2e_step(i)
    ems = timed(e2e_step, a.steps, barrier)
    t = torch.tensor([ems], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ems = float(t.item())
    e2e_value = B * world * a.steps / (ems * 1e-3)

    if rank != 0:
        if world > 1:
            _leave(dist)
        return
    mlp_roof, gae_roof = kernel_rooflines(a, algo, device)
    cpu = None
    if world == 1 and not a.no_cpu_baseline:
        r, sample, cores = cpu_iteration_sample(a, 3)
        cpu = dict(value=r['env_steps'] / r['total_s'], unit='env-steps/s', cores=cores, kind='port',
                   sample=sample)
    cfg = workload_config(a, world)
    cfg['cuda_graph'] = use_graph
    line = dict(metric='ppo_env_steps_per_sec', value=value, unit='env-steps/s', n_gpus=world,
                steps=a.steps, warmup=max(a.warmup, 3), ms_per_step=ms / a.steps, higher_is_better=True,
                scaling='weak', vs_baseline=None, dtype='f32', data='synthetic', config=cfg,
                e2e=dict(value=e2e_value, unit='env-steps/s', ms_per_step=ems / a.steps,
                         h2d_bytes_per_step=io['h2d'], d2h_bytes_per_step=io['d2h']),
                gpu_launches=launches * a.steps, gpu_launches_per_step=launches, clocks=clk,
                roofline=mlp_roof, roofline_gae=gae_roof, cpu_baseline=cpu,
                last_stats={k: float(v) for k, v in algo.last_stats.items()})
    print(json.dumps(line), flush=True)
    if world > 1:
        _leave(dist)


def main():
    a = parse()
    if a.impl == 'reference':
        run_reference(a)
    else:
        run_b200(a)


if __name__ == '__main__':
    main()
