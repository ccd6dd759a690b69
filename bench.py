"""Benchmark of the on-policy hot path (BASELINE.json metric: PPO env-steps/sec over a
full rollout+GAE+update iteration; GAE GB/s vs HBM peak).

    python bench.py --gpus N --steps K --warmup W            # this repo (B200 kernels)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path

One "step" = one PPO iteration: obtain_samples (rollout of n_envs x horizon steps,
EpisodeBatch assembly) + _train_once (baselines, GAE/returns, normalisation, E epochs of
minibatch updates for the policy and for the value function, logging passes).
Workload at N=1 = BASELINE.json configs[1]: Pendulum, 64x64 tanh policy + value,
4,096 envs x 200 steps on one B200; for N>1 every rank owns 4,096 envs (weak scaling, so the
N=1 line stays the BENCH line) and gradients are all-reduced per minibatch inside the loop
kernel over NVLink peer memory.  For N>1 the same run also times BASELINE configs[3] (C4:
65,536 Pendulum envs sharded over the N GPUs) -> key `c4`, and checks the in-kernel all-reduce
against the step-wise NCCL path -> key `dist_parity`.

Prints ONE JSON line on rank 0 (see the keys in `run_b200` / `run_reference`).
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
    ap.add_argument('--no-extras', action='store_true', help='skip the c3 / C5-sweep / c4 / dist_parity records')
    ap.add_argument('--cpu-envs', type=int, default=128, help='envs of the bounded CPU sample')
    ap.add_argument('--cpu-threads', type=int, default=0, help='host threads of the CPU arm (0: min(16, cores))')
    ap.add_argument('--ref-kind', default='auto', choices=['auto', 'reference', 'port'],
                    help="CPU arm: the unmodified reference (oracle/_ref or /root/reference) or the oracle port")
    ap.add_argument('--gae-T', type=int, default=200)
    ap.add_argument('--gae-N', type=int, default=1048576)
    ap.add_argument('--c4-envs', type=int, default=65536, help='total envs of the C4 record (N > 1)')
    ap.add_argument('--c4-minibatches', default='75776,409600',
                    help='per-GPU minibatch sizes of the C4 record, comma separated')
    ap.add_argument('--c4-steps', type=int, default=5)
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
def cpu_threads(a):
    """Host threads of the CPU arm: the same number whatever the launcher (torchrun exports
    OMP_NUM_THREADS=1, which would otherwise give the N > 1 reference lines a different baseline)."""
    import torch
    n = a.cpu_threads or min(16, len(os.sched_getaffinity(0)))
    torch.set_num_threads(n)
    return torch.get_num_threads()


def cpu_arm_kind(a):
    """'reference' when the unmodified garage package can be imported (from /root/reference in the
    build container, from its staged copy oracle/_ref on the GPU box), else 'port'."""
    if a.ref_kind == 'port':
        return 'port'
    from oracle import ref_shims
    if ref_shims.available():
        return 'reference'
    if a.ref_kind == 'reference':
        raise SystemExit('--ref-kind reference: neither /root/reference nor oracle/_ref is present '
                         '(run `python -m oracle.make_ref` where the reference is)')
    return 'port'


class CpuArm:
    """The reference's CPU implementation of the iteration on a BOUNDED sample of the workload:
    cpu_envs x horizon steps per iteration (the GPU arm runs n_envs x horizon), same networks, same
    PPO settings; the minibatch is min(minibatch, sample batch).  kind 'reference' = the unmodified
    garage package (LocalSampler + VecWorker + garage.torch PPO, oracle/ref_arm.py); kind 'port' =
    oracle/port.py, pinned against it by tests/test_port_vs_reference.py."""

    def __init__(self, a):
        self.a = a
        self.threads = cpu_threads(a)
        self.kind = cpu_arm_kind(a)
        self.mb = min(a.minibatch, a.cpu_envs * a.horizon)
        kw = dict(env_name='pendulum', n_envs=a.cpu_envs, horizon=a.horizon, hidden=a.hidden,
                  epochs=a.epochs, minibatch_size=self.mb, gae_lambda=0.95)
        if self.kind == 'reference':
            from oracle import ref_arm
            self._it = ref_arm.ReferenceIteration(**kw)
            self.run = self._it.run
        else:
            from oracle import port
            self.run = lambda: (lambda r: (r['env_steps'], r['rollout_s'], r['train_s']))(
                port.time_iteration(iters=1, **{**kw, 'minibatch_size': self.mb}))

    def config(self):
        a = self.a
        return dict(workload='PPO Pendulum (gym 0.17.2 dynamics), 64x64 tanh policy+value, {} envs x {} steps '
                             '(bounded CPU sample of the {} envs x {} steps GPU workload)'.format(
                                 a.cpu_envs, a.horizon, a.n_envs, a.horizon),
                    env='Pendulum', n_envs_per_gpu=a.cpu_envs, horizon=a.horizon,
                    batch_per_gpu=a.cpu_envs * a.horizon, hidden=[a.hidden, a.hidden], epochs=a.epochs,
                    minibatch_per_gpu=self.mb, discount=0.99, gae_lambda=0.95, clip=0.2, lr=2.5e-4,
                    center_adv=True, parallelism='cpu x{} threads'.format(self.threads),
                    sample_of=dict(n_envs_per_gpu=a.n_envs, horizon=a.horizon, minibatch_per_gpu=a.minibatch),
                    same_config=False)

    def sample(self, iters, roll, train):
        what = ('unmodified garage v2021.03.0 (LocalSampler + VecWorker + garage.torch PPO through oracle/ref_shims; '
                'Pendulum dynamics from the oracle restatement of gym 0.17.2)') if self.kind == 'reference' else \
               'oracle/port.py (reference algorithm restated: VecWorker loop + torch-CPU PPO)'
        return ('{}, {} iteration(s) of {} envs x {} steps, E={}, minibatch {}; rollout {:.2f}s + update {:.2f}s; '
                'rollout single-threaded like LocalSampler, torch intra-op threads {}').format(
                    what, iters, self.a.cpu_envs, self.a.horizon, self.a.epochs, self.mb, roll, train, self.threads)

    def measure(self, iters):
        """(env-steps/s, sample text) over `iters` iterations."""
        steps, roll, train = 0, 0.0, 0.0
        for _ in range(iters):
            s_, r_, t_ = self.run()
            steps += s_
            roll += r_
            train += t_
        return steps / (roll + train), self.sample(iters, roll, train), steps, roll + train


def run_reference(a):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    arm = CpuArm(a)
    for _ in range(a.warmup):
        arm.run()
    t0 = time.perf_counter()
    value, sample, steps, busy = arm.measure(a.steps)
    dt = time.perf_counter() - t0
    value = steps / dt
    line = dict(impl='reference', metric='ppo_env_steps_per_sec', value=value, unit='env-steps/s',
                n_gpus=a.gpus, steps=a.steps, warmup=a.warmup, ms_per_step=1e3 * dt / max(a.steps, 1),
                higher_is_better=True, scaling='weak', vs_baseline=None, dtype='f32',
                data='synthetic', config=arm.config(),
                cpu_baseline=dict(value=value, unit='env-steps/s', cores=arm.threads, kind=arm.kind, sample=sample),
                e2e=dict(value=value, unit='env-steps/s', h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                note='throughput of a CPU iteration does not depend on the number of env copies (VecWorker steps '
                     'them one by one in Python), so the bounded sample measures the rate of the full workload; '
                     'the ratio to the GPU arm is an extrapolation from that sample (config.same_config = false)')
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
    """K iterations bracketed by barrier + synchronize; every iteration additionally gets its own
    event pair (no extra synchronisation), so a run reports the median / min / max iteration next to
    the bracketed total and a clock ramp cannot hide in one number.
    Returns (total ms, [ms of each iteration])."""
    import torch
    barrier()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ev[0].record()
    for i in range(steps):
        flush_l2()
        fn(i)
        ev[i + 1].record()
    torch.cuda.synchronize()
    barrier()
    return ev[0].elapsed_time(ev[steps]), [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]


def spread(per):
    q = sorted(per)
    return dict(median=q[len(q) // 2], min=q[0], max=q[-1], n=len(q))


def warm(fn, n, at_least=30):
    """max(n, at_least) untimed iterations (a count, not a duration: every rank must run the same
    number because each iteration contains collectives).  The SM clock needs a few hundred ms of
    load to reach its steady state after an idle period; 30 iterations are ~0.3 s."""
    import torch
    k = max(n, at_least)
    for i in range(k):
        flush_l2()
        fn(i)
    torch.cuda.synchronize()
    return k


def ncu_traffic(kernel, **shape):
    """dram bytes per launch of `kernel` from the committed `ncu --set full` capture
    (profiles/r2_ncu_traffic.json) when it was taken on this exact shape, else None."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'profiles', 'r2_ncu_traffic.json')) as fh:
            rec = json.load(fh)[kernel]
    except (OSError, KeyError, ValueError):
        return None
    if any(rec.get(k) != v for k, v in shape.items()):
        return None
    return rec['dram_read_bytes'] + rec['dram_write_bytes']


def smem_pipe(kernel, **shape):
    """Shared-memory data-pipe utilisation of `kernel` from the committed ncu capture (same shape only):
    MMA operand fetch + LDS/STS wavefronts as a fraction of the pipe's peak over the whole launch."""
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'profiles', 'r2_ncu_traffic.json')) as fh:
            rec = json.load(fh)[kernel]
    except (OSError, KeyError, ValueError):
        return None
    if any(rec.get(k) != v for k, v in shape.items()) or 'smem_tc_pct_of_peak' not in rec:
        return None
    return dict(source='ncu --set full, profiles/r2_ncu_traffic.json (not measured live)',
                mma_operand_fetch_frac=rec['smem_tc_pct_of_peak'] / 100.0,
                lds_sts_frac=rec['smem_lsu_pct_of_peak'] / 100.0,
                frac=(rec['smem_tc_pct_of_peak'] + rec['smem_lsu_pct_of_peak']) / 100.0,
                note='fraction of the l1tex shared-memory data pipe over the whole launch; the tile phases '
                     '(58 % of a CTA\'s time) carry almost all of it')


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
                   traffic_unit='dram bytes per launch (ncu --set full, profiles/r2_ncu_traffic.json); the '
                                'gathered batch (rows x 40 B per epoch) is L2-resident',
                   kernel='train_loop_kernel<O={},A={}> (tcgen05 kind::tf32, 3xTF32 split: the tensor pipe '
                          'executes 3x the algorithmic FLOPs)'.format(O_, A),
                   us_per_launch=ms * 1e3, minibatch_steps_per_launch=steps, us_per_minibatch_step=ms * 1e3 / steps,
                   rows_per_launch=2 * B * a.epochs, peak_kind=kind + ' bf16 sustained',
                   note='algorithmic FLOPs = 3 x 2 x (O*H + H*H + H*A) per row and pass, both networks; '
                        'K = 64 contractions issued as M64 x N<=72 x K8 MMAs: 46 cycles per instruction from one '
                        'issuing thread, 32 cycles of tensor pipe each when both CTAs of an SM issue '
                        '(tools/ld_probe.cu, profiles/r2_ld_probe.log) = their 4 KB of shared-memory operands at 128 B per '
                        'cycle. The unit that is full is the shared-memory data pipe (see shared_memory_pipe), not the FLOP rate',
                   shared_memory_pipe=smem_pipe('train_loop_kernel', rows=B, minibatch=mb, epochs=a.epochs))
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
               traffic_unit='dram bytes per launch (ncu --set full, profiles/r2_ncu_traffic.json)',
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


# ------------------------------------------------------------ extra records --
def _pendulum_algo(n_envs, horizon, minibatch_global, epochs, rank, mode='auto', seed=1234):
    import torch
    from garage_b200.envs import PendulumEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.algos import PPO
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    torch.manual_seed(1)
    env = PendulumEnv(max_episode_length=horizon)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
    vf = GaussianMLPValueFunction(env.spec, hidden_sizes=(64, 64))
    sampler = GPUSampler(policy, env, max_episode_length=horizon, n_envs=n_envs, seed=seed,
                         env_id_offset=rank * n_envs)
    mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), m, max_optimization_epochs=epochs,
                                    minibatch_size=minibatch_global)
    algo = PPO(env.spec, policy, vf, sampler, policy_optimizer=mk(policy), vf_optimizer=mk(vf), discount=0.99,
               gae_lambda=0.95, lr_clip_range=0.2, center_adv=True, shuffle='device', update_mode=mode)
    return algo, policy, vf, sampler


def dist_parity(rank, world, device):
    """Multi-GPU parity on the driver's clock: the fused loop (gradient all-reduce INSIDE the kernel
    over NVLink peer memory) against the step-wise path (one NCCL all-reduce per minibatch step) on the
    same shards, 2 iterations x 3 epochs x minibatches of 1,024; every rank must hold identical bits."""
    import torch
    import torch.distributed as dist

    def run(mode, use_peer=None):
        algo, policy, vf, sampler = _pendulum_algo(64, 40, 1024, 3, rank, mode)
        algo._use_peer = use_peer  # None: default (peer memory); False: NCCL for the small exchanges
        for it in range(2):
            eps = sampler.obtain_samples(it, 64 * 40, policy)
            algo._train_once(it, eps)
        torch.cuda.synchronize()
        return policy.flat.clone(), vf.flat.clone(), algo._comm is not None, dict(algo.last_stats)

    rel = lambda x, y: float((x.double() - y.double()).norm() / y.double().norm())
    pf, vf_, used_peer, logs_peer = run('fused')
    ps, vs, _, logs_nccl = run('stepwise', use_peer=False)
    # advantage moments + logging scalars: peer-memory exchange kernel vs NCCL + host arithmetic
    log_err = max(abs(float(logs_peer[k]) - float(logs_nccl[k])) / max(1.0, abs(float(logs_nccl[k])))
                  for k in logs_nccl if k in logs_peer and logs_nccl[k] == logs_nccl[k])
    errs = torch.tensor([rel(pf, ps), rel(vf_, vs)], device=device)
    ref = pf.clone()
    dist.broadcast(ref, 0)
    same = torch.tensor([float(torch.equal(ref, pf))], device=device)
    dist.all_reduce(errs, op=dist.ReduceOp.MAX)
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    return dict(fused_vs_nccl_stepwise_rel_err=float(errs.max()), policy_rel_err=float(errs[0]),
                vf_rel_err=float(errs[1]), ranks_bit_identical=bool(same.item()),
                fused_used_peer_memory=bool(used_peer), world=world,
                logs_peer_vs_nccl_max_rel_err=log_err, logged_keys_compared=len(logs_nccl),
                workload='PPO Pendulum 64 envs x 40 steps per rank, 2 iterations, E=3, global minibatch 1,024')


def c4_records(a, rank, world, device, barrier):
    """BASELINE configs[3]: PPO Pendulum, 65,536 envs sharded over the N GPUs (65,536 / N per GPU),
    E = 10, one record per per-GPU minibatch size (the minibatch is a free hyper-parameter of the
    config; the reference default 64 would be 204,800 optimiser steps per epoch)."""
    import torch
    n_envs = a.c4_envs // world
    B = n_envs * a.horizon
    out = []
    for mb in [int(x) for x in a.c4_minibatches.split(',') if x]:
        mb = min(mb, B)
        algo, policy, vf, sampler = _pendulum_algo(n_envs, a.horizon, mb * world, a.epochs, rank)
        m = measure(a, algo, policy, sampler, B, world, device, barrier, a.c4_steps, 3)
        sp = spread(m['per'])
        out.append(dict(n_envs_total=n_envs * world, n_envs_per_gpu=n_envs, horizon=a.horizon,
                        minibatch_per_gpu=mb, epochs=a.epochs, optimiser_steps_per_iteration=a.epochs * -(-B // mb),
                        steps=a.c4_steps, value=B * world * a.c4_steps / (m['ms_total'] * 1e-3), unit='env-steps/s',
                        ms_per_step=m['ms_total'] / a.c4_steps, ms_per_step_spread=sp,
                        e2e=dict(value=B * world * a.c4_steps / (m['e2e_ms_total'] * 1e-3), unit='env-steps/s',
                                 ms_per_step=m['e2e_ms_total'] / a.c4_steps,
                                 h2d_bytes_per_step=m['io']['h2d'], d2h_bytes_per_step=m['io']['d2h']),
                        target='>= 1e9 env-steps/s on 8 GPUs (north star)'))
        del algo, policy, vf, sampler
        torch.cuda.empty_cache()
    return out


def gae_cpu_baseline(kind):
    """The reference's own GAE + returns on the host cores at a size it finishes in seconds
    (garage.torch.compute_advantages, torch/_functions.py:25-85: conv2d with a length-T filter;
    garage.np.discount_cumsum, np/_functions.py:111-128: scipy lfilter per episode), same 17 B/element."""
    try:
        if kind == 'reference':
            from oracle import ref_arm
            r = ref_arm.time_gae_reference(T=200, N=4096)
            r['kind'] = 'reference'
        else:
            import numpy as np
            import torch
            from oracle import port
            T, N = 200, 4096
            rng = np.random.default_rng(0)
            rew = torch.from_numpy(rng.standard_normal((N, T)).astype(np.float32))
            val = torch.from_numpy(rng.standard_normal((N, T)).astype(np.float32))
            t0 = time.perf_counter()
            port.compute_advantages_torch(0.99, 0.95, T, val, rew)
            t1 = time.perf_counter()
            np.stack([port.discount_cumsum(r_, 0.99) for r_ in rew.numpy().astype(np.float64)])
            t2 = time.perf_counter()
            r = dict(T=T, N=N, compute_advantages_s=t1 - t0, discount_cumsum_s=t2 - t1,
                     gbs=17.0 * T * N / (t2 - t0) / 1e9, threads=torch.get_num_threads(), kind='port')
        r['unit'] = 'GB/s'
        return r
    except Exception as e:  # a reported baseline must not take the bench line down
        return dict(error=repr(e))


def c3_record(device, n_envs=16384, horizon=200, iters=5):
    """BASELINE configs[2]: garage.torch VPG (REINFORCE) on CartPole dynamics, 16,384 envs, ragged
    episodes (early termination -> GAE with done masks, episode packing); the reference's default
    optimiser settings for VPG (one full-batch step per network per iteration).  Eager iterations
    (batch shapes change every iteration), CUDA events, median of `iters`."""
    import torch
    from garage_b200.envs import CartPoleEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.algos import VPG
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    try:
        torch.manual_seed(0)
        env = CartPoleEnv(max_episode_length=horizon)
        policy = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
        vf = GaussianMLPValueFunction(env.spec, hidden_sizes=(64, 64))
        sampler = GPUSampler(policy, env, max_episode_length=horizon, n_envs=n_envs, seed=1)
        algo = VPG(env.spec, policy, vf, sampler, discount=0.99, gae_lambda=0.95)
        B = n_envs * horizon
        per, steps, eps_n = [], [], 0
        for it in range(iters + 2):
            flush_l2()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            eps = sampler.obtain_samples(it, B, policy)
            algo._train_once(it, eps)
            e1.record()
            torch.cuda.synchronize()
            if it >= 2:
                per.append(e0.elapsed_time(e1))
                steps.append(int(eps.n_steps))
                eps_n = int(eps.n_episodes)
        sp = spread(per)
        k = per.index(sp['median'])
        return dict(workload='VPG CartPole (gym 0.17.2 dynamics), 64x64, {} envs, batch >= {} steps, ragged episodes'.format(n_envs, B),
                    n_envs=n_envs, value=steps[k] / (sp['median'] * 1e-3), unit='env-steps/s', ms_per_step_spread=sp,
                    env_steps_per_iteration=steps[k], episodes_last_iteration=eps_n, fused_loop=bool(algo._fused_applies()),
                    average_return=float(algo.last_stats['Evaluation/AverageReturn']))
    except Exception as e:
        return dict(error=repr(e))


_clean_buf = None


def flush_l2_clean():
    """flush_l2() (write 256 MB) followed by a read of another 256 MB: the L2 then holds CLEAN lines of
    a foreign buffer.  After the write alone it holds 126 MB of dirty lines, and a short kernel measured
    right behind it pays for writing them back (every miss evicts one): +2..8 us on the 25..50 us GAE
    launches (tools/gae_tune.py), traffic that is not the kernel's."""
    import torch
    global _clean_buf
    flush_l2()
    if _clean_buf is None:
        _clean_buf = torch.zeros(64 << 20, dtype=torch.float32, device='cuda')
    _clean_buf.sum()


def c5_sweep(device):
    """BASELINE configs[4]: standalone GAE / discounted-return sweep, gamma 0.99, lambda 0.95, 1 % of
    the steps terminal, L2 flushed (written, then read: clean lines) before every launch, median of 7;
    17 algorithmic B per element."""
    import torch
    from garage_b200 import ops
    hbm, _, kind = peaks()
    rows = []
    try:
        g = torch.Generator(device=device).manual_seed(0)
        for T in (128, 200, 512, 1000):
            for N in (1024, 16384, 65536, 262144, 1048576):
                if 17 * T * N > 12e9:
                    continue
                rew = torch.randn(T, N, device=device, generator=g)
                val = torch.randn(T, N, device=device, generator=g)
                st = torch.ones(T, N, device=device, dtype=torch.uint8)
                st[torch.rand(T, N, device=device, generator=g) < 0.01] = 2
                st[T - 1] = 3
                adv, ret = torch.empty_like(rew), torch.empty_like(rew)
                ts = []
                for i in range(9):
                    flush_l2_clean()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    ops.gae_time_major(rew, val, st, None, discount=0.99, gae_lambda=0.95, adv_out=adv, ret_out=ret)
                    e1.record()
                    torch.cuda.synchronize()
                    if i >= 2:
                        ts.append(e0.elapsed_time(e1))
                ms = sorted(ts)[len(ts) // 2]
                gbs = 17.0 * T * N / (ms * 1e-3) / 1e9
                rows.append(dict(T=T, N=N, us=ms * 1e3, gbs=gbs, frac=gbs / hbm))
                del rew, val, st, adv, ret
        return dict(peak_gbs=hbm, peak_kind=kind, p_done=0.01,
                    l2='flushed before every launch: 256 MB written, then 256 MB read (clean lines)', rows=rows)
    except Exception as e:
        return dict(error=repr(e), rows=rows)


def make_step(a, algo, policy, sampler, B):
    """One iteration as a callable: a replayed CUDA graph of the whole iteration unless --no-graph."""
    from garage_b200.torch.algos.graphed import GraphedIteration
    if not a.no_graph:
        runner = GraphedIteration(algo, B, warmup=2)
        return lambda i: runner.enqueue()

    def step(i):
        eps = sampler.obtain_samples(i, B, policy)
        algo._enqueue_iteration(eps)
    return step


def make_e2e_step(a, algo, policy, sampler, B, io):
    """The call a user makes, with HOST buffers: parameters in (pinned, H2D), EpisodeBatch + logging
    scalars + updated parameters out (D2H).  With graphs enabled the public
    `GraphedIteration(algo, B, host_params=...)` replays the same work -- copies included -- as one
    CUDA graph and waits once per iteration; --no-graph runs it as eager API calls."""
    host_sd = {k: v.detach().cpu().pin_memory() for k, v in policy.state_dict().items()}
    if not a.no_graph:
        from garage_b200.torch.algos.graphed import GraphedIteration
        runner = GraphedIteration(algo, B, warmup=2, host_params=host_sd)

        def graphed_step(i):
            ret = runner.run(i)
            eps = runner.eps
            _ = (eps.observations, eps.actions, eps.rewards, eps.lengths)  # numpy views of the pinned batch
            io['h2d'], io['d2h'] = runner.h2d_bytes, runner.d2h_bytes
            return ret
        return graphed_step

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
    return e2e_step


def max_over_ranks(x, world, device):
    import torch
    import torch.distributed as dist
    t = torch.tensor(x if isinstance(x, (list, tuple)) else [x], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def measure(a, algo, policy, sampler, B, world, device, barrier, steps, warmup, with_e2e=True):
    """Device-resident and end-to-end timing of one workload.  Returns a dict."""
    step = make_step(a, algo, policy, sampler, B)
    wk = warm(step, max(warmup, 3))
    ms, per = timed(step, steps, barrier)
    ms = max_over_ranks(ms, world, device)[0]
    per = max_over_ranks(per, world, device)
    out = dict(ms_total=ms, per=per, warm_iters=wk)
    if with_e2e:
        io = dict(h2d=0, d2h=0)
        e2e_step = make_e2e_step(a, algo, policy, sampler, B, io)
        warm(e2e_step, max(warmup, 3), at_least=10)
        ems, eper = timed(e2e_step, steps, barrier)
        out.update(e2e_ms_total=max_over_ranks(ems, world, device)[0],
                   e2e_per=max_over_ranks(eper, world, device), io=io)
    return out


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

    algo, policy, vf, sampler = build_algo(a, rank, world, device)
    B = a.n_envs * a.horizon
    use_graph = not a.no_graph  # multi-GPU too: NCCL collectives and the fused loop are capturable

    def eager(i):
        eps = sampler.obtain_samples(i, B, policy)
        algo._enqueue_iteration(eps)
    launches = count_launches(lambda: eager(0))  # eager iteration: counts kernels per step
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    m = measure(a, algo, policy, sampler, B, world, device, barrier, a.steps, a.warmup)
    clk = clocks.stop() if rank == 0 else None
    ms, ems, io = m['ms_total'], m['e2e_ms_total'], m['io']
    value = B * world * a.steps / (ms * 1e-3)
    e2e_value = B * world * a.steps / (ems * 1e-3)

    extras = {}
    if world > 1 and not a.no_extras:
        extras['dist_parity'] = dist_parity(rank, world, device)
        extras['c4'] = c4_records(a, rank, world, device, barrier)

    if rank != 0:
        if world > 1:
            _leave(dist)
        return
    mlp_roof, gae_roof = kernel_rooflines(a, algo, device)
    cpu = None
    if world == 1 and not a.no_cpu_baseline:
        arm = CpuArm(a)
        arm.run()
        v, sample, _, _ = arm.measure(2)
        cpu = dict(value=v, unit='env-steps/s', cores=arm.threads, kind=arm.kind, sample=sample)
        gae_roof['cpu_baseline'] = gae_cpu_baseline(arm.kind)
    if world == 1 and not a.no_extras:
        extras['c3'] = c3_record(device)
        extras['c5_gae_sweep'] = c5_sweep(device)
    cfg = workload_config(a, world)
    cfg['cuda_graph'] = use_graph
    sp, esp = spread(m['per']), spread(m['e2e_per'])
    line = dict(metric='ppo_env_steps_per_sec', value=value, unit='env-steps/s', n_gpus=world,
                steps=a.steps, warmup=max(a.warmup, 3), ms_per_step=ms / a.steps, higher_is_better=True,
                scaling='weak', vs_baseline=None, dtype='f32', data='synthetic', config=cfg,
                ms_per_step_spread=sp, warm_iterations_run=m['warm_iters'],
                value_at_median=B * world / (sp['median'] * 1e-3),
                e2e=dict(value=e2e_value, unit='env-steps/s', ms_per_step=ems / a.steps, ms_per_step_spread=esp,
                         h2d_bytes_per_step=io['h2d'], d2h_bytes_per_step=io['d2h']),
                gpu_launches=launches * a.steps, gpu_launches_per_step=launches, clocks=clk,
                roofline=mlp_roof, roofline_gae=gae_roof, cpu_baseline=cpu,
                last_stats={k: float(v) for k, v in algo.last_stats.items()})
    line.update(extras)
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
