"""Sweep of the GAE kernel configurations (B2G_GAE_FORCE) at the sizes the pipeline uses: one process
per configuration (the choice is read once per process), L2 flushed before every launch, median of 9."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys, torch
sys.path.insert(0, %r)
from garage_b200 import ops
dev = torch.device('cuda:0')
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
clean = torch.zeros(64 << 20, dtype=torch.float32, device=dev)
import os
CLEAN = os.environ.get('FLUSH_CLEAN', '0') == '1'
g = torch.Generator(device=dev).manual_seed(0)
out = []
for T, N in [(200, 4096), (200, 16384), (200, 32768), (200, 65536), (128, 16384), (512, 16384), (1000, 16384), (200, 262144), (1000, 65536)]:
    rew = torch.randn(T, N, device=dev, generator=g); val = torch.randn(T, N, device=dev, generator=g)
    st = torch.ones(T, N, device=dev, dtype=torch.uint8); st[torch.rand(T, N, device=dev, generator=g) < 0.01] = 2; st[T - 1] = 3
    adv = torch.empty_like(rew); ret = torch.empty_like(rew)
    ts = []
    for i in range(12):
        flush.fill_(1)
        if CLEAN: clean.sum()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ops.gae_time_major(rew, val, st, None, discount=0.99, gae_lambda=0.95, adv_out=adv, ret_out=ret); e1.record()
        torch.cuda.synchronize()
        if i >= 3: ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    out.append('%%dx%%d %%.1fus %%.2f' %% (T, N, ms * 1e3, 17.0 * T * N / (ms * 1e-3) / 1e9 / 6540.8))
print(' | '.join(out))
''' % ROOT
for cfg in ['', '1', '16,8', '16,16', '32,16']:
    env = dict(os.environ)
    if cfg:
        env['B2G_GAE_FORCE'] = cfg
    for clean in ('0', '1'):
        env['FLUSH_CLEAN'] = clean
        r = subprocess.run([sys.executable, '-c', CHILD], env=env, capture_output=True, text=True)
        print('%-8s clean=%s' % (cfg or 'default', clean), r.stdout.strip() or r.stderr[-300:], flush=True)
