import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from garage_b200 import ops
dev = torch.device('cuda:0')
T, N = 1000, 16384
g = torch.Generator(device=dev).manual_seed(0)
rew = torch.randn(T, N, device=dev, generator=g); val = torch.randn(T, N, device=dev, generator=g)
st = torch.ones(T, N, device=dev, dtype=torch.uint8); st[T-1] = 3
adv = torch.empty_like(rew); ret = torch.empty_like(rew)
def t(fn, n=20):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / n
print('adv+ret  %.1f us' % t(lambda: ops.gae_time_major(rew, val, st, None, discount=0.99, gae_lambda=0.95, adv_out=adv, ret_out=ret)))
print('adv only %.1f us' % t(lambda: ops.gae_time_major(rew, val, st, None, discount=0.99, gae_lambda=0.95, adv_out=adv, want_ret=False)))
print('ret only %.1f us' % t(lambda: ops.gae_time_major(rew, None, st, None, discount=0.99, gae_lambda=0.95, ret_out=ret, want_adv=False)))
print('no st    %.1f us' % t(lambda: ops.gae_time_major(rew, val, None, None, discount=0.99, gae_lambda=0.95, adv_out=adv, ret_out=ret)))
