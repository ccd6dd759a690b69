"""The bench's GAE roofline case (T x N time-major, adv + returns) run a few times: the target of
`ncu --set full -k regex:gae_time_major` for the dram traffic figure in bench.py's `roofline_gae`."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from garage_b200 import ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--T', type=int, default=200)
ap.add_argument('--N', type=int, default=1048576)
ap.add_argument('--reps', type=int, default=3)
a = ap.parse_args()
dev = torch.device('cuda:0')
g = torch.Generator(device=dev).manual_seed(0)
rew = torch.randn(a.T, a.N, device=dev, generator=g)
val = torch.randn(a.T, a.N, device=dev, generator=g)
st = torch.ones(a.T, a.N, device=dev, dtype=torch.uint8)
st[a.T - 1] = 3
adv = torch.empty_like(rew)
ret = torch.empty_like(rew)
for _ in range(a.reps):
    ops.gae_time_major(rew, val, st, None, discount=0.99, gae_lambda=0.95, adv_out=adv, ret_out=ret)
torch.cuda.synchronize()
print('ok')
