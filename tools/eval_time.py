"""Forward-only pass (mlp_eval_tc_kernel) on the bench shape: 819,200 rows, 64x64, PPO statistics and value baselines."""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from garage_b200 import ops  # noqa: E402


def main():
    dev = torch.device('cuda:0')
    g = torch.Generator(device=dev).manual_seed(0)
    B = int(os.environ.get('ROWS', '819200'))
    sh = ops.MlpShape(3, 64, 1, dev)
    obs = torch.randn(B, 3, device=dev, generator=g)
    act = torch.randn(B, 1, device=dev, generator=g)
    adv = torch.randn(B, device=dev, generator=g)
    olp = torch.randn(B, device=dev, generator=g) * 0.1 - 1.0
    ret = torch.randn(B, device=dev, generator=g)
    flat = torch.randn(sh.P, device=dev, generator=g) * 0.1
    mean_out = torch.empty(B, 1, device=dev)
    cases = dict(ppo_stats=dict(act=act, adv=adv, old_logp=olp, loss_kind=ops.LOSS_PPO, min_log_std=math.log(1e-6)),
                 baselines=dict(ret=ret, loss_kind=ops.LOSS_VF, mean_out=mean_out))
    for name, kw in cases.items():
        for _ in range(3):
            ops.mlp_eval(sh, flat, obs, **kw)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            ops.mlp_eval(sh, flat, obs, **kw)
        e1.record()
        torch.cuda.synchronize()
        us = 1e3 * e0.elapsed_time(e1) / 20
        print('{}: {:.1f} us per pass of {} rows ({:.2f} G rows/s)'.format(name, us, B, B / us / 1e3), flush=True)


if __name__ == '__main__':
    main()
