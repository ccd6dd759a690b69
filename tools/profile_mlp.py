"""Minibatch kernel alone (policy PPO loss, 64x64, 32,768-row minibatches gathered from an
819,200-row batch) -- the ncu target for the dominant kernel.  Prints CUDA-event timings."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from garage_b200 import ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--rows', type=int, default=819200)
    ap.add_argument('--mb', type=int, default=32768)
    ap.add_argument('--reps', type=int, default=20)
    ap.add_argument('--vf', action='store_true')
    a = ap.parse_args()
    dev = torch.device('cuda:0')
    g = torch.Generator(device=dev).manual_seed(0)
    B, mb = a.rows, a.mb
    shape = ops.MlpShape(3, 64, 1, dev)
    flat = torch.randn(shape.P, device=dev, generator=g) * 0.1
    obs = torch.randn(B, 3, device=dev, generator=g)
    act = torch.randn(B, 1, device=dev, generator=g)
    adv = torch.randn(B, device=dev, generator=g)
    olp = torch.randn(B, device=dev, generator=g) * 0.1 - 1.0
    ret = torch.randn(B, device=dev, generator=g)
    idx = ops.permutation(B, 0, device=dev)
    kw = dict(ret=ret, loss_kind=ops.LOSS_VF) if a.vf else dict(act=act, adv=adv, old_logp=olp, loss_kind=ops.LOSS_PPO)

    def run(r):
        s = (r * mb) % max(B - mb + 1, 1)
        ops.mlp_minibatch_grad(shape, flat, obs, idx=idx[s:s + mb], mb=mb, **kw)
    for r in range(3):
        run(r)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(a.reps):
        run(r)
    e1.record()
    torch.cuda.synchronize()
    print('mlp path', ops._lib.load().b2g_get_mlp_path(), 'us per launch', 1e3 * e0.elapsed_time(e1) / a.reps)


if __name__ == '__main__':
    main()
