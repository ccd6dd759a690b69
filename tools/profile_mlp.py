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
    ap.add_argument('--timing', action='store_true')
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
    if a.timing:
        tb = torch.zeros(296, 32, dtype=torch.int64, device=dev)
        ops._lib.load().b2g_mlp_debug_timing(tb.data_ptr())
        run(0)
        torch.cuda.synchronize()
        ops._lib.load().b2g_mlp_debug_timing(None)
        t = tb.cpu().numpy()
        names = {0: 'start', 1: 'prologue done', 28: 'loop done', 29: 'last MMA drained', 30: 'partials written'}
        for it in range(2):
            for k, nm in enumerate(['bar4 waited', 'gather published', 'L1 done', 'GEMM1 issued', 'GEMM1 done',
                                    'EPI1 done', 'GEMM2/3 issued', 'GEMM2 done', 'GEMM3 done', 'GEMM4 issued']):
                names[2 + 10 * it + k] = 'tile{} {}'.format(it, nm)
        names[22] = 'tile0 L1 barrier passed (abs)'; names[23] = 'tile0 EPI1 barrier passed (abs)'; names[24] = 'tile0 GEMM2 issued (abs)'
        import numpy as np
        two = t[:, 12] > 0
        print('CTAs with 2 tiles:', int(two.sum()), 'of', t.shape[0])
        for sel, label in ((two, '2-tile CTAs'), (~two, '1-tile CTAs')):
            if not sel.any():
                continue
            tt = t[sel]
            prev = tt[:, 0]
            print('---', label, '(median cycles since previous stamp / since start)')
            for i in sorted(names):
                if (tt[:, i] > 0).all():
                    print('{:28s} {:8.0f} {:8.0f}'.format(names[i], np.median(tt[:, i] - prev), np.median(tt[:, i] - tt[:, 0])))
                    prev = tt[:, i]
    print('mlp path', ops._lib.load().b2g_get_mlp_path(), 'us per launch', 1e3 * e0.elapsed_time(e1) / a.reps)


if __name__ == '__main__':
    main()
