"""Fused training loop alone on the bench workload's shapes: CUDA-event time per minibatch step and
(--timing) the clock64 stamps of one step."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from garage_b200 import ops  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--rows', type=int, default=819200)
    ap.add_argument('--mb', type=int, default=37888)
    ap.add_argument('--epochs', type=int, default=2)
    ap.add_argument('--timing', action='store_true')
    ap.add_argument('--single', action='store_true', help='policy network only')
    a = ap.parse_args()
    dev = torch.device('cuda:0')
    g = torch.Generator(device=dev).manual_seed(0)
    B = a.rows
    shape = ops.MlpShape(3, 64, 1, dev)
    obs = torch.randn(B, 3, device=dev, generator=g)
    act = torch.randn(B, 1, device=dev, generator=g)
    adv = torch.randn(B, device=dev, generator=g)
    olp = torch.randn(B, device=dev, generator=g) * 0.1 - 1.0
    ret = torch.randn(B, device=dev, generator=g)

    def net(policy):
        d = dict(shape=shape, params=torch.randn(shape.P, device=dev, generator=g) * 0.1,
                 adam_m=torch.zeros(shape.P, device=dev), adam_v=torch.zeros(shape.P, device=dev),
                 step_counter=torch.zeros(1, dtype=torch.int64, device=dev), lr=2.5e-4, beta1=0.9, beta2=0.999,
                 eps=1e-8, obs=obs, minibatch=a.mb, epochs=a.epochs, shuffle=ops.SHUFFLE_DEVICE, shuffle_seed=1)
        if policy:
            d.update(act=act, adv=adv, old_logp=olp, loss_kind=ops.LOSS_PPO)
        else:
            d.update(ret=ret, loss_kind=ops.LOSS_VF)
        return d
    nets = [net(True)] if a.single else [net(True), net(False)]
    steps = a.epochs * -(-B // a.mb)
    ops.mlp_train_loop(nets)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.mlp_train_loop(nets)
    e1.record()
    torch.cuda.synchronize()
    print('loop of {} steps x {} nets: {:.1f} us, {:.2f} us per step'.format(
        steps, len(nets), 1e3 * e0.elapsed_time(e1), 1e3 * e0.elapsed_time(e1) / steps))
    if a.timing:
        lib = ops._lib.load()
        per = 296 // len(nets)
        tb = torch.zeros(2 * 296 * 16 + 2 * 296 * 32, dtype=torch.int64, device=dev)
        lib.b2g_mlp_train_loop_debug_timing(tb.data_ptr(), 5)
        ops.mlp_train_loop(nets)
        torch.cuda.synchronize()
        lib.b2g_mlp_train_loop_debug_timing(None, 0)
        raw = tb.cpu().numpy()
        t = raw[:len(nets) * per * 16].reshape(len(nets), per, 16)
        tiles = raw[2 * 296 * 16:2 * 296 * 16 + len(nets) * per * 32].reshape(len(nets), per, 32)
        names = ['step start', 'tiles done', 'partial written', 'barrier 1 passed', 'slice reduced + Adam',
                 'barrier 2 passed', 'weights reloaded']
        for k in range(len(nets)):
            tt = t[k]
            print('--- net', k, ' (median / max cycles since previous stamp; median since start)')
            for i in range(1, 7):
                d = tt[:, i] - tt[:, i - 1]
                print('{:24s} {:8.0f} {:8.0f} {:8.0f}'.format(names[i], np.median(d), d.max(), np.median(tt[:, i] - tt[:, 0])))
        # stamp ids of run_tiles in program order
        order = [2, 4, 5, 6, 7, 8, 3, 9, 10, 11]
        stage = {2: 'bar4 waited', 4: 'h1 + aug panels stored', 5: 'GEMM1 issued', 6: 'GEMM1 done',
                 7: 'EPI1 done (+ next rows published)', 8: 'GEMM2 issued', 3: 'next L1 (+ GEMM3 issue) done',
                 9: 'GEMM2 done', 10: 'EPI2 math + GEMM3 done', 11: 'dz1 stored, GEMM4 issued'}
        for k in range(len(nets)):
            tt = tiles[k]
            print('--- net', k, 'tile stages (median cycles since previous stamp / since step start)')
            prev = tt[:, 0]
            for it in range(2):
                for q in order:
                    i = q + 10 * it
                    print('tile{} {:36s} {:8.0f} {:8.0f}'.format(it, stage[q], np.median(tt[:, i] - prev), np.median(tt[:, i] - tt[:, 0])))
                    prev = tt[:, i]
        for k in range(len(nets)):
            print('--- net', k, 'between barrier-1 arrive and wait: set_step {:.0f}, prime {:.0f}, remaining wait {:.0f}'.format(
                np.median(t[k][:, 8] - t[k][:, 2]), np.median(t[k][:, 9] - t[k][:, 8]), np.median(t[k][:, 3] - t[k][:, 9])))
        for k in range(len(nets)):
            print('--- net', k, 'tiles done -> write_partial entry {:.0f}, write_partial exit -> stamp 2 {:.0f}'.format(
                np.median(tiles[k][:, 28] - t[k][:, 1]), np.median(t[k][:, 2] - tiles[k][:, 30])))
        for k in range(len(nets)):
            tt = tiles[k]
            print('--- net', k, 'write_partial: drain wait {:.0f}, TMEM->smem->global {:.0f} (median cycles)'.format(
                np.median(tt[:, 29] - tt[:, 28]), np.median(tt[:, 30] - tt[:, 29])))
        sm = t[:, :, 7]
        pairs = sum(1 for s in set(sm[0].tolist()) if len(nets) > 1 and s in set(sm[1].tolist()))
        if len(nets) > 1:
            print('CTA j of both nets on the same SM:', int((sm[0] == sm[1]).sum()), 'of', per)
            inv = {int(s_): j for j, s_ in enumerate(sm[1].tolist())}
            print('net-1 CTA sharing the SM of net-0 CTA j:', [inv.get(int(s_), -1) for s_ in sm[0].tolist()])
            print('SM of net-0 CTA j:', [int(x) for x in sm[0].tolist()])
        print('distinct SMs per net:', [len(set(sm[k].tolist())) for k in range(len(nets))], 'SMs hosting both nets:', pairs)


if __name__ == '__main__':
    main()
