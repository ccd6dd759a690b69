"""C5: standalone GAE/returns sweep, achieved GB/s (17 algorithmic B/elem) vs measured HBM peak.

    python tools/gae_sweep.py [--quick] [--json out.json]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from garage_b200 import ops  # noqa: E402


def peak_gbs():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        return json.load(open(p))['hbm_gbs'], 'measured'
    return 6650.0, 'fallback'


def time_case(T, N, p_done, iters=20, warmup=3):
    dev = torch.device('cuda:0')
    g = torch.Generator(device=dev).manual_seed(0)
    rew = torch.randn(T, N, device=dev, generator=g)
    val = torch.randn(T, N, device=dev, generator=g)
    st = torch.ones(T, N, device=dev, dtype=torch.uint8)
    if p_done > 0:
        st[torch.rand(T, N, device=dev, generator=g) < p_done] = 2
    st[T - 1] = 3
    adv = torch.empty_like(rew)
    ret = torch.empty_like(rew)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    times = []
    for i in range(warmup + iters):
        flush.zero_()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.gae_time_major(rew, val, st, None, discount=0.99, gae_lambda=0.95, adv_out=adv, ret_out=ret)
        e1.record()
        torch.cuda.synchronize()
        if i >= warmup:
            times.append(e0.elapsed_time(e1))
    times.sort()
    ms = times[len(times) // 2]
    return ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--quick', action='store_true')
    ap.add_argument('--json', default=None)
    a = ap.parse_args()
    peak, how = peak_gbs()
    Ts = [200, 1000] if a.quick else [128, 200, 256, 512, 1000]
    Ns = [4096, 65536, 1048576] if a.quick else [1024, 16384, 65536, 262144, 1048576]
    rows = []
    for T in Ts:
        for N in Ns:
            for p in (0.0, 0.01):
                if 17 * T * N > 40e9:
                    continue
                ms = time_case(T, N, p)
                gbs = 17.0 * T * N / (ms * 1e-3) / 1e9
                rows.append(dict(T=T, N=N, p_done=p, ms=ms, gbs=gbs, frac=gbs / peak,
                                 footprint_mb=17.0 * T * N / 1e6))
                print('T={:5d} N={:8d} p={:.2f}  {:8.3f} ms  {:8.1f} GB/s  {:.3f} of {} peak ({:.0f} MB)'.format(
                    T, N, p, ms, gbs, gbs / peak, how, 17.0 * T * N / 1e6), flush=True)
    if a.json:
        json.dump(dict(peak_gbs=peak, peak_kind=how, rows=rows), open(a.json, 'w'), indent=1)


if __name__ == '__main__':
    main()
