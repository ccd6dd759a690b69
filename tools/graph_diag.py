import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests.test_full_size_gpu import _build, N, T
from garage_b200.torch.algos.graphed import GraphedIteration
def mk():
    algo, policy, vf, sampler = _build()
    return algo, policy, vf, GraphedIteration(algo, N * T, warmup=1)
for trial in range(6):
    A, B = mk(), mk()
    bad = None
    for i in range(5):
        for X in (A, B):
            X[3].run(i)
        torch.cuda.synchronize()
        ea, eb = A[3].eps, B[3].eps
        checks = [('tm.obs', ea.tm.obs, eb.tm.obs), ('tm.act', ea.tm.act, eb.tm.act), ('tm.rew', ea.tm.rew, eb.tm.rew),
                  ('packed.obs', ea.dev.observations, eb.dev.observations), ('packed.logp', ea.dev.logp, eb.dev.logp),
                  ('policy', A[1].flat, B[1].flat), ('vf', A[2].flat, B[2].flat)]
        diff = [n for n, x, y in checks if not torch.equal(x, y)]
        sa, sb = A[0].last_stats, B[0].last_stats
        diff += [k for k in sa if sa[k] != sb[k]]
        if diff:
            bad = (i, diff)
            break
    print('trial', trial, 'first mismatch', bad, flush=True)
