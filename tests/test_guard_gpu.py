"""Out-of-bounds WRITE detection without compute-sanitizer (the tool is closed on this GPU pool: see
profiles/r2_sanitizer.md).  Every buffer a kernel writes is carved out of a larger allocation with
4 KB canary zones on both sides; after the launch the canaries must be intact.  Covers the kernels
with hand-computed indexing: rollout (time-major stores), GAE (sequential and time-split), the
minibatch kernels' partial rows and mean outputs (tcgen05, CUDA-core, general shapes), the
Fisher-vector product, and both persistent training loops (parameters, Adam moments, workspace)."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GUARD = 1024  # elements on each side
PAT = {torch.float32: 1.2345e30, torch.float64: 1.2345e300, torch.uint8: 0xA5, torch.int32: 0x5A5A5A5A, torch.int64: 0x5A5A5A5A5A5A}


class Guarded:
    def __init__(self):
        self.items = []

    def make(self, shape, dtype=torch.float32, fill=None):
        n = int(np.prod(shape))
        raw = torch.full((n + 2 * GUARD, ), PAT[dtype], dtype=dtype, device='cuda')
        view = raw[GUARD:GUARD + n].view(*shape)
        if fill is not None:
            view.copy_(fill) if torch.is_tensor(fill) else view.fill_(fill)
        self.items.append((raw, n, dtype))
        return view

    def check(self):
        torch.cuda.synchronize()
        for raw, n, dtype in self.items:
            lo, hi = raw[:GUARD], raw[GUARD + n:]
            want = torch.full((GUARD, ), PAT[dtype], dtype=dtype, device='cuda')
            assert torch.equal(lo, want), 'write below a {} buffer of {} elements'.format(dtype, n)
            assert torch.equal(hi, want), 'write past a {} buffer of {} elements'.format(dtype, n)


@pytest.mark.parametrize('T,N', [(7, 33), (200, 64), (50, 4096), (33, 40000), (128, 1025)])
def test_gae_outputs_stay_in_bounds(T, N):
    from garage_b200 import ops
    g = Guarded()
    rew, val = torch.randn(T, N, device='cuda'), torch.randn(T, N, device='cuda')
    st = torch.ones(T, N, dtype=torch.uint8, device='cuda')
    st[T - 1] = 3
    adv, ret = g.make((T, N)), g.make((T, N))
    ops.gae_time_major(rew, val, st, discount=0.99, gae_lambda=0.95, adv_out=adv, ret_out=ret)
    g.check()
    assert bool(torch.isfinite(adv).all()) and bool(torch.isfinite(ret).all())


@pytest.mark.parametrize('env_name,n_envs,T', [('pendulum', 33, 17), ('cartpole', 64, 40), ('point', 100, 25)])
def test_rollout_buffers_stay_in_bounds(env_name, n_envs, T):
    from garage_b200 import ops
    from garage_b200.envs import CartPoleEnv, PendulumEnv, PointEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.policies import GaussianMLPPolicy
    env = dict(pendulum=PendulumEnv, cartpole=CartPoleEnv, point=PointEnv)[env_name](max_episode_length=T)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
    s = GPUSampler(policy, env, max_episode_length=T, n_envs=n_envs, seed=2)
    s._ensure_state()
    s._ensure_buffers(T)
    g = Guarded()
    b = s._buf
    for name in ('obs', 'act', 'rew', 'logp', 'step_type', 'mean', 'success', 'last_obs', 'ep_len'):
        t = getattr(b, name)
        if t is not None:
            setattr(b, name, g.make(tuple(t.shape), t.dtype, fill=0))
    eps = s.obtain_samples(0, n_envs * T, policy)
    g.check()
    assert eps.n_steps >= n_envs * T or env_name != 'pendulum'


@pytest.mark.parametrize('Od,H,A,path', [(3, 64, 1, 0), (3, 64, 2, 0), (3, 32, 2, 1), (6, 128, 3, 3), (8, 32, 4, 3), (4, 64, 1, 1)])
@pytest.mark.parametrize('n', [1, 129, 20001])
def test_minibatch_kernels_stay_in_bounds(Od, H, A, path, n):
    from garage_b200 import ops
    from tests.test_mlp_gpu import make_case
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, n, seed=n)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    prev = ops.mlp_path()
    try:
        ops.set_mlp_path(path)
        sh = ops.MlpShape(Od, H, A, torch.device('cuda:0'))
        g = Guarded()
        sh.ws = g.make((sh.ws.numel(), ))
        mean_out = g.make((n, A))
        grad, stats = g.make((sh.P, )), g.make((8, ), torch.float64)
        nb = ops.mlp_minibatch_grad(sh, t(flat), t(obs), act=t(act), adv=t(adv), old_logp=t(old_logp),
                                    loss_kind=ops.LOSS_PPO, min_log_std=math.log(1e-6))
        ops.mlp_reduce(sh, nb, 1.0 / n, grad_out=grad, stats_out=stats)
        nb = ops.mlp_eval(sh, t(flat), t(obs), act=t(act), adv=t(adv), old_logp=t(old_logp), loss_kind=ops.LOSS_PPO,
                          min_log_std=math.log(1e-6), mean_out=mean_out)
        ops.mlp_reduce(sh, nb, 1.0, stats_out=stats)
        out, scratch = g.make((sh.P, )), g.make((n * A, ))
        ops.mlp_fvp(sh, t(flat), torch.randn(sh.P, device='cuda'), t(obs), min_log_std=math.log(1e-6), out=out, scratch=scratch)
        g.check()
        assert bool(torch.isfinite(grad).all()) and bool(torch.isfinite(out).all())
    finally:
        ops.set_mlp_path(prev)


@pytest.mark.parametrize('Od,H,A,path,B,mb', [(3, 64, 1, 0, 5000, 1024), (3, 64, 2, 0, 700, None), (3, 32, 2, 0, 1000, 64),
                                               (6, 128, 3, 3, 900, 300), (4, 64, 1, 3, 3000, 512)])
def test_training_loops_stay_in_bounds(Od, H, A, path, B, mb):
    from garage_b200 import ops
    from tests.test_mlp_gpu import make_case
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, B, seed=B)
    vflat = make_case(Od, H, 1, 8, seed=1)[0]
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    prev = ops.mlp_path()
    try:
        ops.set_mlp_path(path)
        g = Guarded()
        dev = torch.device('cuda:0')

        def net(sh, f0, policy):
            d = dict(shape=sh, params=g.make((sh.P, ), fill=t(f0)), adam_m=g.make((sh.P, ), fill=0.0), adam_v=g.make((sh.P, ), fill=0.0),
                     step_counter=g.make((1, ), torch.int64, fill=0), lr=2.5e-4, beta1=0.9, beta2=0.999, eps=1e-8, obs=t(obs),
                     minibatch=mb, epochs=2, shuffle=ops.SHUFFLE_DEVICE if mb else ops.SHUFFLE_NONE, shuffle_seed=3)
            if policy:
                d.update(act=t(act), adv=t(adv), old_logp=t(old_logp), loss_kind=ops.LOSS_PPO, min_log_std=math.log(1e-6))
            else:
                d.update(ret=t(ret), loss_kind=ops.LOSS_VF)
            return d
        nets = [net(ops.MlpShape(Od, H, A, dev), flat, True), net(ops.MlpShape(Od, H, 1, dev), vflat, False)]
        need = int(ops._lib.load().b2g_mlp_train_loop_ws_floats())
        ops._loop_ws[dev] = g.make((need, ))
        try:
            ops.mlp_train_loop(nets)
            g.check()
        finally:
            ops._loop_ws.pop(dev, None)
        steps = 2 * (-(-B // mb) if mb else 1)
        assert int(nets[0]['step_counter'].item()) == steps
        assert bool(torch.isfinite(nets[0]['params']).all()) and bool(torch.isfinite(nets[1]['params']).all())
    finally:
        ops.set_mlp_path(prev)
