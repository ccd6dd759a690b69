"""GPU parity: env reset/step kernels, the persistent rollout kernel and the
EpisodeBatch pack kernels vs the numpy oracle under REPLAYED actions
(input-replay parity, SURVEY.md section 7: the reference's torch/numpy global RNGs
cannot be reproduced by a counter RNG, so the kernel's own actions are replayed
through the oracle).  Integer bookkeeping is compared bit-exactly."""
import math

import numpy as np
import pytest
import torch

from oracle import np_oracle as O

pytestmark = pytest.mark.gpu

ENVS = {
    'point': dict(id=0, params=[1., 1., 5., 0., 0.], kw=dict(goal=(1., 1.))),
    'point_bonus': dict(id=0, params=[0.3, -0.2, 5., 7.5, 0.], kw=dict(goal=(0.3, -0.2), done_bonus=7.5)),
    'pendulum': dict(id=1, params=[8., 2., .05, 10., 1., 1.], kw={}),
    'cartpole': dict(id=2, params=[9.8, 1.0, 0.1, 0.5, 10.0, 0.02, 2.4, 12 * 2 * math.pi / 360, 0], kw={}),
}


def oracle_env(name):
    return 'point' if name.startswith('point') else name


def make_policy(Od, H, A, seed, log_std=0.0, scale=1.0):
    rng = np.random.default_rng(seed)
    P = O.mlp_param_count(Od, H, A)
    flat = (rng.standard_normal(P) * scale / math.sqrt(H)).astype(np.float32)
    flat[0] = log_std
    return flat


@pytest.mark.parametrize('name', ['point', 'point_bonus', 'pendulum', 'cartpole'])
def test_env_reset_and_single_steps(name):
    from garage_b200 import ops
    cfg = ENVS[name]
    dev = torch.device('cuda:0')
    N, max_len, seed = 257, 9, 1234
    es = ops.EnvState(cfg['id'], N, dev)
    prm = ops.env_params_array(cfg['params'])
    with pytest.raises(RuntimeError):  # point_env.py:113-114
        ops.env_step(es, prm, max_len, torch.zeros(N, ops.env_dims(cfg['id'])[1], device=dev))
    ops.env_reset(es, prm, seed, env_id_offset=5)
    st0, obs0 = O.env_reset_states(oracle_env(name), seed, np.arange(5, 5 + N), np.zeros(N, np.int64),
                                   cfg['kw'].get('goal', (1., 1.)))
    assert np.allclose(es.state.cpu().numpy().T, st0, rtol=0, atol=1e-12)
    assert np.allclose(es.obs.cpu().numpy(), obs0, rtol=1e-6, atol=1e-7)
    assert (es.step_cnt.cpu().numpy() == 0).all() and (es.episode.cpu().numpy() == 1).all()
    rng = np.random.default_rng(0)
    state = st0
    A = ops.env_dims(cfg['id'])[1]
    alive = np.ones(N, bool)
    cnt = np.zeros(N, np.int64)
    for t in range(max_len):
        act = (rng.standard_normal((N, A)) * (0.2 if cfg['id'] == 0 else 1.5)).astype(np.float32)
        if not alive.all():
            break
        obs, rew, st, succ = ops.env_step(es, prm, max_len, torch.from_numpy(act).to(dev), want_success=True)
        state, o_obs, o_rew, o_done, o_succ = O.env_step_any(oracle_env(name), state, act, cfg['kw'])
        cnt += 1
        o_st = np.array([O.get_step_type(int(c), max_len, bool(d)) for c, d in zip(cnt, o_done)], np.uint8)
        assert np.array_equal(st.cpu().numpy(), o_st)                      # bit-exact
        assert np.array_equal(succ.cpu().numpy().astype(bool), o_succ)     # bit-exact
        assert np.allclose(obs.cpu().numpy(), o_obs, rtol=1e-5, atol=1e-6)
        assert np.allclose(rew.cpu().numpy(), o_rew, rtol=1e-5, atol=1e-7)
        alive &= o_st < O.TERMINAL
    # after a TERMINAL/TIMEOUT step the env needs reset() again
    with pytest.raises(RuntimeError):
        ops.env_step(es, prm, max_len, torch.zeros(N, A, device=dev))


def test_point_env_reference_unit_cases():
    """tests/garage/envs/test_point_env.py: reset obs is (0,0,dist); stepping with
    a=goal direction terminates with success (test_done); action not modified."""
    from garage_b200 import ops
    dev = torch.device('cuda:0')
    es = ops.EnvState(0, 1, dev)
    prm = ops.env_params_array([1., 1., 5., 0., 0.])
    ops.env_reset(es, prm, 0)
    assert np.allclose(es.obs.cpu().numpy(), [[0, 0, math.sqrt(2)]])
    a = torch.tensor([[1., 1.]], device=dev)
    a0 = a.clone()
    last = None
    for _ in range(20):
        obs, rew, st, succ = ops.env_step(es, prm, 0, a, want_success=True)
        last = int(st.item())
        if last == O.TERMINAL:
            break
    assert last == O.TERMINAL and bool(succ.item())
    assert torch.equal(a, a0)


@pytest.mark.parametrize('name,H,N,T,max_len', [
    ('pendulum', 64, 300, 50, 20), ('pendulum', 32, 33, 30, 200), ('cartpole', 64, 1500, 120, 60),
    ('point', 32, 700, 60, 25), ('point_bonus', 64, 5000, 40, 40), ('cartpole', 32, 10000, 64, 500),
    # hidden 128: W2 resident in shared memory; N picks 1, 2, 4 and 8 environments per warp
    ('pendulum', 128, 300, 40, 20), ('cartpole', 128, 3000, 40, 30), ('point', 128, 6000, 30, 25),
    ('pendulum', 128, 9001, 24, 10),
])
def test_rollout_replay_parity(name, H, N, T, max_len):
    from garage_b200 import ops
    cfg = ENVS[name]
    dev = torch.device('cuda:0')
    Od, A, S = ops.env_dims(cfg['id'])
    seed, gid0, tc0 = 99, 1000, 17
    log_std = -1.2 if cfg['id'] == 0 else 0.1
    flat = make_policy(Od, H, A, seed=H + N, log_std=log_std, scale=2.0)
    es = ops.EnvState(cfg['id'], N, dev)
    prm = ops.env_params_array(cfg['params'])
    ops.env_reset(es, prm, seed, env_id_offset=gid0)
    buf = ops.RolloutBuffers(cfg['id'], T, N, dev, want_mean=True, want_success=True)
    buf.last_obs.zero_(); buf.ep_len.zero_()
    pol = torch.from_numpy(flat).to(dev)
    # two chunks: state persists between launches, the Philox counter continues
    T1 = T // 3
    ops.rollout(es, prm, max_len, pol, H, math.log(1e-6), float('inf'), seed, gid0, tc0, buf, 0, T1)
    ops.rollout(es, prm, max_len, pol, H, math.log(1e-6), float('inf'), seed, gid0, tc0 + T1, buf, T1, T - T1)
    torch.cuda.synchronize()
    g = {k: getattr(buf, k).cpu().numpy() for k in
         ('obs', 'act', 'rew', 'logp', 'step_type', 'mean', 'success', 'last_obs', 'ep_len')}
    # 1. bookkeeping + dynamics under replayed actions
    ref = O.replay_rollout(oracle_env(name), g['act'], max_len, seed, gid0=gid0, env_kw=cfg['kw'])
    assert np.array_equal(g['step_type'], ref['step_type'])                 # bit-exact
    done = ref['step_type'] >= O.TERMINAL
    assert np.array_equal(g['ep_len'][done], ref['ep_len'][done])           # bit-exact
    assert np.array_equal(g['success'].astype(bool), ref['success'])        # bit-exact
    assert done.any() or T < max_len
    assert np.allclose(g['obs'], ref['obs'], rtol=1e-5, atol=1e-6)
    assert np.allclose(g['rew'], ref['rew'], rtol=1e-5, atol=1e-6)
    assert np.allclose(g['last_obs'][done], ref['last_obs'][done], rtol=1e-5, atol=1e-6)
    # 2. policy forward, sampling and log-prob
    p = O.unflatten_params(flat.astype(np.float64), Od, H, A)
    mean, _, _ = O.mlp_forward(p, g['obs'].reshape(-1, Od))
    assert np.allclose(g['mean'].reshape(-1, A), mean, rtol=1e-5, atol=2e-6)
    ls, _ = O.clamped_log_std(p)
    gid = np.arange(gid0, gid0 + N)
    for t in (0, T1 - 1, T1, T - 1):
        eps = O.action_noise(seed, gid, tc0 + t, A)
        assert np.allclose(g['act'][t], g['mean'][t] + np.float32(math.exp(ls)) * eps, rtol=1e-5, atol=2e-6)
    want_lp = O.normal_log_prob(g['act'].reshape(-1, A).astype(np.float64), g['mean'].reshape(-1, A).astype(np.float64), ls)
    assert np.allclose(g['logp'].reshape(-1), want_lp, rtol=1e-5, atol=1e-5)
    # 3. the noise is standard normal
    z = (g['act'] - g['mean']) / math.exp(ls)
    assert abs(z.mean()) < 5 / math.sqrt(z.size) and abs(z.std() - 1) < 0.02
    # 4. EpisodeBatch packing: order, cut, lengths bit-exact; fields equal
    for num_samples in (1, N * max(1, min(T, max_len)) // 2, 10**9):
        ws, (t_star, n_eps, n_steps, reached) = ops.pack_plan(buf.step_type, buf.ep_len, T, num_samples)
        lengths_o, packed_o, order = O.pack_episodes(g['step_type'], g['ep_len'], num_samples,
                                                     [g['obs'], g['act'], g['rew'], g['step_type']])
        assert n_eps == len(lengths_o) and n_steps == int(lengths_o.sum())
        assert reached == (n_steps >= num_samples)
        if n_eps == 0:
            continue
        lengths, ep_off, last_observations, ep_env, ep_end = ops.pack_index(
            buf.step_type, buf.ep_len, buf.last_obs, ws, T, t_star, n_eps)
        assert np.array_equal(lengths.cpu().numpy(), lengths_o)             # bit-exact
        assert [(int(t), int(n)) for t, n in zip(ep_end.cpu(), ep_env.cpu())] == order
        outs = ops.pack_fields(buf.step_type, buf.ep_len, ep_off, t_star, n_steps,
                               [buf.obs, buf.act, buf.rew, buf.step_type])
        for got, want in zip(outs, packed_o):
            assert np.array_equal(got.cpu().numpy(), want)
        # tiled (row map + transpose) and single-kernel scatter forms agree, also for a field wider
        # than one 4-word pass and for a shared row map
        wide = torch.arange(buf.rew.numel() * 6, dtype=torch.float32, device=dev).reshape(*buf.rew.shape, 6)
        row_map = ops.pack_rowmap(buf.step_type, buf.ep_len, ep_off, t_star)
        assert int((row_map >= 0).sum()) == n_steps
        a1 = ops.pack_fields(buf.step_type, buf.ep_len, ep_off, t_star, n_steps, [wide, buf.step_type, buf.rew],
                             row_map=row_map)
        a2 = ops.pack_fields_scatter(buf.step_type, buf.ep_len, ep_off, t_star, n_steps, [wide, buf.step_type, buf.rew])
        for x, y in zip(a1, a2):
            assert torch.equal(x, y)
        want_last = np.stack([g['last_obs'][t, n] for t, n in order])
        assert np.array_equal(last_observations.cpu().numpy(), want_last)


def test_rollout_independent_of_sharding():
    """Results depend on the GLOBAL env id only: two half-size shards == one launch."""
    from garage_b200 import ops
    cfg = ENVS['pendulum']
    dev = torch.device('cuda:0')
    N, T, H, seed = 512, 25, 64, 7
    flat = torch.from_numpy(make_policy(3, H, 1, seed=1)).to(dev)
    prm = ops.env_params_array(cfg['params'])

    def run(n, off):
        es = ops.EnvState(1, n, dev)
        ops.env_reset(es, prm, seed, env_id_offset=off)
        buf = ops.RolloutBuffers(1, T, n, dev)
        ops.rollout(es, prm, 10, flat, H, math.log(1e-6), float('inf'), seed, off, 0, buf, 0, T)
        return buf
    full, a, b = run(N, 0), run(N // 2, 0), run(N // 2, N // 2)
    for k in ('obs', 'act', 'rew', 'logp', 'step_type'):
        assert torch.equal(getattr(full, k), torch.cat([getattr(a, k), getattr(b, k)], dim=1)), k


def test_static_pack_plan_is_reused_without_changing_the_batch():
    """Never-terminating envs: the pack plan / row map of a shape is computed once and reused; the
    second batch equals the one a sampler without the cached plan produces, and a different
    num_samples (another shape) re-plans."""
    from garage_b200.envs import PendulumEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.policies import GaussianMLPPolicy
    env = PendulumEnv(max_episode_length=25)
    torch.manual_seed(4)
    pol = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
    a = GPUSampler(pol, env, max_episode_length=25, n_envs=96, seed=9)
    b = GPUSampler(pol, env, max_episode_length=25, n_envs=96, seed=9)
    a.obtain_samples(0, 96 * 25, pol)
    b.obtain_samples(0, 96 * 25, pol)
    assert a._static_pack is not None
    map_before = a._static_pack[2]
    b._static_pack = None                      # b plans again, a reuses
    ea, eb = a.obtain_samples(1, 96 * 25, pol), b.obtain_samples(1, 96 * 25, pol)
    assert a._static_pack[2] is map_before and ea.row_map is map_before
    for name in ('observations', 'actions', 'rewards', 'logp', 'step_types', 'mean', 'last_observations',
                 'lengths', 'ep_env', 'ep_end'):
        assert torch.equal(getattr(ea.dev, name), getattr(eb.dev, name)), name
    assert torch.equal(ea.row_map, eb.row_map)
    e2 = a.obtain_samples(2, 96 * 50, pol)    # two episodes per env: another shape
    assert a._static_pack[2] is not map_before and e2.n_episodes == 192 and list(e2.lengths[:3]) == [25] * 3
