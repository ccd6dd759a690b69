"""GPU parity of `normalize(env)` fused into the env kernels (reference envs/normalized_env.py):
 * the single-env Environment API against the reference's own run (tests/golden/normalized_env.npz),
 * the persistent rollout kernel against the oracle under replayed actions (integers bit-exact),
 * the sampler plumbing (running estimates persist across obtain_samples calls)."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import np_oracle as O

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


@pytest.mark.parametrize('case', 'abcd')
def test_normalize_point_env_vs_reference_run(case):
    from garage_b200.envs import PointEnv, normalize
    d = np.load(os.path.join(G, 'normalized_env.npz'))
    gx, gy, arena, bonus, never, max_len = d[case + '_env']
    scale_reward, n_obs, n_rew, es, oa, ra = d[case + '_norm']
    env = normalize(PointEnv(goal=np.array([gx, gy], np.float32), arena_size=arena, done_bonus=bonus,
                             never_done=bool(never), max_episode_length=int(max_len)),
                    scale_reward=scale_reward, normalize_obs=bool(n_obs), normalize_reward=bool(n_rew),
                    expected_action_scale=es, obs_alpha=oa, reward_alpha=ra)
    acts, obs, first = d[case + '_actions'], d[case + '_obs'], list(d[case + '_first'])
    with pytest.raises(RuntimeError):  # the wrapped env's own check (point_env.py:113-114)
        env.step(acts[0])
    k = 0
    for t, a in enumerate(acts):
        if t in first:
            o0, info = env.reset()
            assert np.allclose(o0, obs[k], rtol=1e-5, atol=1e-6)
            assert np.allclose(info['goal'], [gx, gy])
            k += 1
        s = env.step(a)
        assert s.action is a                                              # stored un-scaled (:111)
        assert int(s.step_type) == d[case + '_step_types'][t]             # bit-exact
        assert np.allclose(s.observation, obs[k], rtol=1e-5, atol=1e-6)
        assert abs(s.reward - d[case + '_rewards'][t]) <= 1e-5 * max(1, abs(d[case + '_rewards'][t]))
        k += 1
    om, ov, rm, rv = env.statistics()
    if n_obs:
        assert np.allclose(om, d[case + '_obs_mean'], rtol=1e-5, atol=1e-7)
        assert np.allclose(ov, d[case + '_obs_var'], rtol=1e-5, atol=1e-7)
    if n_rew:
        assert np.allclose([rm, rv], d[case + '_reward_stats'], rtol=1e-5, atol=1e-7)


ENVS = {
    'point': dict(id=0, params=[0.4, 0.3, 5., 2.5, 0.], kw=dict(goal=(0.4, 0.3), done_bonus=2.5), lo=[-.1, -.1], hi=[.1, .1]),
    'pendulum': dict(id=1, params=[8., 2., .05, 10., 1., 1.], kw={}, lo=[-2.], hi=[2.]),
    'cartpole': dict(id=2, params=[9.8, 1.0, 0.1, 0.5, 10.0, 0.02, 2.4, 12 * 2 * math.pi / 360, 1], kw=dict(act_mode=1),
                     lo=[-1.], hi=[1.]),
}
NORMS = {
    'action_only': dict(),
    'obs_reward': dict(normalize_obs=True, normalize_reward=True, scale_reward=0.25, obs_alpha=0.02, reward_alpha=0.05),
    'obs': dict(normalize_obs=True, expected_action_scale=1.5),
    'reward': dict(normalize_reward=True, reward_alpha=0.01, scale_reward=2.0),
}


@pytest.mark.parametrize('name', ['point', 'pendulum', 'cartpole'])
@pytest.mark.parametrize('norm', list(NORMS))
def test_rollout_with_normalize_replay_parity(name, norm):
    from garage_b200 import ops
    from tests.test_rollout_gpu import make_policy
    cfg, nk = ENVS[name], NORMS[norm]
    dev = torch.device('cuda:0')
    Od, A, S = ops.env_dims(cfg['id'])
    N, T, H, max_len, seed, gid0 = 300, 40, 64, 13, 5, 64
    flags = ops.NORM_ACTION | (ops.NORM_OBS if nk.get('normalize_obs') else 0) | \
        (ops.NORM_REWARD if nk.get('normalize_reward') else 0)
    nc = ops.NormConfig(flags, cfg['lo'], cfg['hi'], nk.get('expected_action_scale', 1.), nk.get('scale_reward', 1.),
                        nk.get('obs_alpha', 0.001), nk.get('reward_alpha', 0.001))
    flat = make_policy(Od, H, A, seed=3, log_std=-0.5, scale=2.0)
    pol = torch.from_numpy(flat).to(dev)
    es = ops.EnvState(cfg['id'], N, dev)
    prm = ops.env_params_array(cfg['params'])
    ops.env_reset(es, prm, seed, env_id_offset=gid0, norm=nc)
    buf = ops.RolloutBuffers(cfg['id'], T, N, dev, want_mean=True, want_success=True)
    buf.last_obs.zero_(); buf.ep_len.zero_()
    T1 = 17  # two launches: the running estimates persist in es.norm_stats
    ops.rollout(es, prm, max_len, pol, H, math.log(1e-6), float('inf'), seed, gid0, 0, buf, 0, T1, norm=nc)
    ops.rollout(es, prm, max_len, pol, H, math.log(1e-6), float('inf'), seed, gid0, T1, buf, T1, T - T1, norm=nc)
    torch.cuda.synchronize()
    g = {k: getattr(buf, k).cpu().numpy() for k in ('obs', 'act', 'rew', 'step_type', 'success', 'last_obs', 'ep_len', 'mean')}
    w = O.NormalizedEnvOracle(N, Od, cfg['lo'], cfg['hi'], **nk)
    ref = O.replay_rollout(name, g['act'], max_len, seed, gid0=gid0, env_kw=cfg['kw'], norm=w)
    assert np.array_equal(g['step_type'], ref['step_type'])                 # bit-exact
    done = ref['step_type'] >= O.TERMINAL
    assert done.any()
    assert np.array_equal(g['ep_len'][done], ref['ep_len'][done])           # bit-exact
    assert np.array_equal(g['success'].astype(bool), ref['success'])        # bit-exact
    assert np.allclose(g['obs'], ref['obs'], rtol=1e-5, atol=2e-6)
    assert np.allclose(g['rew'], ref['rew'], rtol=1e-5, atol=2e-6)
    assert np.allclose(g['last_obs'][done], ref['last_obs'][done], rtol=1e-5, atol=2e-6)
    # the policy saw the NORMALISED observation
    p = O.unflatten_params(flat.astype(np.float64), Od, H, A)
    mean, _, _ = O.mlp_forward(p, g['obs'].reshape(-1, Od))
    assert np.allclose(g['mean'].reshape(-1, A), mean, rtol=1e-5, atol=2e-6)
    if es.norm_stats is not None:
        st = es.norm_stats.cpu().numpy()
        if nk.get('normalize_obs'):
            assert np.allclose(st[:Od].T, w.obs_mean, rtol=1e-9, atol=1e-12)
            assert np.allclose(st[Od:2 * Od].T, w.obs_var, rtol=1e-9, atol=1e-12)
        else:  # untouched
            assert (st[:Od] == 0).all() and (st[Od:2 * Od] == 1).all()
        if nk.get('normalize_reward'):
            assert np.allclose(st[2 * Od], w.reward_mean, rtol=1e-9, atol=1e-12)
            assert np.allclose(st[2 * Od + 1], w.reward_var, rtol=1e-9, atol=1e-12)
    else:
        assert norm == 'action_only'


def test_bare_env_unchanged_by_the_norm_entry_points():
    """norm=None / default arguments == the plain kernels, bit for bit."""
    from garage_b200 import ops
    from tests.test_rollout_gpu import make_policy
    dev = torch.device('cuda:0')
    N, T, H, seed = 256, 20, 64, 11
    prm = ops.env_params_array(ENVS['pendulum']['params'])
    pol = torch.from_numpy(make_policy(3, H, 1, seed=2)).to(dev)

    def run(norm):
        es = ops.EnvState(1, N, dev)
        ops.env_reset(es, prm, seed, norm=norm)
        buf = ops.RolloutBuffers(1, T, N, dev)
        ops.rollout(es, prm, 7, pol, H, math.log(1e-6), float('inf'), seed, 0, 0, buf, 0, T, norm=norm)
        return buf
    a, b = run(None), run(ops.NormConfig(0, [0.], [0.]))
    for k in ('obs', 'act', 'rew', 'logp', 'step_type'):
        assert torch.equal(getattr(a, k), getattr(b, k)), k


def test_sampler_with_normalized_pendulum():
    """GPUSampler(envs=normalize(PendulumEnv())): the description reaches the kernels, every env
    copy keeps its running estimates across obtain_samples calls, the batch is well-formed."""
    from garage_b200.envs import PendulumEnv, normalize
    from garage_b200.sampler import LocalSampler
    from garage_b200.torch.policies import GaussianMLPPolicy
    torch.manual_seed(0)
    env = normalize(PendulumEnv(max_episode_length=50), normalize_obs=True, normalize_reward=True, obs_alpha=0.01)
    assert env.spec.max_episode_length == 50 and env.never_terminates
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
    sampler = LocalSampler(agents=policy, envs=env, max_episode_length=50, n_envs=128, seed=3)
    eps = sampler.obtain_samples(0, 128 * 50, policy.get_param_values())
    assert eps.lengths.sum() == 128 * 50 and (np.asarray(eps.lengths) == 50).all()
    s1 = sampler._es.norm_stats.clone()
    assert (s1[:3] != 0).any() and (s1[3:6] != 1).any() and (s1[7] != 1).any()
    raw = GPU_raw_rewards(policy, 50, 128)
    # normalised rewards differ from the bare env's, and they are finite
    assert np.isfinite(eps.rewards).all() and not np.allclose(eps.rewards, raw)
    sampler.obtain_samples(1, 128 * 50, policy.get_param_values())
    assert not torch.equal(s1, sampler._es.norm_stats)
    # |normalised obs| stays moderate once the estimates moved (cos/sin/thdot scaled by ~1/std)
    assert np.abs(eps.observations).max() < 50


def GPU_raw_rewards(policy, L, n):
    from garage_b200.envs import PendulumEnv
    from garage_b200.sampler import LocalSampler
    s = LocalSampler(agents=policy, envs=PendulumEnv(max_episode_length=L), max_episode_length=L, n_envs=n, seed=3)
    return s.obtain_samples(0, n * L, policy.get_param_values()).rewards
