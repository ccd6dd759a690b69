"""BASELINE configs[1] at its full size (Pendulum, 64x64, 4,096 envs x 200 steps, minibatch 32,768):
too large for the numpy oracle to referee in seconds, so the checks are size-independent properties
(recurrences that define GAE / returns, the packing being the transposition, fused == step-wise,
graph replay == eager) -- the small-size tests pin the same kernels to the oracle and the golden
fixtures."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

N, T, MB, E = 4096, 200, 32768, 2


def _build(update_mode='auto', seed=3):
    from garage_b200.envs import PendulumEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.algos import PPO
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    torch.manual_seed(seed)
    env = PendulumEnv(max_episode_length=T)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
    vf = GaussianMLPValueFunction(env.spec, hidden_sizes=(64, 64))
    sampler = GPUSampler(policy, env, max_episode_length=T, n_envs=N, seed=11)
    mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), m, max_optimization_epochs=E, minibatch_size=MB)
    algo = PPO(env.spec, policy, vf, sampler, policy_optimizer=mk(policy), vf_optimizer=mk(vf), discount=0.99,
               gae_lambda=0.95, update_mode=update_mode)
    return algo, policy, vf, sampler


def test_rollout_bookkeeping_and_packing_full_size():
    """Integer bookkeeping is exact at full size: every episode is FIRST, MID x 198, TIMEOUT with length
    200, and the packed (episode-major) batch is the transposition of the time-major buffers."""
    algo, policy, vf, sampler = _build()
    eps = sampler.obtain_samples(0, N * T, policy)
    tm = eps.tm
    st = tm.step_type[:T]
    assert bool((st[0] == 0).all()) and bool((st[1:T - 1] == 1).all()) and bool((st[T - 1] == 3).all())
    assert eps.n_episodes == N and eps.n_steps == N * T
    assert torch.equal(eps.dev.lengths, torch.full((N,), T, dtype=torch.int64, device=st.device))
    # episodes complete in the same row -> ordered by env index: packed row n*T + t == (t, n)
    assert torch.equal(eps.dev.observations.reshape(N, T, 3), tm.obs[:T].transpose(0, 1))
    assert torch.equal(eps.dev.actions.reshape(N, T, 1), tm.act[:T].transpose(0, 1))
    assert torch.equal(eps.dev.rewards.reshape(N, T), tm.rew[:T].t())
    assert torch.equal(eps.dev.step_types.reshape(N, T), st.t())
    # obs = [cos th, sin th, thdot]: on the unit circle, speed clipped at 8 (gym pendulum)
    o = eps.dev.observations
    assert float(((o[:, 0]**2 + o[:, 1]**2) - 1).abs().max()) < 1e-5 and float(o[:, 2].abs().max()) <= 8.0
    # rewards are -(angle^2 + .1 thdot^2 + .001 u^2) <= 0, bounded by pi^2 + 6.4 + 0.004
    assert float(eps.dev.rewards.max()) <= 0 and float(eps.dev.rewards.min()) > -(np.pi**2 + 6.4 + 0.0041)


def test_gae_and_returns_recurrences_full_size():
    """A_t = delta_t + gamma lambda A_{t+1},  R_t = r_t + gamma R_{t+1}, with A_T = R_T = V_T = 0
    (torch/_functions.py:25-85, np/_functions.py:111-128), checked on all 819,200 steps."""
    from garage_b200 import ops
    algo, policy, vf, sampler = _build()
    eps = sampler.obtain_samples(0, N * T, policy)
    tm = eps.tm
    g, lam = 0.99, 0.95
    val = torch.randn(T, N, device=tm.rew.device)
    adv, ret = ops.gae_time_major(tm.rew[:T], val, tm.step_type[:T], discount=g, gae_lambda=lam)
    r, v = tm.rew[:T].double(), val.double()
    v_next = torch.cat([v[1:], torch.zeros(1, N, dtype=torch.float64, device=v.device)])
    delta = r + g * v_next - v
    a = adv.double()
    a_next = torch.cat([a[1:], torch.zeros(1, N, dtype=torch.float64, device=v.device)])
    assert float((a - (delta + g * lam * a_next)).abs().max()) < 1e-4 * float(a.abs().max())
    R = ret.double()
    R_next = torch.cat([R[1:], torch.zeros(1, N, dtype=torch.float64, device=v.device)])
    assert float((R - (r + g * R_next)).abs().max()) < 1e-5 * float(R.abs().max())


def test_fused_loop_matches_stepwise_full_size():
    """One full PPO iteration (2 epochs x 25 minibatches of 32,768 rows, both networks): the persistent
    tensor-core loop and the per-minibatch launches end at the same parameters (1e-5 relative)."""
    out = {}
    for mode in ('fused', 'stepwise'):
        algo, policy, vf, sampler = _build(mode)
        eps = sampler.obtain_samples(0, N * T, policy)
        algo._train_once(0, eps)
        torch.cuda.synchronize()
        out[mode] = (policy.flat.clone(), vf.flat.clone(), int(algo._policy_optimizer.state()[2].item()), dict(algo.last_stats))
    pf, vf_, sf, lf = out['fused']
    ps, vs, ss, ls = out['stepwise']
    assert sf == ss == E * (N * T // MB)
    rel = lambda x, y: float((x.double() - y.double()).norm() / y.double().norm())
    assert rel(pf, ps) < 1e-5 and rel(vf_, vs) < 1e-5
    for k in ('GaussianMLPPolicy/LossAfter', 'GaussianMLPPolicy/KL', 'GaussianMLPValueFunction/LossAfter',
              'Evaluation/AverageReturn'):
        assert abs(lf[k] - ls[k]) <= 1e-5 * max(1.0, abs(ls[k])) + 1e-7, k


def test_graph_replay_matches_eager_full_size():
    """bench.py replays the whole iteration as one CUDA graph: three replays leave the same parameters as
    three eager iterations (every counter that advances lives on the device)."""
    from garage_b200.torch.algos.graphed import GraphedIteration
    res = []
    for graphed in (False, True):
        algo, policy, vf, sampler = _build()
        if graphed:
            runner = GraphedIteration(algo, N * T, warmup=1)
            for i in range(3):
                runner.run(i)
        else:
            for i in range(3):
                eps = sampler.obtain_samples(i, N * T, policy)
                algo._train_once(i, eps)
        torch.cuda.synchronize()
        res.append((policy.flat.clone(), vf.flat.clone()))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])


def test_cartpole_16384_envs_ragged_bookkeeping_full_size():
    """BASELINE configs[2]: 16,384 CartPole envs, early termination.  Integer bookkeeping at full size:
    lengths sum to the batch, every episode is FIRST .. (TERMINAL | TIMEOUT), the row map is a bijection
    onto the packed rows and the packed fields are the mapped time-major ones (bit-exact)."""
    from garage_b200.envs import CartPoleEnv
    from garage_b200.sampler import GPUSampler
    from garage_b200.torch.policies import GaussianMLPPolicy
    n, t_max = 16384, 200
    torch.manual_seed(0)
    env = CartPoleEnv(max_episode_length=t_max)
    policy = GaussianMLPPolicy(env.spec, hidden_sizes=(64, 64))
    sampler = GPUSampler(policy, env, max_episode_length=t_max, n_envs=n, seed=5)
    eps = sampler.obtain_samples(0, n * t_max, policy)
    lengths = eps.dev.lengths
    assert int(lengths.sum()) == eps.n_steps >= n * t_max and lengths.numel() == eps.n_episodes
    assert int(lengths.min()) >= 1 and int(lengths.max()) <= t_max
    st = eps.dev.step_types.long()
    ends = torch.cumsum(lengths, 0) - 1
    starts = ends - lengths + 1
    is_end = torch.zeros(eps.n_steps, dtype=torch.bool, device=st.device)
    is_end[ends] = True
    assert bool((st[ends] >= 2).all()) and bool((st[~is_end] <= 1).all())       # TERMINAL/TIMEOUT only at the ends
    multi = lengths > 1
    assert bool((st[starts[multi]] == 0).all())                                 # FIRST (a 1-step episode is TERMINAL)
    assert bool((st[ends[lengths == t_max]] == 3).all())                        # the time limit wins: TIMEOUT
    assert bool((st[ends[lengths < t_max]] == 2).all())                         # early end: TERMINAL
    rm = eps.row_map
    valid = rm >= 0
    assert int(valid.sum()) == eps.n_steps
    assert torch.equal(torch.sort(rm[valid].long())[0], torch.arange(eps.n_steps, device=rm.device))
    rows = rm[valid].long()
    tm = eps.tm
    T_used = rm.shape[0]
    assert torch.equal(eps.dev.observations[rows], tm.obs[:T_used][valid])
    assert torch.equal(eps.dev.rewards[rows], tm.rew[:T_used][valid])
    assert torch.equal(eps.dev.step_types[rows], tm.step_type[:T_used][valid])


def test_graphed_iterations_are_bit_reproducible_full_size():
    """Two identically seeded runners replayed in lockstep must agree bit for bit after every iteration
    (rollout buffers, packed batch, both parameter vectors).  Back-to-back graph replays are where a
    missing barrier inside the persistent loop kernel showed up as a run-to-run difference."""
    from garage_b200.torch.algos.graphed import GraphedIteration
    for trial in range(4):
        runs = []
        for _ in range(2):
            algo, policy, vf, sampler = _build()
            runs.append((policy, vf, GraphedIteration(algo, N * T, warmup=1)))
        for i in range(3):
            for r in runs:
                r[2].run(i)
            torch.cuda.synchronize()
            a, b = runs
            assert torch.equal(a[2].eps.tm.act, b[2].eps.tm.act), (trial, i)
            assert torch.equal(a[2].eps.dev.observations, b[2].eps.dev.observations), (trial, i)
            assert torch.equal(a[0].flat, b[0].flat), (trial, i)
            assert torch.equal(a[1].flat, b[1].flat), (trial, i)
