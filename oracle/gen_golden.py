"""Generate tests/golden/*.npz by RUNNING the unmodified reference
(/root/reference/src/garage, v2021.03.0) in the build container through the import
shims of oracle/ref_shims.py.  Test infrastructure only.

    python -m oracle.gen_golden          # rewrites tests/golden/

The fixtures travel to the GPU box (the reference cannot) and pin both the oracle
(tests/test_golden.py, CPU) and the CUDA path (tests/test_golden_gpu.py).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, 'tests', 'golden')


def flat_of(module_sd, prefix):
    """state_dict -> garage_b200's flat order [log_std|W1|b1|W2|b2|W3|b3]."""
    keys = ['_init_std', '_mean_module._layers.0.linear.weight', '_mean_module._layers.0.linear.bias',
            '_mean_module._layers.1.linear.weight', '_mean_module._layers.1.linear.bias',
            '_mean_module._output_layers.0.linear.weight', '_mean_module._output_layers.0.linear.bias']
    return np.concatenate([module_sd[prefix + k].detach().numpy().reshape(-1) for k in keys]).astype(np.float32)


def flat_grad_of(module, prefix):
    named = dict(module.named_parameters())
    keys = ['_init_std', '_mean_module._layers.0.linear.weight', '_mean_module._layers.0.linear.bias',
            '_mean_module._layers.1.linear.weight', '_mean_module._layers.1.linear.bias',
            '_mean_module._output_layers.0.linear.weight', '_mean_module._output_layers.0.linear.bias']
    return np.concatenate([named[prefix + k].grad.numpy().reshape(-1) for k in keys]).astype(np.float32)


def make_spec(obs_dim, act_dim, max_len):
    import akro
    from garage import EnvSpec
    return EnvSpec(observation_space=akro.Box(low=-np.inf, high=np.inf, shape=(obs_dim, )),
                   action_space=akro.Box(low=-1., high=1., shape=(act_dim, )),
                   max_episode_length=max_len)


def gen_gae():
    from garage.np import discount_cumsum
    from garage.torch import compute_advantages
    rng = np.random.default_rng(0)
    out = {}
    for name, (N, P, g, l) in dict(a=(5, 50, 0.99, 0.95), b=(3, 200, 0.99, 0.97), c=(4, 33, 0.9, 1.0),
                                   d=(2, 1000, 0.995, 0.9)).items():
        b = rng.standard_normal((N, P)).astype(np.float32)
        r = rng.standard_normal((N, P)).astype(np.float32)
        adv = compute_advantages(g, l, P, torch.from_numpy(b), torch.from_numpy(r)).numpy()
        ret = np.stack([discount_cumsum(row.astype(np.float64), g) for row in r])
        out.update({name + '_baselines': b, name + '_rewards': r, name + '_adv': adv,
                    name + '_ret': ret, name + '_cfg': np.array([g, l])})
    np.savez_compressed(os.path.join(OUT, 'gae.npz'), **out)


def gen_point_env():
    from garage.envs import PointEnv
    rng = np.random.default_rng(1)
    out = {}
    cfgs = dict(a=dict(goal=(1., 1.), max_episode_length=30),
                b=dict(goal=(0.3, -0.2), done_bonus=7.5, max_episode_length=40),
                c=dict(goal=(0.25, 0.25), never_done=True, max_episode_length=12))
    for name, kw in cfgs.items():
        env = PointEnv(**{k: (np.array(v, dtype=np.float32) if k == 'goal' else v) for k, v in kw.items()})
        obs0, info = env.reset()
        T = kw['max_episode_length']
        goal = np.array(kw['goal'], np.float32)
        acts, obs, rews, sts, succ = [], [obs0], [], [], []
        for t in range(T):
            # head towards the goal with noise so that some runs terminate early
            a = (0.5 * (goal - obs[-1][:2]) + rng.standard_normal(2) * 0.08).astype(np.float32)
            es = env.step(a)
            acts.append(a); obs.append(es.observation); rews.append(es.reward)
            sts.append(int(es.step_type)); succ.append(bool(es.env_info['success']))
            if es.last:
                break
        out.update({name + '_actions': np.array(acts, np.float32), name + '_obs': np.array(obs, np.float32),
                    name + '_rewards': np.array(rews, np.float64), name + '_step_types': np.array(sts, np.uint8),
                    name + '_success': np.array(succ), name + '_params': np.array(
                        [goal[0], goal[1], 5., kw.get('done_bonus', 0.), float(kw.get('never_done', False)),
                         kw['max_episode_length']], np.float64)})
    np.savez_compressed(os.path.join(OUT, 'point_env.npz'), **out)


def gen_normalized_env():
    """`normalize(PointEnv(...), ...)` of the reference over several episodes (the wrapper's
    running estimates persist across reset()): actions in, observations / rewards / final
    estimates out."""
    from garage.envs import PointEnv, normalize
    rng = np.random.default_rng(5)
    out = {}
    cfgs = dict(
        a=dict(env=dict(goal=(1., 1.), max_episode_length=20), norm=dict()),  # default: action scaling only
        b=dict(env=dict(goal=(0.5, -0.4), max_episode_length=25),
               norm=dict(normalize_obs=True, normalize_reward=True, scale_reward=0.5, obs_alpha=0.05,
                         reward_alpha=0.02)),
        c=dict(env=dict(goal=(0.3, 0.3), done_bonus=2., max_episode_length=30),
               norm=dict(normalize_obs=True, expected_action_scale=2., obs_alpha=0.01)),
        d=dict(env=dict(goal=(2., 2.), max_episode_length=15),
               norm=dict(normalize_reward=True, scale_reward=3., reward_alpha=0.1)))
    for name, c in cfgs.items():
        kw = c['env']
        env = normalize(PointEnv(**{k: (np.array(v, dtype=np.float32) if k == 'goal' else v) for k, v in kw.items()}),
                        **c['norm'])
        goal = np.array(kw['goal'], np.float32)
        scale = c['norm'].get('expected_action_scale', 1.)
        acts, obs, rews, sts, first = [], [], [], [], []
        for ep in range(3):
            o, _ = env.reset()
            first.append(len(acts))
            obs.append(np.asarray(o, np.float64))
            pos = np.zeros(2)
            for t in range(kw['max_episode_length']):
                # in units of expected_action_scale: towards the goal, noisy, sometimes out of range
                a = (scale * np.clip(3. * (goal - pos), -1.3, 1.3) + rng.standard_normal(2) * 0.3 * scale).astype(np.float32)
                es = env.step(a)
                pos = np.asarray(env._env._point, np.float64)
                acts.append(a); obs.append(np.asarray(es.observation, np.float64)); rews.append(float(es.reward))
                sts.append(int(es.step_type))
                if es.last:
                    break
        nk = c['norm']
        out.update({name + '_actions': np.array(acts, np.float32), name + '_obs': np.array(obs, np.float64),
                    name + '_rewards': np.array(rews, np.float64), name + '_step_types': np.array(sts, np.uint8),
                    name + '_first': np.array(first, np.int64),
                    name + '_env': np.array([goal[0], goal[1], 5., kw.get('done_bonus', 0.), 0.,
                                             kw['max_episode_length']], np.float64),
                    name + '_norm': np.array([nk.get('scale_reward', 1.), float(nk.get('normalize_obs', False)),
                                              float(nk.get('normalize_reward', False)),
                                              nk.get('expected_action_scale', 1.), nk.get('obs_alpha', 0.001),
                                              nk.get('reward_alpha', 0.001)], np.float64),
                    name + '_obs_mean': np.asarray(env._obs_mean, np.float64),
                    name + '_obs_var': np.asarray(env._obs_var, np.float64),
                    name + '_reward_stats': np.array([env._reward_mean, env._reward_var], np.float64)})
    np.savez_compressed(os.path.join(OUT, 'normalized_env.npz'), **out)


def gen_losses():
    """PPO / VPG / value-function losses and per-parameter gradients from the reference's
    own autograd graph (`_compute_loss_with_adv().backward()`, `compute_loss().backward()`),
    then three Adam steps over fixed minibatch index lists."""
    from garage.torch.algos import PPO, VPG
    from garage.torch.policies import GaussianMLPPolicy
    from garage.torch.value_functions import GaussianMLPValueFunction
    out = {}
    for k, (name, (O_, A, H, n)) in enumerate(dict(p64=(3, 1, 64, 600), p32=(3, 2, 32, 257),
                                                   c64=(4, 1, 64, 300)).items()):
        torch.manual_seed(100 + k)
        rng = np.random.default_rng(200 + k)
        spec = make_spec(O_, A, 200)
        policy = GaussianMLPPolicy(spec, hidden_sizes=(H, H))
        vf = GaussianMLPValueFunction(spec, hidden_sizes=(H, H))
        # move away from the zero-bias init so every gradient path is exercised
        with torch.no_grad():
            for p in list(policy.parameters()) + list(vf.parameters()):
                p.add_(torch.randn_like(p) * 0.1)
        algo = PPO(spec, policy, vf, sampler=None)
        vpg = VPG(spec, policy, vf, sampler=None)
        # the old policy differs from the current one (ratio != 1, some rows clipped)
        with torch.no_grad():
            for p in algo._old_policy.parameters():
                p.add_(torch.randn_like(p) * 0.05)
        obs = torch.Tensor(rng.standard_normal((n, O_)))
        act = torch.Tensor(rng.standard_normal((n, A)))
        adv = torch.Tensor(rng.standard_normal(n))
        ret = torch.Tensor(rng.standard_normal(n) * 3)
        rew = torch.zeros(n)
        with torch.no_grad():
            old_ll = algo._old_policy(obs)[0].log_prob(act)
        out[name + '_dims'] = np.array([O_, A, H, n])
        out[name + '_policy_flat'] = flat_of(policy.state_dict(), '_module.')
        out[name + '_vf_flat'] = flat_of(vf.state_dict(), 'module.')
        out[name + '_obs'], out[name + '_act'] = obs.numpy(), act.numpy()
        out[name + '_adv'], out[name + '_ret'], out[name + '_old_logp'] = adv.numpy(), ret.numpy(), old_ll.numpy()
        for tag, al in (('ppo', algo), ('vpg', vpg)):
            policy.zero_grad()
            loss = al._compute_loss_with_adv(obs, act, rew, adv)
            loss.backward()
            out['{}_{}_loss'.format(name, tag)] = np.float64(loss.item())
            out['{}_{}_grad'.format(name, tag)] = flat_grad_of(policy, '_module.')
        vf.zero_grad()
        loss = vf.compute_loss(obs, ret)
        loss.backward()
        out[name + '_vf_loss'] = np.float64(loss.item())
        out[name + '_vf_grad'] = flat_grad_of(vf, 'module.')
        with torch.no_grad():
            new_dist, old_dist = policy(obs)[0], algo._old_policy(obs)[0]
            out[name + '_kl'] = np.float64(torch.distributions.kl.kl_divergence(old_dist, new_dist).mean().item())
            out[name + '_entropy'] = np.float64(new_dist.entropy().mean().item())
            out[name + '_old_mean'] = old_dist.mean.numpy()
            out[name + '_old_log_std'] = np.float64((old_dist.variance**.5).log().reshape(-1)[0].item())
        # Adam: 3 PPO steps + 3 VF steps on fixed minibatches (lr of ppo.py:65-76)
        idx = np.stack([rng.permutation(n)[:n // 2] for _ in range(3)]).astype(np.int32)
        for mb in idx:
            algo._train_policy(obs[mb], act[mb], rew[mb], adv[mb])
            algo._train_value_function(obs[mb], ret[mb])
        out[name + '_adam_idx'] = idx
        out[name + '_policy_flat_after'] = flat_of(policy.state_dict(), '_module.')
        out[name + '_vf_flat_after'] = flat_of(vf.state_dict(), 'module.')
    np.savez_compressed(os.path.join(OUT, 'losses.npz'), **out)


def gen_train_once():
    """C1 end to end: reference LocalSampler (DefaultWorker) on PointEnv -> PPO._train_once."""
    from garage.envs import PointEnv
    from garage.experiment import deterministic
    from garage.sampler import LocalSampler
    from garage.torch.algos import PPO
    from garage.torch.optimizers import OptimizerWrapper
    from garage.torch.policies import GaussianMLPPolicy
    from garage.torch.value_functions import GaussianMLPValueFunction
    from oracle import ref_shims
    out = {}
    for name, (max_len, goal, nsamp) in dict(fixed=(25, (3., 3.), 500), early=(40, (0.35, 0.3), 600)).items():
        import random
        random.seed(7); np.random.seed(7); torch.manual_seed(7)  # deterministic.set_seed minus the TF call
        env = PointEnv(goal=np.array(goal, dtype=np.float32), max_episode_length=max_len)
        policy = GaussianMLPPolicy(env.spec, hidden_sizes=(32, 32))
        vf = GaussianMLPValueFunction(env.spec, hidden_sizes=(32, 32))
        sampler = LocalSampler(agents=policy, envs=env, max_episode_length=max_len, n_workers=2)
        mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), m,
                                        max_optimization_epochs=3, minibatch_size=64)
        algo = PPO(env.spec, policy, vf, sampler, policy_optimizer=mk(policy), vf_optimizer=mk(vf),
                   gae_lambda=0.95)
        eps = sampler.obtain_samples(0, nsamp, policy.get_param_values())
        out[name + '_cfg'] = np.array([max_len, goal[0], goal[1]], np.float64)
        out[name + '_policy_flat'] = flat_of(policy.state_dict(), '_module.')
        out[name + '_vf_flat'] = flat_of(vf.state_dict(), 'module.')
        out[name + '_obs'], out[name + '_act'] = eps.observations, eps.actions
        out[name + '_rew'], out[name + '_lengths'] = eps.rewards, eps.lengths.astype(np.int64)
        out[name + '_step_types'] = np.asarray(eps.step_types, dtype=np.uint8)
        out[name + '_last_obs'] = eps.last_observations
        out[name + '_success'] = eps.env_infos['success']
        np.random.seed(11)  # the minibatch shuffles of BatchDataset
        ref_shims.tabular.clear()
        ret = algo._train_once(0, eps)
        out[name + '_return'] = np.float64(ret)
        out[name + '_policy_flat_after'] = flat_of(policy.state_dict(), '_module.')
        out[name + '_vf_flat_after'] = flat_of(vf.state_dict(), 'module.')
        keys = sorted(ref_shims.tabular.data)
        out[name + '_log_keys'] = np.array(keys)
        out[name + '_log_vals'] = np.array([float(ref_shims.tabular.data[k]) for k in keys])
    np.savez_compressed(os.path.join(OUT, 'train_once.npz'), **out)


def gen_gridworld():
    """GridWorldEnv + ScriptedPolicy through the reference's own VecWorker / LocalSampler
    (the setting of tests/garage/sampler/test_vec_worker.py): integer observations, actions,
    step types and lengths for the bit-exact GPU parity test.  One case per built-in map; the
    4x4 cases use the action table of the reference test, the chain walks left to the goal, the
    others use seeded random tables."""
    from garage.envs import GridWorldEnv
    from garage.envs.grid_world_env import MAPS
    from garage.np.policies import ScriptedPolicy
    from garage.sampler import LocalSampler, VecWorker, WorkerFactory
    out = {}
    rng = np.random.RandomState(7)
    ref_table = [2, 2, 1, 0, 3, 1, 1, 1, 2, 2, 1, 1, 1, 2, 2, 1]  # test_vec_worker.py:22-31
    cases = [('4x4', ref_table, 16, 3), ('4x4', ref_table, 4, 3), ('8x8', None, 25, 4),
             ('chain', [0] * 29, 40, 2), ('4x4_safe', None, 9, 5), ('8x8', None, 200, 2)]
    for k, (desc, table, max_len, n_envs) in enumerate(cases):
        n_cells = sum(len(r) for r in MAPS[desc])
        if table is None:
            # random tables mostly wander until the time limit: for the odd cases draw (at most 500
            # times) until the walk from the start ends in a hole or at the goal
            for _ in range(500):
                table = [int(v) for v in rng.randint(0, 4, size=n_cells)]
                probe = GridWorldEnv(desc=desc, max_episode_length=max_len)
                obs, _ = probe.reset()
                for _ in range(max_len):
                    es = probe.step(table[obs])
                    obs = es.observation
                    if es.last:
                        break
                if k % 2 == 0 or es.terminal:
                    break
        env = GridWorldEnv(desc=desc, max_episode_length=max_len)
        policy = ScriptedPolicy(scripted_actions=table)
        wf = WorkerFactory(max_episode_length=max_len, n_workers=1, worker_class=VecWorker,
                           worker_args=dict(n_envs=n_envs))
        sampler = LocalSampler.from_worker_factory(wf, policy, env)
        num_samples = 6 * max_len * n_envs // 2 + 1
        eps = sampler.obtain_samples(0, num_samples, None)
        pre = 'c{}_'.format(k)
        out[pre + 'desc'] = np.array(desc)
        out[pre + 'cells'] = np.array([list(r) for r in MAPS[desc]])
        out[pre + 'table'] = np.asarray(table, dtype=np.int32)
        out[pre + 'max_len'] = np.int64(max_len)
        out[pre + 'n_envs'] = np.int64(n_envs)
        out[pre + 'num_samples'] = np.int64(num_samples)
        out[pre + 'observations'] = np.asarray(eps.observations).astype(np.int64).reshape(-1)
        out[pre + 'last_observations'] = np.asarray(eps.last_observations).astype(np.int64).reshape(-1)
        out[pre + 'actions'] = np.asarray(eps.actions).astype(np.int64).reshape(-1)
        out[pre + 'rewards'] = np.asarray(eps.rewards, dtype=np.float64)
        out[pre + 'step_types'] = np.asarray([int(s) for s in eps.step_types], dtype=np.uint8)
        out[pre + 'lengths'] = np.asarray(eps.lengths, dtype=np.int64)
        print('gridworld case', k, desc, 'episodes', len(eps.lengths), 'steps', int(eps.lengths.sum()),
              'obs shape', np.asarray(eps.observations).shape, np.asarray(eps.observations).dtype,
              'returns', float(eps.rewards.sum()))
    out['n_cases'] = np.int64(len(cases))
    np.savez_compressed(os.path.join(OUT, 'gridworld.npz'), **out)


def gen_trpo():
    """reference TRPO (torch/algos/trpo.py) with its default optimisers --
    ConjugateGradientOptimizer(max_constraint_value=0.01) for the policy, Adam minibatches for the
    value function -- on a reference LocalSampler batch of PointEnv.  Records the batch, both
    parameter vectors before / after `_train_once`, the dowel values, and known answers of the
    optimiser's pieces on the same batch: the loss gradient, Hessian-vector products of the KL
    constraint (conjugate_gradient_optimizer.py:18-66) for fixed vectors, the CG solution (:69-110)
    and the accepted step of the line search (:141-260).

    With the default `hvp_reg_coeff=1e-5` the linear system is nearly singular (the Fisher matrix of a
    few hundred samples has a large null space in which A = 1e-5 I), so 10 CG iterations in float32
    amplify rounding noise by ~1e5: the reference's own step changes by tens of per cent with the
    summation order (a float64 replay of the same algorithm lands 0.14 away on a step of 0.13).  The
    cases `fixed` / `early` therefore pin the well-posed PIECES (gradient, Hessian-vector products,
    logged values); the cases `*_reg` run the same reference code with
    `ConjugateGradientOptimizer(hvp_reg_coeff=1.0)`, where CG converges and the whole `_train_once`
    (line search included) is reproducible to ~1e-5."""
    import random

    from garage.envs import PointEnv
    from garage.sampler import LocalSampler
    from garage.torch.algos import TRPO
    from garage.torch.optimizers import conjugate_gradient_optimizer as cgo
    from garage.torch.policies import GaussianMLPPolicy
    from garage.torch.value_functions import GaussianMLPValueFunction
    from oracle import ref_shims
    out = {}
    from garage.torch.optimizers import ConjugateGradientOptimizer, OptimizerWrapper
    cases = dict(fixed=(25, (3., 3.), 500, 32, 1e-5), early=(40, (0.35, 0.3), 600, 64, 1e-5),
                 fixed_reg=(25, (3., 3.), 500, 32, 1.0), early_reg=(40, (0.35, 0.3), 600, 64, 1.0))
    for name, (max_len, goal, nsamp, hid, reg) in cases.items():
        random.seed(9); np.random.seed(9); torch.manual_seed(9)
        env = PointEnv(goal=np.array(goal, dtype=np.float32), max_episode_length=max_len)
        policy = GaussianMLPPolicy(env.spec, hidden_sizes=(hid, hid))
        vf = GaussianMLPValueFunction(env.spec, hidden_sizes=(hid, hid))
        sampler = LocalSampler(agents=policy, envs=env, max_episode_length=max_len, n_workers=2)
        if reg == 1e-5:
            algo = TRPO(env.spec, policy, vf, sampler)  # every default of trpo.py:47-88
        else:
            algo = TRPO(env.spec, policy, vf, sampler, policy_optimizer=OptimizerWrapper(
                (ConjugateGradientOptimizer, dict(max_constraint_value=0.01, hvp_reg_coeff=reg)), policy))
        eps = sampler.obtain_samples(0, nsamp, policy.get_param_values())
        out[name + '_cfg'] = np.array([max_len, goal[0], goal[1], hid, reg], np.float64)
        out[name + '_policy_flat'] = flat_of(policy.state_dict(), '_module.')
        out[name + '_vf_flat'] = flat_of(vf.state_dict(), 'module.')
        out[name + '_obs'], out[name + '_act'] = eps.observations, eps.actions
        out[name + '_rew'], out[name + '_lengths'] = eps.rewards, eps.lengths.astype(np.int64)
        out[name + '_step_types'] = np.asarray(eps.step_types, dtype=np.uint8)
        out[name + '_last_obs'] = eps.last_observations
        out[name + '_success'] = eps.env_infos['success']
        # ---- known answers of the optimiser's pieces at the initial parameters
        obs = torch.from_numpy(np.asarray(eps.observations, np.float32))
        params = [p for p in policy.parameters()]
        rng = np.random.default_rng(4)
        P = sum(p.numel() for p in params)
        f_Ax = cgo._build_hessian_vector_product(lambda: algo._compute_kl_constraint(obs), params, reg)
        vs = rng.standard_normal((3, P)).astype(np.float32)
        out[name + '_hvp_vec'] = vs
        out[name + '_hvp_out'] = np.stack([f_Ax(torch.from_numpy(v)).detach().numpy() for v in vs])
        b = torch.from_numpy((rng.standard_normal(P) * 0.01).astype(np.float32))
        out[name + '_cg_b'] = b.numpy().copy()
        out[name + '_cg_x'] = cgo._conjugate_gradient(f_Ax, b, 10).detach().numpy()
        np.random.seed(13)  # the value function's BatchDataset shuffles
        ref_shims.tabular.clear()
        # record (not alter) what the reference feeds its policy step and the gradient it starts from
        orig_train_policy, rec = algo._train_policy, {}

        def recording_train_policy(obs_, actions_, rewards_, advantages_):
            rec['adv'] = advantages_.detach().numpy().copy()
            rec['n'] = rec.get('n', 0) + 1
            loss_ = orig_train_policy(obs_, actions_, rewards_, advantages_)
            rec['grad'] = flat_grad_of(policy, '_module.')
            return loss_
        algo._train_policy = recording_train_policy
        ret = algo._train_once(0, eps)
        assert rec['n'] == 1  # full batch, one epoch (OptimizerWrapper defaults)
        out[name + '_adv'] = rec['adv']
        out[name + '_loss_grad'] = rec['grad']
        out[name + '_return'] = np.float64(ret)
        out[name + '_policy_flat_after'] = flat_of(policy.state_dict(), '_module.')
        out[name + '_vf_flat_after'] = flat_of(vf.state_dict(), 'module.')
        keys = sorted(ref_shims.tabular.data)
        out[name + '_log_keys'] = np.array(keys)
        out[name + '_log_vals'] = np.array([float(ref_shims.tabular.data[k]) for k in keys])
        d = out[name + '_policy_flat_after'] - out[name + '_policy_flat']
        print('trpo case', name, 'P', P, 'episodes', len(eps.lengths), 'steps', int(eps.lengths.sum()),
              '|step|', float(np.abs(d).max()), 'KL', ref_shims.tabular.data.get('GaussianMLPPolicy/KL'))
    np.savez_compressed(os.path.join(OUT, 'trpo.npz'), **out)


def main():
    from oracle import ref_shims
    if not ref_shims.install():
        sys.exit('/root/reference is not available here')
    os.makedirs(OUT, exist_ok=True)
    gen_gae()
    gen_point_env()
    gen_normalized_env()
    gen_losses()
    gen_train_once()
    gen_gridworld()
    gen_trpo()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == '__main__':
    main()
