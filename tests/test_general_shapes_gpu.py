"""GPU parity of the general-shape MLP kernels (csrc/mlp_general.cuh: any obs_dim <= 8, out_dim <= 4,
hidden in {32, 64, 128} -- what the reference accepts through hidden_sizes / the env spec,
torch/policies/gaussian_mlp_policy.py:53-87, torch/modules/multi_headed_mlp_module.py:49-106) against
the float64 oracle: loss, every gradient group, forward means / KL / entropy, Fisher-vector products,
and a whole PPO `_train_once` on a host EpisodeBatch of a 6-d observation / 3-d action task compared
with the oracle port of the reference's `_train_once`."""
import math

import numpy as np
import pytest
import torch

from oracle import np_oracle as O
from tests.test_mlp_gpu import check_grads, make_case, rel_err

pytestmark = pytest.mark.gpu
RTOL = 1e-5
# (obs_dim, hidden, out_dim): new shapes + the old ones forced through the general kernels
NEW = [(6, 128, 3), (8, 32, 4), (5, 64, 2), (1, 64, 1), (7, 128, 1), (2, 32, 3), (8, 128, 4), (6, 64, 3)]
OLD = [(3, 64, 1), (3, 32, 2), (4, 64, 1)]


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


class general_path:
    """Force every shape through the general kernels (path 3)."""

    def __enter__(self):
        from garage_b200 import ops
        self.prev = ops.mlp_path()
        ops.set_mlp_path(3)

    def __exit__(self, *a):
        from garage_b200 import ops
        ops.set_mlp_path(self.prev)


@pytest.mark.parametrize('Od,H,A', NEW + OLD)
@pytest.mark.parametrize('n', [1, 77, 1000, 20001])
@pytest.mark.parametrize('kind', ['ppo', 'vpg'])
def test_policy_loss_and_grad_general(Od, H, A, n, kind):
    from garage_b200 import ops
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, n, seed=n + H + A + Od)
    want_loss, want_g = O.policy_loss_and_grad(flat, Od, H, A, obs, act, adv, old_logp, kind=kind)
    with general_path():
        sh = ops.MlpShape(Od, H, A, torch.device('cuda:0'))
        nb = ops.mlp_minibatch_grad(sh, _t(flat), _t(obs), act=_t(act), adv=_t(adv), old_logp=_t(old_logp),
                                    loss_kind=ops.LOSS_PPO if kind == 'ppo' else ops.LOSS_VPG,
                                    min_log_std=math.log(1e-6))
        grad = torch.empty(sh.P, device='cuda')
        stats = torch.empty(8, dtype=torch.float64, device='cuda')
        ops.mlp_reduce(sh, nb, 1.0 / n, grad_out=grad, stats_out=stats)
    stats = stats.cpu().numpy()
    assert stats[4] == n
    assert abs(stats[0] / n - want_loss) <= RTOL * max(1.0, abs(want_loss))
    tol = RTOL if n < 20000 else 3e-5  # bias groups of 20,001 zero-mean rows: see test_bench_shape_gpu.py
    check_grads(grad.cpu().numpy().astype(np.float64), want_g, Od, H, A, tol=tol)


@pytest.mark.parametrize('Od,H', [(6, 128), (8, 32), (5, 64), (3, 64)])
@pytest.mark.parametrize('n', [5, 300, 4096])
def test_vf_loss_and_grad_general(Od, H, n):
    from garage_b200 import ops
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, 1, n, seed=7 * n + H)
    want_loss, want_g = O.vf_loss_and_grad(flat, Od, H, obs, ret)
    with general_path():
        sh = ops.MlpShape(Od, H, 1, torch.device('cuda:0'))
        nb = ops.mlp_minibatch_grad(sh, _t(flat), _t(obs), ret=_t(ret), loss_kind=ops.LOSS_VF)
        grad = torch.empty(sh.P, device='cuda')
        stats = torch.empty(8, dtype=torch.float64, device='cuda')
        ops.mlp_reduce(sh, nb, 1.0 / n, grad_out=grad, stats_out=stats)
    assert abs(stats[0].item() / n - want_loss) <= RTOL * max(1.0, abs(want_loss))
    check_grads(grad.cpu().numpy().astype(np.float64), want_g, Od, H, 1)


@pytest.mark.parametrize('Od,H,A', [(6, 128, 3), (8, 32, 4), (3, 64, 2)])
def test_forward_means_kl_entropy_general(Od, H, A):
    from garage_b200 import ops
    n = 3001
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, n, seed=3)
    p = O.unflatten_params(flat.astype(np.float64), Od, H, A)
    mean, _, _ = O.mlp_forward(p, obs)
    ls, _ = O.clamped_log_std(p)
    old_mean = (mean + np.random.default_rng(1).standard_normal(mean.shape) * 0.1).astype(np.float32)
    idx = np.random.default_rng(2).permutation(n).astype(np.int32)
    with general_path():
        sh = ops.MlpShape(Od, H, A, torch.device('cuda:0'))
        mo = torch.empty(n, A, device='cuda')
        nb = ops.mlp_eval(sh, _t(flat), _t(obs), idx=_t(idx), act=_t(act), adv=_t(adv), old_logp=_t(old_logp),
                          old_mean=_t(old_mean), old_log_std=torch.tensor([ls - 0.05], dtype=torch.float32, device='cuda'),
                          loss_kind=ops.LOSS_PPO, min_log_std=math.log(1e-6), mean_out=mo)
        st = torch.empty(8, dtype=torch.float64, device='cuda')
        ops.mlp_reduce(sh, nb, 1.0, stats_out=st)
    st = st.cpu().numpy()
    assert np.allclose(mo.cpu().numpy(), mean[idx], rtol=1e-5, atol=1e-5)
    want_kl = O.kl_normal(old_mean.astype(np.float64), ls - 0.05, mean, ls).mean()
    assert abs(st[1] / n - want_kl) <= 1e-5 * max(1, abs(want_kl))
    assert abs(st[2] / n - O.normal_entropy(ls, A)) <= 1e-5
    want_loss, _ = O.policy_loss_and_grad(flat, Od, H, A, obs, act, adv, old_logp)
    assert abs(st[0] / n - want_loss) <= 1e-5 * max(1, abs(want_loss))


@pytest.mark.parametrize('Od,H,A', NEW[:5] + OLD)
@pytest.mark.parametrize('n', [33, 5000])
def test_fisher_vector_product_general(Od, H, A, n):
    from garage_b200 import ops
    flat, obs, act, adv, old_logp, ret = make_case(Od, H, A, n, seed=n + Od)
    rng = np.random.default_rng(Od + H)
    sh = ops.MlpShape(Od, H, A, torch.device('cuda:0'))
    for reg in (1e-5, 0.3):
        v = rng.standard_normal(sh.P).astype(np.float32)
        want = O.fisher_vector_product(flat, Od, H, A, obs, v, reg)
        got = ops.mlp_fvp(sh, _t(flat), _t(v), _t(obs), min_log_std=math.log(1e-6), reg_coeff=reg)
        g, w = O.unflatten_params(got.cpu().numpy().astype(np.float64), Od, H, A), O.unflatten_params(want, Od, H, A)
        for k in w:
            assert rel_err(g[k], w[k]) < 1e-5, (k, rel_err(g[k], w[k]))


def test_clamped_log_std_blocks_fvp_entry():
    from garage_b200 import ops
    Od, H, A, n = 4, 32, 2, 200
    flat, obs, *_ = make_case(Od, H, A, n, seed=1)
    flat[0] = -20.0  # below log(1e-6): the clamp holds, no gradient reaches log_std
    v = np.random.default_rng(0).standard_normal(flat.size).astype(np.float32)
    sh = ops.MlpShape(Od, H, A, torch.device('cuda:0'))
    got = ops.mlp_fvp(sh, _t(flat), _t(v), _t(obs), min_log_std=math.log(1e-6), reg_coeff=0.0).cpu().numpy()
    want = O.fisher_vector_product(flat, Od, H, A, obs, v, 0.0)
    assert got[0] == 0.0 and want[0] == 0.0
    assert rel_err(got.astype(np.float64), want) < 1e-5


def test_unsupported_shapes_fail_loudly():
    from garage_b200 import ops
    for shape in [(9, 64, 1), (3, 64, 5), (3, 48, 1), (3, 256, 1)]:
        with pytest.raises(NotImplementedError):
            ops.MlpShape(*shape, torch.device('cuda:0'))


@pytest.mark.parametrize('H', [128, 32])
def test_train_once_6d_obs_3d_act_vs_oracle_port(H):
    """`GaussianMLPPolicy(hidden_sizes=(128, 128))` on a 6-d observation / 3-d action task: a host
    EpisodeBatch (as the reference's sampler would hand over) through PPO._train_once with the
    reference's numpy shuffles vs the oracle port of the reference's `_train_once` on the same batch."""
    from garage_b200 import EnvSpec, EpisodeBatch
    from garage_b200._dtypes import step_types_from_codes
    from garage_b200.spaces import Box
    from garage_b200.torch.algos import PPO
    from garage_b200.torch.optimizers import OptimizerWrapper
    from garage_b200.torch.policies import GaussianMLPPolicy
    from garage_b200.torch.value_functions import GaussianMLPValueFunction
    from oracle import port
    rng = np.random.default_rng(H)
    Od, A, max_len = 6, 3, 30
    lengths = np.array([30, 12, 30, 7, 30, 21, 30, 30, 18, 30], np.int64)
    n = int(lengths.sum())
    spec = EnvSpec(observation_space=Box(low=-np.inf, high=np.inf, shape=(Od, ), dtype=np.float32),
                   action_space=Box(low=-1.0, high=1.0, shape=(A, ), dtype=np.float32), max_episode_length=max_len)
    torch.manual_seed(H)
    policy = GaussianMLPPolicy(spec, hidden_sizes=(H, H))
    vf = GaussianMLPValueFunction(spec, hidden_sizes=(H, H))
    obs = rng.standard_normal((n, Od)).astype(np.float32)
    rew = rng.standard_normal(n).astype(np.float64)
    pol = port.GaussianMLP(Od, A, (H, H))
    vfp = port.GaussianMLP(Od, 1, (H, H), min_std=None)
    pol.load_flat(policy.flat.cpu().numpy())
    vfp.load_flat(vf.flat.cpu().numpy())
    with torch.no_grad():  # actions drawn from the policy itself, so that the likelihood ratio starts at 1
        dist = pol(torch.from_numpy(obs))
        act = dist.sample().numpy().astype(np.float32)
    st = []
    for L in lengths:
        st += [O.FIRST] + [O.MID] * (L - 2) + [O.TIMEOUT if L == max_len else O.TERMINAL]
    eps = EpisodeBatch(env_spec=spec, episode_infos={}, observations=obs, last_observations=obs[np.cumsum(lengths) - 1],
                       actions=act, rewards=rew, env_infos={}, agent_infos={}, step_types=step_types_from_codes(np.array(st, np.uint8)),
                       lengths=lengths)
    mk = lambda m: OptimizerWrapper((torch.optim.Adam, dict(lr=2.5e-4)), m, max_optimization_epochs=3, minibatch_size=64)
    algo = PPO(spec, policy, vf, None, policy_optimizer=mk(policy), vf_optimizer=mk(vf), gae_lambda=0.95,
               shuffle='numpy', pad_value_mode='reference')
    np.random.seed(21)
    ret = algo._train_once(0, eps)
    palgo = port.PPOPort(pol, vfp, max_len, kind='ppo', epochs=3, minibatch_size=64, gae_lambda=0.95)
    np.random.seed(21)
    out = palgo.train_once(dict(observations=obs, actions=act, rewards=rew, lengths=lengths))
    assert abs(ret - out['average_return']) <= 1e-6 * max(1.0, abs(ret))  # rewards travel as float32
    assert np.allclose(policy.flat.cpu().numpy(), pol.flat(), rtol=1e-5, atol=2e-6)
    assert np.allclose(vf.flat.cpu().numpy(), vfp.flat(), rtol=1e-5, atol=2e-6)
