"""Known-answer tests for the two gym==0.17.2 step functions that are NOT in /root/reference
(Pendulum-v0, CartPole-v1; SURVEY 8c: "parity unpinned").  gym cannot be run on either box, so
these pin what CAN be pinned without it: closed-form consequences of the published step() code,
worked out by hand below (the expected numbers are literals with their derivation, not calls into
the oracle), at the edge cases where a transcription usually goes wrong:

  Pendulum  (g=10, m=l=1, dt=.05, max_speed=8, max_torque=2)
    thdot' = thdot + (-3g/(2l) sin(th+pi) + 3/(m l^2) u) dt ;  th' = th + thdot' dt  (UNCLIPPED thdot')
    thdot' clipped to +-8 afterwards;  cost = angle_normalize(th)^2 + .1 thdot^2 + .001 u^2 with the
    CLIPPED torque and the OLD state;  angle_normalize(x) = ((x+pi) mod 2pi) - pi, python modulo.
  CartPole  (gravity 9.8, masscart 1, masspole .1, length .5, force 10, tau .02, euler)
    positions advance with the OLD velocities; done is evaluated on the NEW state with strict
    inequalities at 2.4 and 12 degrees.

The same table is checked against the numpy oracle here (CPU) and against the CUDA step kernel
(`b2g_env_step`, csrc/envs.cuh) in the `gpu`-marked test.
"""
import math

import numpy as np
import pytest

from oracle import np_oracle as O

PI = math.pi

# (theta, theta_dot, action) -> (theta', theta_dot', reward)
PENDULUM = [
    # upright, at rest, no torque: sin(pi) = 1.2246e-16 in float64 -> thdot' = -15 * 1.2246e-16 * .05
    ((0.0, 0.0, 0.0), (-9.184850993605148e-17 * 0.05, -9.184850993605148e-17, 0.0)),
    # horizontal: sin(3pi/2) = -1 -> thdot' = 15 * .05 = .75 ; th' = pi/2 + .75 * .05 ; cost (pi/2)^2
    ((PI / 2, 0.0, 0.0), (PI / 2 + 0.0375, 0.75, -(PI / 2)**2)),
    # torque 5 is clipped to 2: thdot' = 3 * 2 * .05 = .3 (+ the sin(pi) residue) ; cost .001 * 4
    ((0.0, 0.0, 5.0), (0.3 * 0.05, 0.3, -0.004)),
    ((0.0, 0.0, -5.0), (-0.3 * 0.05, -0.3, -0.004)),
    # speed clip ORDER: thdot' = 7.9 + .3 = 8.2 moves the angle (th' = .41), THEN it is clipped to 8
    ((0.0, 7.9, 2.0), (8.2 * 0.05, 8.0, -(0.1 * 7.9**2 + 0.004))),
    ((0.0, -7.9, -2.0), (-8.2 * 0.05, -8.0, -(0.1 * 7.9**2 + 0.004))),
    # angle_normalize: pi -> -pi, -pi -> -pi (python modulo), 3pi/2 -> -pi/2, -3pi/2 -> +pi/2, 2pi+.5 -> .5
    ((PI, 0.0, 0.0), (None, None, -PI**2)),
    ((-PI, 0.0, 0.0), (None, None, -PI**2)),
    ((3 * PI / 2, 0.0, 0.0), (None, None, -(PI / 2)**2)),
    ((-3 * PI / 2, 0.0, 0.0), (None, None, -(PI / 2)**2)),
    ((2 * PI + 0.5, 1.0, 0.0), (None, None, -(0.25 + 0.1))),
    ((-7 * PI, 0.0, 1.0), (None, None, -(PI**2 + 0.001))),
]

THETA_LIM = 12 * 2 * PI / 360  # 0.20943951023931953
# (x, x_dot, theta, theta_dot, action) -> (x', x_dot', theta', theta_dot', done)
# From rest at the origin with a push to the right (force 10):
#   temp = 10 / 1.1 = 100/11 ; thetaacc = -(100/11) / (.5 (4/3 - .1/1.1)) = -(100/11)(66/41) = -600/41
#   xacc = 100/11 + .05 (600/41) / 1.1 = 4400/451 = 400/41
#   x_dot' = .02 * 400/41 = 8/41 ; theta_dot' = -.02 * 600/41 = -12/41 ; x, theta unchanged (old velocities 0)
CARTPOLE = [
    ((0.0, 0.0, 0.0, 0.0, 1.0), (0.0, 8 / 41, 0.0, -12 / 41, False)),
    ((0.0, 0.0, 0.0, 0.0, -0.3), (0.0, -8 / 41, 0.0, 12 / 41, False)),   # any a <= 0 pushes left
    ((0.0, 0.0, 0.0, 0.0, 0.0), (0.0, -8 / 41, 0.0, 12 / 41, False)),    # a == 0 is gym's action 0
    # euler order: positions advance with the OLD velocities
    ((0.0, 1.0, 0.0, 0.5, 1.0), (0.02, None, 0.01, None, False)),
    # thresholds, strict, on the NEW state: exactly 2.4 / 12 degrees is still alive, a hair beyond is done
    ((2.4, 0.0, 0.0, 0.0, 1.0), (2.4, 8 / 41, 0.0, -12 / 41, False)),
    ((2.4, 1e-6, 0.0, 0.0, 1.0), (2.4 + 2e-8, None, 0.0, None, True)),
    ((-2.4, -1e-6, 0.0, 0.0, 1.0), (-2.4 - 2e-8, None, 0.0, None, True)),
    ((0.0, 0.0, THETA_LIM, 0.0, 1.0), (0.0, None, THETA_LIM, None, False)),
    ((0.0, 0.0, THETA_LIM, 1e-6, 1.0), (0.0, None, THETA_LIM + 2e-8, None, True)),
    ((0.0, 0.0, -THETA_LIM, -1e-6, 1.0), (0.0, None, -THETA_LIM - 2e-8, None, True)),
]


def _check_pendulum(step):
    """`step(state [n,2] f64, action [n,1] f32) -> (new_state, obs f32, reward f64)`."""
    st = np.array([c[0][:2] for c in PENDULUM], np.float64)
    ac = np.array([[c[0][2]] for c in PENDULUM], np.float32)
    ns, obs, rew = step(st, ac)
    for i, (_, (th, thd, r)) in enumerate(PENDULUM):
        assert rew[i] == pytest.approx(r, rel=1e-12, abs=1e-15), ('reward', i)
        if th is not None:
            assert ns[i, 0] == pytest.approx(th, rel=1e-12, abs=1e-17), ('theta', i)
            assert ns[i, 1] == pytest.approx(thd, rel=1e-12, abs=1e-17), ('theta_dot', i)
        assert np.allclose(obs[i], [math.cos(ns[i, 0]), math.sin(ns[i, 0]), ns[i, 1]], rtol=1e-6, atol=1e-7)


def _check_cartpole(step):
    """`step(state [n,4] f64, action [n,1] f32) -> (new_state, obs f32, done)`."""
    st = np.array([c[0][:4] for c in CARTPOLE], np.float64)
    ac = np.array([[c[0][4]] for c in CARTPOLE], np.float32)
    ns, obs, done = step(st, ac)
    for i, (_, want) in enumerate(CARTPOLE):
        for j in range(4):
            if want[j] is not None:
                assert ns[i, j] == pytest.approx(want[j], rel=1e-12, abs=1e-15), (i, j)
        assert bool(done[i]) == want[4], ('done', i)
        assert np.array_equal(obs[i], ns[i].astype(np.float32))


def test_pendulum_kat_oracle():
    def step(st, ac):
        ns, obs, rew, _ = O.pendulum_step(st, ac)
        return ns, obs, rew
    _check_pendulum(step)


def test_cartpole_kat_oracle():
    def step(st, ac):
        ns, obs, _, done = O.cartpole_step(st, ac)
        return ns, obs, done
    _check_cartpole(step)


def _gpu_step(env, st, ac, max_len=1000):
    import torch
    from garage_b200 import ops
    dev = torch.device('cuda:0')
    n = st.shape[0]
    es = ops.EnvState(env.env_id, n, dev)
    es.state.copy_(torch.from_numpy(np.ascontiguousarray(st.T)))
    es.step_cnt.zero_()
    obs, rew, stype, _ = ops.env_step(es, env.params_array(), max_len, torch.from_numpy(ac).to(dev))
    return es.state.cpu().numpy().T.copy(), obs.cpu().numpy(), rew.cpu().numpy(), stype.cpu().numpy()


@pytest.mark.gpu
def test_pendulum_kat_kernel():
    from garage_b200.envs import PendulumEnv
    env = PendulumEnv(max_episode_length=1000)

    def step(st, ac):
        ns, obs, rew, stype = _gpu_step(env, st, ac)
        assert (stype == O.FIRST).all()
        return ns, obs, rew
    _check_pendulum(step)


@pytest.mark.gpu
def test_cartpole_kat_kernel():
    from garage_b200.envs import CartPoleEnv
    env = CartPoleEnv(max_episode_length=1000)

    def step(st, ac):
        ns, obs, rew, stype = _gpu_step(env, st, ac)
        assert (rew == 1.0).all()
        done = stype == O.TERMINAL
        assert ((stype == O.TERMINAL) | (stype == O.FIRST)).all()
        return ns, obs, done
    _check_cartpole(step)
