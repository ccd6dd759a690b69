"""`GymEnv('<id>')` for launcher compatibility (reference envs/gym_env.py:80-141).

gym itself is absent here and the accelerated path needs the dynamics on the device anyway, so
the ids whose dynamics exist as kernels (csrc/envs.cuh) map to the batched environments with
gym's registered episode limits; any other id fails loudly instead of silently running
something else."""
from garage_b200.envs.batched import CartPoleEnv, PendulumEnv

# id -> (class, gym 0.17.2 registry max_episode_steps)
_REGISTRY = {
    'Pendulum-v0': (PendulumEnv, 200),
    'CartPole-v0': (CartPoleEnv, 200),
    'CartPole-v1': (CartPoleEnv, 500),
}


def GymEnv(env, is_image=False, max_episode_length=None):
    """Returns the batched GPU environment for a gym id.

    Args:
        env (str): gym id; only ids in the registry above are available.
        is_image (bool): must be False (no pixel environments on this path).
        max_episode_length (int): overrides gym's registered limit (gym_env.py:116-141).
    """
    if not isinstance(env, str):
        raise ValueError('GymEnv on the GPU path takes a gym id string, not a constructed '
                         'gym.Env: the dynamics run as CUDA kernels')
    if is_image:
        raise NotImplementedError('image observations are not part of the accelerated path')
    if env not in _REGISTRY:
        raise NotImplementedError(
            'no device kernel for gym id {!r}; available: {}'.format(env, sorted(_REGISTRY)))
    cls, limit = _REGISTRY[env]
    return cls(max_episode_length=max_episode_length or limit)
