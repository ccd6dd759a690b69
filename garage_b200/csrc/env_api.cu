// Stand-alone env reset / single-step kernels (one thread per env): the
// replay-parity entry points.  The rollout kernel (rollout.cu) inlines the same
// device functions from envs.cuh.
#include "envs.cuh"

extern "C" int b2g_env_obs_dim(int env_id) { return env_obs_dim_(env_id); }
extern "C" int b2g_env_act_dim(int env_id) { return env_act_dim_(env_id); }
extern "C" int b2g_env_state_dim(int env_id) { return env_state_dim_(env_id); }

template <int ENV, bool NORM>
__global__ void __launch_bounds__(128)
env_reset_kernel(EnvParams P, NormParams Q, uint32_t seed_lo, uint32_t seed_hi, int64_t gid0,
                 double* __restrict__ state, float* __restrict__ obs,
                 int32_t* __restrict__ step_cnt, int32_t* __restrict__ episode, int64_t N) {
    constexpr int O = EnvDims<ENV>::O, S = EnvDims<ENV>::S;
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    double s[S];
    float o[O];
    const int32_t ep = episode[n];
    env_reset<ENV>(P, seed_lo, seed_hi, (uint32_t)(gid0 + n), (uint32_t)ep, s, o);
    if (NORM) {  // normalized_env.py:71-73
        NormStats<O> ns;
        ns.load(Q, N, n);
        ns.obs(Q, o);
        ns.store(Q, N, n);
    }
#pragma unroll
    for (int i = 0; i < S; ++i) state[(int64_t)i * N + n] = s[i];
#pragma unroll
    for (int i = 0; i < O; ++i) obs[n * O + i] = o[i];
    step_cnt[n] = 0;
    episode[n] = ep + 1;
}

template <int ENV, bool NORM>
__global__ void __launch_bounds__(128)
env_step_kernel(EnvParams P, NormParams Q, double* __restrict__ state, const float* __restrict__ actions,
                float* __restrict__ obs_out, double* __restrict__ reward_out,
                uint8_t* __restrict__ step_type_out, uint8_t* __restrict__ success_out,
                int32_t* __restrict__ step_cnt, int32_t* __restrict__ err_flag, int64_t N) {
    constexpr int O = EnvDims<ENV>::O, A = EnvDims<ENV>::A, S = EnvDims<ENV>::S;
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    int cnt = step_cnt[n];
    if (cnt < 0) {  // point_env.py:113-114: reset() must be called before step()
        if (err_flag) atomicMax(err_flag, (int32_t)(n + 1 > 0x7fffffff ? 0x7fffffff : n + 1));
        return;
    }
    double s[S];
    float a[A], o[O];
#pragma unroll
    for (int i = 0; i < S; ++i) s[i] = state[(int64_t)i * N + n];
#pragma unroll
    for (int i = 0; i < A; ++i) a[i] = actions[n * A + i];
    StepOut r;
    if (NORM) {
        float sa[A];
        norm_action<A>(Q, a, sa);
        r = env_step<ENV>(P, s, sa, o);
        NormStats<O> ns;
        ns.load(Q, N, n);
        ns.obs(Q, o);
        r.reward = ns.reward(Q, r.reward);
        ns.store(Q, N, n);
    } else {
        r = env_step<ENV>(P, s, a, o);
    }
    cnt += 1;
    const uint8_t st = b2g_step_type(cnt, P.max_len, r.done);
#pragma unroll
    for (int i = 0; i < S; ++i) state[(int64_t)i * N + n] = s[i];
#pragma unroll
    for (int i = 0; i < O; ++i) obs_out[n * O + i] = o[i];
    reward_out[n] = r.reward;
    step_type_out[n] = st;
    if (success_out) success_out[n] = r.success ? 1 : 0;
    step_cnt[n] = (st >= ST_TERMINAL) ? -1 : cnt;
}

extern "C" int b2g_env_reset(int env_id, const double* params, uint64_t seed,
                             int64_t env_id_offset, double* state, float* obs,
                             int32_t* step_cnt, int32_t* episode, int64_t N, void* stream) {
    return b2g_env_reset_norm(env_id, params, nullptr, seed, env_id_offset, state, obs, step_cnt, episode, N, stream);
}

extern "C" int b2g_env_reset_norm(int env_id, const double* params, const b2g_norm* norm, uint64_t seed,
                                  int64_t env_id_offset, double* state, float* obs,
                                  int32_t* step_cnt, int32_t* episode, int64_t N, void* stream) {
    EnvParams P;
    NormParams Q;
    B2G_CHECK_ARG(env_params_from_host(env_id, params, 0, &P), "bad env_id or params");
    B2G_CHECK_ARG(norm_from_host(norm, &Q), "normalize_obs / normalize_reward need the stats buffer");
    B2G_CHECK_ARG(N >= 0, "negative N");
    if (N == 0) return B2G_OK;
    B2G_CHECK_ARG(state && obs && step_cnt && episode, "NULL pointer");
    const unsigned blocks = (unsigned)((N + 127) / 128);
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t lo = (uint32_t)seed, hi = (uint32_t)(seed >> 32);
#define B2G_RESET(E)                                                                                                  \
    if (Q.flags & B2G_NORM_OBS) env_reset_kernel<E, true><<<blocks, 128, 0, s>>>(P, Q, lo, hi, env_id_offset, state, obs, step_cnt, episode, N); \
    else env_reset_kernel<E, false><<<blocks, 128, 0, s>>>(P, Q, lo, hi, env_id_offset, state, obs, step_cnt, episode, N);
    switch (env_id) {
        case B2G_ENV_POINT: B2G_RESET(B2G_ENV_POINT) break;
        case B2G_ENV_PENDULUM: B2G_RESET(B2G_ENV_PENDULUM) break;
        default: B2G_RESET(B2G_ENV_CARTPOLE) break;
    }
#undef B2G_RESET
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

extern "C" int b2g_env_step(int env_id, const double* params, int max_episode_length,
                            double* state, const float* actions, float* obs_out,
                            double* reward_out, uint8_t* step_type_out, uint8_t* success_out,
                            int32_t* step_cnt, int32_t* err_flag, int64_t N, void* stream) {
    return b2g_env_step_norm(env_id, params, nullptr, max_episode_length, state, actions, obs_out, reward_out,
                             step_type_out, success_out, step_cnt, err_flag, N, stream);
}

extern "C" int b2g_env_step_norm(int env_id, const double* params, const b2g_norm* norm, int max_episode_length,
                                 double* state, const float* actions, float* obs_out,
                                 double* reward_out, uint8_t* step_type_out, uint8_t* success_out,
                                 int32_t* step_cnt, int32_t* err_flag, int64_t N, void* stream) {
    EnvParams P;
    NormParams Q;
    B2G_CHECK_ARG(env_params_from_host(env_id, params, max_episode_length, &P), "bad env_id or params");
    B2G_CHECK_ARG(norm_from_host(norm, &Q), "normalize_obs / normalize_reward need the stats buffer");
    B2G_CHECK_ARG(N >= 0, "negative N");
    if (N == 0) return B2G_OK;
    B2G_CHECK_ARG(state && actions && obs_out && reward_out && step_type_out && step_cnt, "NULL pointer");
    const unsigned blocks = (unsigned)((N + 127) / 128);
    cudaStream_t s = (cudaStream_t)stream;
#define B2G_STEP(E)                                                                                                  \
    if (norm_active(Q)) env_step_kernel<E, true><<<blocks, 128, 0, s>>>(P, Q, state, actions, obs_out, reward_out, step_type_out, success_out, step_cnt, err_flag, N); \
    else env_step_kernel<E, false><<<blocks, 128, 0, s>>>(P, Q, state, actions, obs_out, reward_out, step_type_out, success_out, step_cnt, err_flag, N);
    switch (env_id) {
        case B2G_ENV_POINT: B2G_STEP(B2G_ENV_POINT) break;
        case B2G_ENV_PENDULUM: B2G_STEP(B2G_ENV_PENDULUM) break;
        default: B2G_STEP(B2G_ENV_CARTPOLE) break;
    }
#undef B2G_STEP
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
