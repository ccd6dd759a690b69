// Persistent rollout kernel: policy MLP forward + Gaussian sampling + log-prob +
// env step + StepType / episode bookkeeping + auto-reset, fused over n_steps.
//   replaces the per-step Python loop of garage sampler/vec_worker.py:176-204
//   (agent.get_actions -> torch/policies/stochastic_policy.py:46-89 ->
//    torch/modules/gaussian_mlp_module.py:158-192, 288-305; env.step per env;
//    _gather_episode :139-174 with the immediate env.reset() at :171).
//
// Mapping: one WARP owns G environments for the whole rollout.  Each lane owns
// H/32 hidden units: its rows of W1, W2 and columns of W3 live in REGISTERS for
// the whole kernel (loaded once, coalesced through shared memory), activations
// are exchanged through a per-warp shared-memory tile and re-read as broadcast
// float4 (G LDS.128 feed 4*G*H/32 FMAs).  Lanes 0..G-1 additionally own one env
// each: state (float64) in registers, Philox4x32-10 counter RNG keyed by
// (seed; global env id, global step), outputs written time-major [T,N].
// WS form (hidden 128, where W2's rows no longer fit in registers): W2 stays resident in shared memory
// (row stride H + 4 floats: the lanes' LDS.128 of one column group fall in distinct banks) and is
// re-read per step next to the broadcast activations: UPL + G LDS.128 feed 4 G UPL FMAs.
#include "envs.cuh"
#include "mlp.cuh"
#include <cstdlib>

// tanh x = 1 - 2 / (1 + e^{2x}) on the SFU (ex2 / rcp), absolute error ~2e-7: the same activation
// the tensor-core update kernels evaluate (mlp_tc.cuh), a quarter of tanhf's instructions.  The
// rollout is serial over time with ~8 warps per SM, so instruction count is what it pays for.
__device__ __forceinline__ float ro_tanh(float x) {
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * 2.8853900817779268f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.0f));
    return fmaf(-2.0f, r, 1.0f);
}

struct RolloutArgs {
    EnvParams P;
    NormParams Q;        // normalize(env) wrapper (envs.cuh); used by the NORM instantiations only
    uint32_t seed_lo, seed_hi;
    int64_t gid0;
    uint32_t t_counter;  // global step counter of row 0 (Philox counter word)
    const unsigned long long* t_counter_dev;  // optional device-side addend (CUDA-graph replays)
    const float* policy;
    float min_log_std, max_log_std;
    double* state;       // [S,N]
    float* cur_obs;      // [N,O]
    int32_t* step_cnt;   // [N]
    int32_t* episode;    // [N]
    int n_steps;
    int64_t N;
    float* obs;          // [T,N,O]
    float* act;          // [T,N,A]
    float* rew;          // [T,N]
    float* logp;         // [T,N]
    uint8_t* step_type;  // [T,N]
    float* mean;         // [T,N,A] or NULL
    uint8_t* success;    // [T,N] or NULL
    float* last_obs;     // [T,N,O] or NULL, written at TERMINAL/TIMEOUT rows only
    int32_t* ep_len;     // [T,N] or NULL, episode length, written at TERMINAL/TIMEOUT rows only
};

#ifndef B2G_RO_WARPS
#define B2G_RO_WARPS 4
#endif
constexpr int RO_WARPS = B2G_RO_WARPS;  // warps per CTA (one warp owns G envs)

template <int ENV, int H, int G, bool NORM, bool WS = false>
__global__ void __launch_bounds__(RO_WARPS * 32)
rollout_kernel(const RolloutArgs a) {
    constexpr int O = EnvDims<ENV>::O, A = EnvDims<ENV>::A, S = EnvDims<ENV>::S;
    constexpr int UPL = H / 32;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t warp = (int64_t)blockIdx.x * RO_WARPS + wib;
    const int64_t e0 = warp * G;
    const int64_t N = a.N;
    const MlpLayout L{O, H, A};

    __shared__ __align__(16) float sh_h[RO_WARPS][G][H];
    // padded staging: coalesced global read, conflict-free rows (WS: resident for the whole kernel)
    constexpr int WSTR = WS ? H + 4 : H + 1;
    __shared__ float sh_w_static[WS ? 1 : H * (H + 1)];
    extern __shared__ __align__(16) float sh_w_dyn[];
    float* const sh_w = WS ? sh_w_dyn : sh_w_static;

    // ---- weights -> registers (W2 through smem so that global reads coalesce)
    float w1[UPL][O], b1[UPL], w2[WS ? 1 : UPL][WS ? 1 : H], b2[UPL], w3[A][UPL], b3[A];
    for (int i = threadIdx.x; i < H * H; i += blockDim.x)
        sh_w[(i / H) * WSTR + (i % H)] = a.policy[L.off_w2() + i];
    __syncthreads();
    if constexpr (!WS) {
#pragma unroll
        for (int u = 0; u < UPL; ++u) {
            const int j = lane + 32 * u;
#pragma unroll
            for (int k = 0; k < H; ++k) w2[u][k] = sh_w[j * WSTR + k];
        }
    }
#pragma unroll
    for (int u = 0; u < UPL; ++u) {
        const int j = lane + 32 * u;
#pragma unroll
        for (int i = 0; i < O; ++i) w1[u][i] = a.policy[L.off_w1() + j * O + i];
        b1[u] = a.policy[L.off_b1() + j];
        b2[u] = a.policy[L.off_b2() + j];
#pragma unroll
        for (int q = 0; q < A; ++q) w3[q][u] = a.policy[L.off_w3() + q * H + j];
    }
#pragma unroll
    for (int q = 0; q < A; ++q) b3[q] = a.policy[L.off_b3() + q];
    // gaussian_mlp_module.py:171-176: clamp(log_std, min, max); std = exp(log_std)
    const float log_std = fminf(fmaxf(a.policy[L.off_logstd()], a.min_log_std), a.max_log_std);
    const float stdv = expf(log_std);
    const float var = stdv * stdv;
    const float log_scale = logf(stdv);  // Normal.log_prob uses scale.log()

    if (e0 >= N) return;  // whole warp idle (after the block-wide weight staging)

    // ---- per-env owner state
    const bool owner = lane < G && (e0 + lane) < N;
    const int64_t n = e0 + (lane < G ? lane : 0);
    double s[S];
    float ob[O];
    int cnt = 0, ep = 0;
    NormStats<NORM ? O : 1> ns;
    if (NORM && owner) ns.load(a.Q, N, n);
    if (owner) {
#pragma unroll
        for (int i = 0; i < S; ++i) s[i] = a.state[(int64_t)i * N + n];
#pragma unroll
        for (int i = 0; i < O; ++i) ob[i] = a.cur_obs[n * O + i];
        cnt = a.step_cnt[n];
        ep = a.episode[n];
    } else {
#pragma unroll
        for (int i = 0; i < S; ++i) s[i] = 0.0;
#pragma unroll
        for (int i = 0; i < O; ++i) ob[i] = 0.0f;
    }
    const uint32_t gid = (uint32_t)(a.gid0 + n);
    const uint32_t t_base = a.t_counter + (a.t_counter_dev ? (uint32_t)*a.t_counter_dev : 0u);

    for (int t = 0; t < a.n_steps; ++t) {
        // ---- layer 1: h1 = tanh(W1 x + b1), x broadcast from the owner lanes
        float x[G][O];
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
            for (int i = 0; i < O; ++i) x[g][i] = __shfl_sync(0xffffffffu, ob[i], g);
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
            for (int u = 0; u < UPL; ++u) {
                float z = b1[u];
#pragma unroll
                for (int i = 0; i < O; ++i) z = fmaf(w1[u][i], x[g][i], z);
                sh_h[wib][g][lane + 32 * u] = ro_tanh(z);
            }
        __syncwarp();
        // ---- layer 2: h2 = tanh(W2 h1 + b2); h1 re-read as broadcast float4
        float acc[G][UPL];
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
            for (int u = 0; u < UPL; ++u) acc[g][u] = b2[u];
#pragma unroll(WS ? 4 : H / 4)
        for (int k4 = 0; k4 < H / 4; ++k4) {
            float4 wv[UPL];
#pragma unroll
            for (int u = 0; u < UPL; ++u) {
                if constexpr (WS) wv[u] = *reinterpret_cast<const float4*>(&sh_w[(lane + 32 * u) * WSTR + 4 * k4]);
                else wv[u] = make_float4(w2[u][4 * k4 + 0], w2[u][4 * k4 + 1], w2[u][4 * k4 + 2], w2[u][4 * k4 + 3]);
            }
#pragma unroll
            for (int g = 0; g < G; ++g) {
                const float4 hv = *reinterpret_cast<const float4*>(&sh_h[wib][g][4 * k4]);
#pragma unroll
                for (int u = 0; u < UPL; ++u) {
                    acc[g][u] = fmaf(wv[u].x, hv.x, acc[g][u]);
                    acc[g][u] = fmaf(wv[u].y, hv.y, acc[g][u]);
                    acc[g][u] = fmaf(wv[u].z, hv.z, acc[g][u]);
                    acc[g][u] = fmaf(wv[u].w, hv.w, acc[g][u]);
                }
            }
        }
        __syncwarp();
        // ---- layer 3: mean = W3 h2 + b3 (warp all-reduce of per-lane partials)
        float m[G][A];
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
            for (int q = 0; q < A; ++q) {
                float p = 0.0f;
#pragma unroll
                for (int u = 0; u < UPL; ++u) p = fmaf(w3[q][u], ro_tanh(acc[g][u]), p);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) p += __shfl_xor_sync(0xffffffffu, p, o);
                m[g][q] = p + b3[q];
            }
        // ---- owners: sample, log-prob, env step, bookkeeping
        if (owner) {
            float mu[A];
#pragma unroll
            for (int q = 0; q < A; ++q) mu[q] = m[0][q];
#pragma unroll
            for (int g = 1; g < G; ++g)
                if (lane == g) {
#pragma unroll
                    for (int q = 0; q < A; ++q) mu[q] = m[g][q];
                }
            float eps[A], act[A];
            action_noise<A>(a.seed_lo, a.seed_hi, gid, t_base + (uint32_t)t, eps);
            float lp = 0.0f;
#pragma unroll
            for (int q = 0; q < A; ++q) {
                act[q] = fmaf(stdv, eps[q], mu[q]);
                lp += normal_logp_term(act[q], mu[q], var, log_scale);
            }
            const int64_t row = (int64_t)t * N + n;
#pragma unroll
            for (int i = 0; i < O; ++i) a.obs[row * O + i] = ob[i];
#pragma unroll
            for (int q = 0; q < A; ++q) a.act[row * A + q] = act[q];
            if (a.mean) {
#pragma unroll
                for (int q = 0; q < A; ++q) a.mean[row * A + q] = mu[q];
            }
            a.logp[row] = lp;
            StepOut r;
            if constexpr (NORM) {  // normalized_env.py:75-115
                float sa[A];
                norm_action<A>(a.Q, act, sa);
                r = env_step<ENV>(a.P, s, sa, ob);
                ns.obs(a.Q, ob);
                r.reward = ns.reward(a.Q, r.reward);
            } else {
                r = env_step<ENV>(a.P, s, act, ob);
            }
            cnt += 1;
            const uint8_t st = b2g_step_type(cnt, a.P.max_len, r.done);
            a.rew[row] = (float)r.reward;
            a.step_type[row] = st;
            if (a.success) a.success[row] = r.success ? 1 : 0;
            if (st >= ST_TERMINAL) {
                if (a.ep_len) a.ep_len[row] = cnt;
                if (a.last_obs) {
#pragma unroll
                    for (int i = 0; i < O; ++i) a.last_obs[row * O + i] = ob[i];
                }
                // vec_worker.py:171: the finished env is reset immediately
                env_reset<ENV>(a.P, a.seed_lo, a.seed_hi, gid, (uint32_t)ep, s, ob);
                if constexpr (NORM) ns.obs(a.Q, ob);  // reset() normalises too (:71-73)
                ep += 1;
                cnt = 0;
            }
        }
    }
    if (owner) {
#pragma unroll
        for (int i = 0; i < S; ++i) a.state[(int64_t)i * N + n] = s[i];
#pragma unroll
        for (int i = 0; i < O; ++i) a.cur_obs[n * O + i] = ob[i];
        a.step_cnt[n] = cnt;
        a.episode[n] = ep;
        if constexpr (NORM) ns.store(a.Q, N, n);
    }
}

template <int ENV, int H, bool WS>
static int launch_rollout_ws(const RolloutArgs& a, cudaStream_t s) {
    // pick G (envs per warp) so that one wave of resident warps covers N when possible: ~8 warps per SM
    // at ~220 registers with W2 in registers, 12 (three 4-warp CTAs of 66 KB) with W2 in shared memory
    const int64_t cap = (int64_t)b2g_sm_count() * (WS ? 12 : 8);
    int G = 8;
    if (a.N <= cap) G = 1;
    else if (a.N <= cap * 2) G = 2;
    else if (a.N <= cap * 4) G = 4;
    if (const char* e = getenv("B2G_RO_G")) {  // experiment hook (tools/rollout_time.py)
        const int g = atoi(e);
        if (g == 1 || g == 2 || g == 4 || g == 8) G = g;
    }
    constexpr size_t dyn = WS ? sizeof(float) * H * (H + 4) : 0;
#define B2G_RO_K(GG, NN)                                                             \
    {                                                                                \
        auto k = rollout_kernel<ENV, H, GG, NN, WS>;                                 \
        if (WS) B2G_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn)); \
        k<<<blocks, RO_WARPS * 32, dyn, s>>>(a);                                     \
    }
#define B2G_RO(GG)                                                                   \
    {                                                                                \
        const int64_t warps = (a.N + GG - 1) / GG;                                   \
        const unsigned blocks = (unsigned)((warps + RO_WARPS - 1) / RO_WARPS);       \
        if (norm_active(a.Q)) B2G_RO_K(GG, true) else B2G_RO_K(GG, false)            \
    }
    switch (G) {
        case 1: B2G_RO(1); break;
        case 2: B2G_RO(2); break;
        case 4: B2G_RO(4); break;
        default: B2G_RO(8); break;
    }
#undef B2G_RO
#undef B2G_RO_K
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

template <int ENV, int H>
static int launch_rollout(const RolloutArgs& a, cudaStream_t s) {
    if constexpr (H > 64) {
        return launch_rollout_ws<ENV, H, true>(a, s);
    } else {
        static const bool ws = [] { const char* e = getenv("B2G_RO_WS"); return e && atoi(e) != 0; }();
        return ws ? launch_rollout_ws<ENV, H, true>(a, s) : launch_rollout_ws<ENV, H, false>(a, s);
    }
}

extern "C" int b2g_rollout(int env_id, const double* params, int max_episode_length,
                           const float* policy_params, int hidden, float min_log_std,
                           float max_log_std, uint64_t seed, int64_t env_id_offset,
                           uint32_t t_counter, const unsigned long long* t_counter_dev,
                           double* state, float* cur_obs,
                           int32_t* step_cnt, int32_t* episode, int n_steps, int64_t N,
                           float* obs, float* act, float* rew, float* logp,
                           uint8_t* step_type, float* mean, uint8_t* success,
                           float* last_obs, int32_t* ep_len, void* stream) {
    return b2g_rollout_norm(env_id, params, nullptr, max_episode_length, policy_params, hidden, min_log_std,
                            max_log_std, seed, env_id_offset, t_counter, t_counter_dev, state, cur_obs, step_cnt,
                            episode, n_steps, N, obs, act, rew, logp, step_type, mean, success, last_obs, ep_len,
                            stream);
}

extern "C" int b2g_rollout_norm(int env_id, const double* params, const b2g_norm* norm, int max_episode_length,
                                const float* policy_params, int hidden, float min_log_std,
                                float max_log_std, uint64_t seed, int64_t env_id_offset,
                                uint32_t t_counter, const unsigned long long* t_counter_dev,
                                double* state, float* cur_obs,
                                int32_t* step_cnt, int32_t* episode, int n_steps, int64_t N,
                                float* obs, float* act, float* rew, float* logp,
                                uint8_t* step_type, float* mean, uint8_t* success,
                                float* last_obs, int32_t* ep_len, void* stream) {
    RolloutArgs a;
    B2G_CHECK_ARG(env_params_from_host(env_id, params, max_episode_length, &a.P), "bad env_id or params");
    B2G_CHECK_ARG(norm_from_host(norm, &a.Q), "normalize_obs / normalize_reward need the stats buffer");
    B2G_CHECK_ARG(n_steps >= 0 && N >= 0, "negative shape");
    if (hidden != 32 && hidden != 64 && hidden != 128) {
        b2g_set_error("b2g_rollout: hidden_sizes=(%d,%d) not built (supported: (32,32), (64,64), (128,128))", hidden, hidden);
        return B2G_ERR_UNSUPPORTED;
    }
    if (n_steps == 0 || N == 0) return B2G_OK;
    B2G_CHECK_ARG(policy_params && state && cur_obs && step_cnt && episode && obs && act && rew && logp && step_type, "NULL pointer");
    a.seed_lo = (uint32_t)seed; a.seed_hi = (uint32_t)(seed >> 32);
    a.gid0 = env_id_offset; a.t_counter = t_counter; a.t_counter_dev = t_counter_dev; a.policy = policy_params;
    a.min_log_std = min_log_std; a.max_log_std = max_log_std;
    a.state = state; a.cur_obs = cur_obs; a.step_cnt = step_cnt; a.episode = episode;
    a.n_steps = n_steps; a.N = N;
    a.obs = obs; a.act = act; a.rew = rew; a.logp = logp; a.step_type = step_type;
    a.mean = mean; a.success = success; a.last_obs = last_obs; a.ep_len = ep_len;
    cudaStream_t s = (cudaStream_t)stream;
#define B2G_ENVH(E)                                                           \
    (hidden == 32 ? launch_rollout<E, 32>(a, s) : hidden == 64 ? launch_rollout<E, 64>(a, s) : launch_rollout<E, 128>(a, s))
    switch (env_id) {
        case B2G_ENV_POINT: return B2G_ENVH(B2G_ENV_POINT);
        case B2G_ENV_PENDULUM: return B2G_ENVH(B2G_ENV_PENDULUM);
        default: return B2G_ENVH(B2G_ENV_CARTPOLE);
    }
#undef B2G_ENVH
}

extern "C" int64_t b2g_mlp_param_count(int obs_dim, int hidden, int out_dim) {
    return MlpLayout{obs_dim, hidden, out_dim}.total();
}

// ---- stand-alone sampling for StochasticPolicy.get_actions
// (torch/policies/stochastic_policy.py:46-89: dist.sample() and agent_info), same
// Philox stream as the rollout kernel: action = mean + std * eps(gid0 + i, t_counter).
template <int A>
__global__ void __launch_bounds__(128)
policy_sample_kernel(const float* __restrict__ mean, float log_std, uint32_t seed_lo,
                     uint32_t seed_hi, int64_t gid0, uint32_t t_counter,
                     float* __restrict__ act, float* __restrict__ logp, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float stdv = expf(log_std), var = stdv * stdv, log_scale = logf(stdv);
    float eps[A];
    action_noise<A>(seed_lo, seed_hi, (uint32_t)(gid0 + i), t_counter, eps);
    float lp = 0.0f;
#pragma unroll
    for (int q = 0; q < A; ++q) {
        const float mu = mean[i * A + q];
        const float av = fmaf(stdv, eps[q], mu);
        act[i * A + q] = av;
        lp += normal_logp_term(av, mu, var, log_scale);
    }
    if (logp) logp[i] = lp;
}

extern "C" int b2g_policy_sample(const float* mean, int act_dim, float log_std, uint64_t seed,
                                 int64_t env_id_offset, uint32_t t_counter, float* act_out,
                                 float* logp_out, int64_t n, void* stream) {
    B2G_CHECK_ARG(n >= 0, "negative n");
    if (n == 0) return B2G_OK;
    B2G_CHECK_ARG(mean && act_out, "NULL pointer");
    const unsigned blocks = (unsigned)((n + 127) / 128);
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t lo = (uint32_t)seed, hi = (uint32_t)(seed >> 32);
    if (act_dim == 1) policy_sample_kernel<1><<<blocks, 128, 0, s>>>(mean, log_std, lo, hi, env_id_offset, t_counter, act_out, logp_out, n);
    else if (act_dim == 2) policy_sample_kernel<2><<<blocks, 128, 0, s>>>(mean, log_std, lo, hi, env_id_offset, t_counter, act_out, logp_out, n);
    else { b2g_set_error("b2g_policy_sample: act_dim %d not built (1 or 2)", act_dim); return B2G_ERR_UNSUPPORTED; }
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
