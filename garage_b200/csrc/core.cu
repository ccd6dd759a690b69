// Library plumbing (error string, device query) and small full-batch reductions.
#include <stdarg.h>

#include "common.cuh"

static thread_local char g_err[512] = "";

void b2g_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int b2g_sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

extern "C" int b2g_version(void) { return B2G_VERSION; }
extern "C" const char* b2g_last_error(void) { return g_err; }
extern "C" int b2g_device_sm_count(void) { return b2g_sm_count(); }

__global__ void counter_add_kernel(unsigned long long* c, unsigned long long d) { *c += d; }
extern "C" int b2g_counter_add(unsigned long long* counter, unsigned long long delta, void* stream) {
    B2G_CHECK_ARG(counter != nullptr, "NULL pointer");
    counter_add_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(counter, delta);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

// ---------------------------------------------------------------- moments --
// mean / unbiased variance / min of a float vector with fp64 accumulation, two
// passes (sum -> mean, then sum of squared deviations), deterministic: fixed
// grid, per-block partials reduced in block order by the last block to finish.
// Workspace layout (doubles): [0..G) partial sums, [G..2G) partial mins,
// [2G] counter (as uint32), G = MOM_BLOCKS.
static const int MOM_BLOCKS = 296;  // 2 x 148 SMs
static const int MOM_THREADS = 512;

__device__ __forceinline__ double block_reduce_sum(double v, double* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        v = (l < (blockDim.x >> 5)) ? sh[l] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    }
    return v;  // valid in thread 0
}
__device__ __forceinline__ float block_reduce_min(float v, float* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_down_sync(0xffffffffu, v, o));
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        v = (l < (blockDim.x >> 5)) ? sh[l] : INFINITY;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_down_sync(0xffffffffu, v, o));
    }
    return v;
}

// PASS 0: sum and min -> out[0] = mean, out[2] = min.  PASS 1: sum (x-mean)^2 ->
// out[1] = / (n-1).
template <int PASS>
__global__ void __launch_bounds__(MOM_THREADS)
moments_kernel(const float* __restrict__ x, int64_t n, double* __restrict__ out,
               double* __restrict__ ws) {
    __shared__ double shd[32];
    __shared__ float shf[32];
    __shared__ bool is_last;
    const double mean = PASS ? out[0] : 0.0;
    double s = 0.0;
    float mn = INFINITY;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        const float v = x[i];
        if (PASS == 0) { s += (double)v; mn = fminf(mn, v); }
        else { const double d = (double)v - mean; s += d * d; }
    }
    s = block_reduce_sum(s, shd);
    if (PASS == 0) mn = block_reduce_min(mn, shf);
    unsigned* counter = reinterpret_cast<unsigned*>(ws + 2 * MOM_BLOCKS);
    if (threadIdx.x == 0) {
        ws[blockIdx.x] = s;
        if (PASS == 0) ws[MOM_BLOCKS + blockIdx.x] = (double)mn;
        __threadfence();
        const unsigned done = atomicAdd(counter, 1u);
        is_last = (done == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    double t = 0.0;
    float m = INFINITY;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) {
        t += __ldcg(ws + b);
        if (PASS == 0) m = fminf(m, (float)__ldcg(ws + MOM_BLOCKS + b));
    }
    // fixed order inside the block: thread-strided then tree -> deterministic
    t = block_reduce_sum(t, shd);
    if (PASS == 0) m = block_reduce_min(m, shf);
    if (threadIdx.x == 0) {
        if (PASS == 0) { out[0] = t / (double)n; out[2] = (double)m; }
        else out[1] = (n > 1) ? t / (double)(n - 1) : NAN;  // torch.var of 1 element is nan
        *counter = 0u;
    }
}

extern "C" int64_t b2g_moments_ws_bytes(void) { return (2 * MOM_BLOCKS + 2) * sizeof(double); }

extern "C" int b2g_moments(const float* x, int64_t n, double* out3, void* ws, void* stream) {
    B2G_CHECK_ARG(n > 0 && x && out3 && ws, "empty input or NULL pointer");
    cudaStream_t s = (cudaStream_t)stream;
    double* w = reinterpret_cast<double*>(ws);
    B2G_CUDA(cudaMemsetAsync(w + 2 * MOM_BLOCKS, 0, sizeof(double), s));
    moments_kernel<0><<<MOM_BLOCKS, MOM_THREADS, 0, s>>>(x, n, out3, w);
    moments_kernel<1><<<MOM_BLOCKS, MOM_THREADS, 0, s>>>(x, n, out3, w);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

// ---- b2g_moments_static: the same two kernels with the scratch in a file-scope device array, for
// callers that may not allocate (CUDA graph capture, serving loops).  No counterpart in the reference.
// Every call on a device shares the array: calls that overlap on the GPU (different streams) race.
static __device__ __align__(256) double g_moments_scratch[2 * MOM_BLOCKS + 2];
static const int64_t MOMENTS_STATIC_MAX_N = (int64_t)1 << 28;
static const int MOMENTS_MAX_DEVICES = 64;

extern "C" int64_t b2g_moments_static_max_n(void) { return MOMENTS_STATIC_MAX_N; }

// device address of the array on the CURRENT device, resolved once per device (the symbol's name in
// host code is a host-side stand-in, not a device pointer)
static int moments_scratch_ptr(double** out) {
    static double* ptrs[MOMENTS_MAX_DEVICES] = {};
    int dev = -1;
    B2G_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MOMENTS_MAX_DEVICES) {
        b2g_set_error("b2g_moments_static: device ordinal %d outside [0, %d)", dev, MOMENTS_MAX_DEVICES);
        return B2G_ERR_INVALID;
    }
    double* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
    if (!p) {
        B2G_CUDA(cudaGetSymbolAddress((void**)&p, g_moments_scratch));
        __atomic_store_n(&ptrs[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return B2G_OK;
}

extern "C" int b2g_moments_static(const float* x, int64_t n, double* out3, void* stream) {
    B2G_CHECK_ARG(n > 0 && x && out3, "empty input or NULL pointer");
    if (n > MOMENTS_STATIC_MAX_N) {
        b2g_set_error("b2g_moments_static: n = %lld exceeds the built-in limit of %lld elements (use b2g_moments "
                      "with a caller-owned workspace)", (long long)n, (long long)MOMENTS_STATIC_MAX_N);
        return B2G_ERR_INVALID;
    }
    double* w = nullptr;
    const int rc = moments_scratch_ptr(&w);
    if (rc != B2G_OK) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    // the arrival counter must start at zero; the per-block slots are fully rewritten by every launch
    B2G_CUDA(cudaMemsetAsync(w + 2 * MOM_BLOCKS, 0, sizeof(double), s));
    moments_kernel<0><<<MOM_BLOCKS, MOM_THREADS, 0, s>>>(x, n, out3, w);
    moments_kernel<1><<<MOM_BLOCKS, MOM_THREADS, 0, s>>>(x, n, out3, w);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

__global__ void __launch_bounds__(256)
normalize_adv_kernel(const float* __restrict__ x, const double* __restrict__ st, int center,
                     int positive, float* __restrict__ y, int64_t n) {
    // torch computes in fp32: mean and var are fp32 tensors (vpg.py:371-377).
    const float mean = (float)st[0];
    const float denom = (float)st[1] + 1e-8f;
    const float mn = (float)st[2];
    float sub = 0.0f;
    if (positive) sub = center ? (mn - mean) / denom : mn;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * blockDim.x) {
        float v = x[i];
        if (center) v = (v - mean) / denom;
        y[i] = v - sub;
    }
}

extern "C" int b2g_normalize_adv(const float* x, const double* stats3, int center_adv,
                                 int positive_adv, float* y, int64_t n, void* stream) {
    B2G_CHECK_ARG(n >= 0, "negative n");
    if (n == 0) return B2G_OK;
    B2G_CHECK_ARG(x && y && stats3, "NULL pointer");
    const int threads = 256;
    int64_t blocks = (n + threads - 1) / threads;
    const int64_t cap = (int64_t)b2g_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    normalize_adv_kernel<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream>>>(x, stats3, center_adv, positive_adv, y, n);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

// ------------------------------------------------------------ permutation --
// Stands in for np.random.shuffle(self._ids) of np/optimizers/minibatch_dataset.py:35 in
// throughput mode (parity mode feeds the reference's own index lists instead); the Feistel
// network itself lives in common.cuh (b2g_perm_apply).
__global__ void __launch_bounds__(256)
permutation_kernel(uint32_t seed_lo, uint32_t seed_hi, const unsigned long long* __restrict__ counter_dev,
                   uint64_t counter, int64_t n, int half_bits, int32_t* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t c = counter + (counter_dev ? (uint64_t)*counter_dev : 0ull);
    uint32_t keys[8];
    b2g_perm_keys(seed_lo, seed_hi, c, keys);
    out[i] = (int32_t)b2g_perm_apply(keys, half_bits, (uint64_t)n, (uint64_t)i);
}

extern "C" int b2g_permutation(uint64_t seed, const unsigned long long* counter_dev,
                               uint64_t counter, int64_t n, int32_t* out, void* stream) {
    B2G_CHECK_ARG(n >= 0 && n <= 0x7fffffff, "n out of range for int32 ids");
    if (n == 0) return B2G_OK;
    B2G_CHECK_ARG(out != nullptr, "NULL pointer");
    const int half_bits = b2g_perm_half_bits(n);
    permutation_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        (uint32_t)seed, (uint32_t)(seed >> 32), counter_dev, counter, n, half_bits, out);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

// ---------------------------------------------------------- episode stats --
// log_performance (_functions.py:233-275) over an episode-major packed batch: per episode
// the undiscounted return sum_t r_t, the discounted return sum_t gamma^t r_t (what
// discount_cumsum(rewards, gamma)[0] is, np/_functions.py:111-128, float64 like the reference's
// rewards), whether any of its steps is TERMINAL and whether any step reports success.
// One warp per episode.  Lane l owns steps l, l+32, ...: a reverse Horner pass in gamma^32 gives
// sum_k r[l+32k] gamma^(32k); the lanes are then combined as sum_l gamma^l (.) by an xor
// butterfly, so the order of additions depends on the episode length only (deterministic).
// fixed_len > 0: episode e spans [e*fixed_len, (e+1)*fixed_len); else offsets[e], lengths[e].
template <typename R>
__global__ void __launch_bounds__(256)
episode_stats_kernel(const R* __restrict__ rew, const uint8_t* __restrict__ step_type,
                     const uint8_t* __restrict__ success, const int64_t* __restrict__ lengths,
                     const int64_t* __restrict__ offsets, int64_t fixed_len, double gamma,
                     int64_t n_eps, double* __restrict__ per_ep) {
    const int lane = threadIdx.x & 31;
    const int64_t e = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (e >= n_eps) return;
    const int64_t len = fixed_len > 0 ? fixed_len : lengths[e];
    const int64_t off = fixed_len > 0 ? e * fixed_len : offsets[e];
    double g32 = gamma;
#pragma unroll
    for (int i = 0; i < 5; ++i) g32 *= g32;
    double und = 0.0, disc = 0.0;
    int succ = 0, term = 0;
    if (lane < len) {
        const int64_t k_last = (len - 1 - lane) / 32;
        for (int64_t k = k_last; k >= 0; --k) {
            const int64_t t = off + lane + 32 * k;
            const double r = (double)rew[t];
            und += r;
            disc = fma(disc, g32, r);
            if (success) succ |= success[t] != 0;
            if (step_type) term |= step_type[t] == ST_TERMINAL;  // any(step_type == TERMINAL), :254-256
        }
    }
    double gl = 1.0;  // gamma^lane by repeated squaring over the bits of lane
    {
        double b = gamma;
        for (int m = lane; m; m >>= 1) {
            if (m & 1) gl *= b;
            b *= b;
        }
    }
    disc *= gl;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        und += __shfl_xor_sync(0xffffffffu, und, o);
        disc += __shfl_xor_sync(0xffffffffu, disc, o);
        succ |= __shfl_xor_sync(0xffffffffu, succ, o);
        term |= __shfl_xor_sync(0xffffffffu, term, o);
    }
    if (lane == 0) {
        per_ep[e] = und;
        per_ep[n_eps + e] = disc;
        per_ep[2 * n_eps + e] = term ? 1.0 : 0.0;
        per_ep[3 * n_eps + e] = succ ? 1.0 : 0.0;
    }
}

// out7 = [mean disc, mean und, population std und (np.std), max und, min und, termination rate,
// success rate (NaN without a success field)]: one block, fixed strides and tree => deterministic.
__global__ void __launch_bounds__(1024)
episode_summary_kernel(const double* __restrict__ per_ep, int64_t n_eps, int has_success,
                       double* __restrict__ out7) {
    __shared__ double sh[32];
    __shared__ double bcast;
    const int tid = threadIdx.x;
    double s_und = 0, s_disc = 0, s_term = 0, s_succ = 0, mx = -INFINITY, mn = INFINITY;
    for (int64_t e = tid; e < n_eps; e += blockDim.x) {
        const double u = per_ep[e];
        s_und += u;
        s_disc += per_ep[n_eps + e];
        s_term += per_ep[2 * n_eps + e];
        s_succ += per_ep[3 * n_eps + e];
        mx = fmax(mx, u);
        mn = fmin(mn, u);
    }
    const double n = (double)n_eps;
    s_und = block_reduce_sum(s_und, sh);
    if (tid == 0) bcast = s_und / n;
    __syncthreads();
    const double mean = bcast;
    double ss = 0;
    for (int64_t e = tid; e < n_eps; e += blockDim.x) {
        const double d = per_ep[e] - mean;
        ss += d * d;
    }
    ss = block_reduce_sum(ss, sh);
    s_disc = block_reduce_sum(s_disc, sh);
    s_term = block_reduce_sum(s_term, sh);
    s_succ = block_reduce_sum(s_succ, sh);
    // max / min: order-independent, reuse the sum tree's staging with fmax/fmin
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mx = fmax(mx, __shfl_down_sync(0xffffffffu, mx, o));
        mn = fmin(mn, __shfl_down_sync(0xffffffffu, mn, o));
    }
    __shared__ double shx[32], shn[32];
    if ((tid & 31) == 0) { shx[tid >> 5] = mx; shn[tid >> 5] = mn; }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) { mx = fmax(mx, shx[w]); mn = fmin(mn, shn[w]); }
        out7[0] = s_disc / n;
        out7[1] = mean;
        out7[2] = sqrt(ss / n);
        out7[3] = mx;
        out7[4] = mn;
        out7[5] = s_term / n;
        out7[6] = has_success ? s_succ / n : nan("");
    }
}

extern "C" int b2g_episode_stats(const void* rewards, int rewards_f64, const uint8_t* step_type,
                                 const uint8_t* success, const int64_t* lengths, const int64_t* offsets,
                                 int64_t fixed_len, double discount, int64_t n_eps, double* per_episode4,
                                 double* out7, void* stream) {
    B2G_CHECK_ARG(n_eps > 0, "n_eps must be positive");
    B2G_CHECK_ARG(rewards && per_episode4, "NULL pointer");
    B2G_CHECK_ARG(fixed_len > 0 || (lengths && offsets), "ragged episodes need lengths and offsets");
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned blocks = (unsigned)((n_eps * 32 + 255) / 256);
    if (rewards_f64)
        episode_stats_kernel<double><<<blocks, 256, 0, s>>>((const double*)rewards, step_type, success, lengths,
                                                            offsets, fixed_len, discount, n_eps, per_episode4);
    else
        episode_stats_kernel<float><<<blocks, 256, 0, s>>>((const float*)rewards, step_type, success, lengths,
                                                           offsets, fixed_len, discount, n_eps, per_episode4);
    B2G_CHECK_LAUNCH();
    if (out7) {
        episode_summary_kernel<<<1, 1024, 0, s>>>(per_episode4, n_eps, success != nullptr, out7);
        B2G_CHECK_LAUNCH();
    }
    return B2G_OK;
}
