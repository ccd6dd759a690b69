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
