// GAE advantages + discounted returns: reverse (segmented) linear-recurrence scans.
//   replaces garage torch/_functions.py:25-85 (compute_advantages, pad + conv2d,
//   O(N*P^2)) and np/_functions.py:111-128 (discount_cumsum, scipy lfilter per row).
// HBM-bound: 17 algorithmic bytes per (t,n) element when fused
//   (r 4 + V 4 + step_type 1 read, A 4 + R 4 written).
#include "common.cuh"
#include <stdlib.h>
#include <stdio.h>

// ------------------------------------------------------ time-major kernel --
// One thread per environment column n, sequential over t (the recurrence), the
// warp covers 32 consecutive n => every access is one full 128 B line.  The
// loads do not depend on the carry, so U rows are prefetched into registers
// (double buffered) to keep U*9 B per thread in flight; U is picked by the host
// from N so that bytes-in-flight cover HBM latency even when N is small.
template <int U, bool HAS_ST, bool HAS_ADV, bool HAS_RET>
__global__ void __launch_bounds__(256)
gae_time_major_kernel(const float* __restrict__ rew, const float* __restrict__ val,
                      const uint8_t* __restrict__ st, const float* __restrict__ last_val,
                      float gamma, float gl, float* __restrict__ adv,
                      float* __restrict__ ret, int T, int64_t N) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    float next_v = last_val ? last_val[n] : 0.0f;
    float ca = 0.0f;
    double cr = 0.0;  // returns accumulate in fp64 like the reference (float64 lfilter, then cast)
    const double gamma_d = (double)gamma;

    float r0[U], v0[U], r1[U], v1[U];
    uint8_t s0[U], s1[U];

    auto load = [&](int t_hi, float* r, float* v, uint8_t* s) {
#pragma unroll
        for (int j = 0; j < U; ++j) {
            const int t = t_hi - 1 - j;
            if (t >= 0) {
                const int64_t i = (int64_t)t * N + n;
                r[j] = __ldcs(rew + i);
                if (HAS_ADV) v[j] = __ldcs(val + i);
                if (HAS_ST) s[j] = __ldcs(st + i);
            }
        }
    };
    auto consume = [&](int t_hi, const float* r, const float* v, const uint8_t* s) {
#pragma unroll
        for (int j = 0; j < U; ++j) {
            const int t = t_hi - 1 - j;
            if (t >= 0) {
                const int64_t i = (int64_t)t * N + n;
                if (HAS_ST && s[j] >= ST_TERMINAL) {  // episode ends here: cut carries
                    ca = 0.0f; cr = 0.0; next_v = 0.0f;
                }
                if (HAS_ADV) {
                    const float delta = r[j] + gamma * next_v - v[j];
                    ca = delta + gl * ca;
                    next_v = v[j];
                    __stcs(adv + i, ca);
                }
                if (HAS_RET) {
                    cr = (double)r[j] + gamma_d * cr;
                    __stcs(ret + i, (float)cr);
                }
            }
        }
    };

    int t_hi = T;
    load(t_hi, r0, v0, s0);
    while (t_hi > 0) {
        load(t_hi - U, r1, v1, s1);
        consume(t_hi, r0, v0, s0);
        t_hi -= U;
        if (t_hi <= 0) break;
        load(t_hi - U, r0, v0, s0);
        consume(t_hi, r1, v1, s1);
        t_hi -= U;
    }
}

// ---------------------------------------------- time-major, time-split form --
// When N is small the kernel above has too few threads to keep HBM busy (one thread per env, the
// bytes in flight are N * U * 9).  Here a CTA owns 32 consecutive envs and its W warps split TIME:
// warp w holds L rows of its envs in registers (coalesced 128 B lines as before), reduces them to
// the affine map  (ca, next_v, cr)_in -> (ca, next_v, cr)_out  of its chunk, the maps of the warps
// above it are folded through shared memory, and a second pass over the REGISTERS (not memory)
// writes the outputs from the true carry: 17 B per element of traffic, W x the loads in flight.
// The result differs from the sequential kernel only by the rounding of the folded carries.
template <int L, int MAXW, bool HAS_ST, bool HAS_ADV, bool HAS_RET>
__global__ void __launch_bounds__(MAXW * 32)
gae_time_split_kernel(const float* __restrict__ rew, const float* __restrict__ val,
                      const uint8_t* __restrict__ st, const float* __restrict__ last_val,
                      float gamma, float gl, float* __restrict__ adv, float* __restrict__ ret,
                      int T, int64_t N) {
    __shared__ float sA0[MAXW][32], sPA[MAXW][32], sPV[MAXW][32], sVF[MAXW][32], sCN[MAXW][32];
    __shared__ double sB0[MAXW][32], sPR[MAXW][32];
    __shared__ float cCa[32], cNv[32];
    __shared__ double cCr[32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, W = blockDim.x >> 5;
    const int64_t n = (int64_t)blockIdx.x * 32 + lane;
    const bool on = n < N;
    const int64_t nc = on ? n : N - 1;  // lanes past N read the last column and write nothing
    const double gamma_d = (double)gamma;
    if (w == 0) {
        cCa[lane] = 0.0f;
        cNv[lane] = (on && last_val) ? last_val[n] : 0.0f;
        cCr[lane] = 0.0;
    }
    __syncthreads();
    for (int top = T; top > 0; top -= W * L) {  // super-chunks of W*L rows, from the end of time
        const int hi = top - w * L;              // this warp's rows: [max(hi - L, 0), hi)
        const int cnt = hi >= L ? L : (hi > 0 ? hi : 0);
        float r[L], v[L];
        uint8_t sb[L];
        const int64_t i0 = (int64_t)(hi - 1) * N + nc;  // row hi-1 (only dereferenced when cnt > 0)
        if (cnt == L) {  // full chunk: no per-row guards, pointers step back one row at a time
            const float* pr_ = rew + i0;
            const float* pv_ = val + i0;
            const uint8_t* ps_ = st + i0;
#pragma unroll
            for (int j = 0; j < L; ++j) {
                r[j] = __ldcs(pr_); pr_ -= N;
                if (HAS_ADV) { v[j] = __ldcs(pv_); pv_ -= N; }
                if (HAS_ST) { sb[j] = __ldcs(ps_); ps_ -= N; }
            }
        } else {
#pragma unroll
            for (int j = 0; j < L; ++j) {
                r[j] = 0.0f; v[j] = 0.0f; sb[j] = 0;
                if (j < cnt) {
                    const int64_t i = i0 - (int64_t)j * N;
                    r[j] = __ldcs(rew + i);
                    if (HAS_ADV) v[j] = __ldcs(val + i);
                    if (HAS_ST) sb[j] = __ldcs(st + i);
                }
            }
        }
        // ---- the chunk as an affine map of the incoming carry (zero carry in: a0, b0)
        float a0 = 0.0f, nv = 0.0f, pa = 1.0f, pv = 0.0f, cn = 1.0f;
        double b0 = 0.0, pr = 1.0;
#pragma unroll
        for (int j = 0; j < L; ++j) {
            if (j < cnt) {
                if (HAS_ST && sb[j] >= ST_TERMINAL) {
                    a0 = 0.0f; nv = 0.0f; pa = 0.0f; pv = 0.0f; cn = 0.0f; b0 = 0.0; pr = 0.0;
                }
                if (HAS_ADV) {
                    a0 = (r[j] + gamma * nv - v[j]) + gl * a0;
                    pv = gamma * cn + gl * pv;
                    pa = gl * pa;
                    nv = v[j];
                    cn = 0.0f;
                }
                if (HAS_RET) {
                    b0 = (double)r[j] + gamma_d * b0;
                    pr = gamma_d * pr;
                }
            }
        }
        sA0[w][lane] = a0; sPA[w][lane] = pa; sPV[w][lane] = pv; sVF[w][lane] = nv; sCN[w][lane] = cn;
        sB0[w][lane] = b0; sPR[w][lane] = pr;
        __syncthreads();
        // ---- true carry into this chunk: fold the maps of the warps holding later rows
        float ca = cCa[lane], next_v = cNv[lane];
        double cr = cCr[lane];
        for (int u = 0; u < w; ++u) {
            const float ca2 = sA0[u][lane] + sPA[u][lane] * ca + sPV[u][lane] * next_v;
            next_v = sVF[u][lane] + sCN[u][lane] * next_v;
            ca = ca2;
            cr = sB0[u][lane] + sPR[u][lane] * cr;
        }
        // ---- second pass over the registers: outputs
        float* pa_ = adv + i0;
        float* pt_ = ret + i0;
#pragma unroll
        for (int j = 0; j < L; ++j) {
            if (j < cnt) {
                if (HAS_ST && sb[j] >= ST_TERMINAL) {
                    ca = 0.0f; cr = 0.0; next_v = 0.0f;
                }
                if (HAS_ADV) {
                    const float delta = r[j] + gamma * next_v - v[j];
                    ca = delta + gl * ca;
                    next_v = v[j];
                    if (on) __stcs(pa_, ca);
                    pa_ -= N;
                }
                if (HAS_RET) {
                    cr = (double)r[j] + gamma_d * cr;
                    if (on) __stcs(pt_, (float)cr);
                    pt_ -= N;
                }
            }
        }
        __syncthreads();  // every warp has read the maps and the super-chunk carry
        if (w == W - 1) { cCa[lane] = ca; cNv[lane] = next_v; cCr[lane] = cr; }
        __syncthreads();
    }
}

template <int L, int MAXW>
static int launch_time_split(const float* rew, const float* val, const uint8_t* st,
                             const float* last_val, float gamma, float lam, float* adv,
                             float* ret, int T, int64_t N, cudaStream_t s) {
    int W = (T + L - 1) / L;
    if (W > MAXW) W = MAXW;
    const unsigned blocks = (unsigned)((N + 31) / 32);
    const float gl = gamma * lam;
#define B2G_GAE_LAUNCH(ST, A, R)                                                                   \
    gae_time_split_kernel<L, MAXW, ST, A, R><<<blocks, W * 32, 0, s>>>(rew, val, st, last_val, gamma, \
                                                                      gl, adv, ret, T, N)
    const bool hs = st != nullptr, ha = adv != nullptr, hr = ret != nullptr;
    if (hs && ha && hr) B2G_GAE_LAUNCH(true, true, true);
    else if (hs && ha && !hr) B2G_GAE_LAUNCH(true, true, false);
    else if (hs && !ha && hr) B2G_GAE_LAUNCH(true, false, true);
    else if (!hs && ha && hr) B2G_GAE_LAUNCH(false, true, true);
    else if (!hs && ha && !hr) B2G_GAE_LAUNCH(false, true, false);
    else B2G_GAE_LAUNCH(false, false, true);
#undef B2G_GAE_LAUNCH
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

static bool gae_split_enabled() {
    static int on = -1;
    if (on < 0) {
        const char* e = getenv("B2G_GAE_SPLIT");
        on = (e && e[0] == '0') ? 0 : 1;
    }
    return on != 0;
}

template <int U>
static int launch_time_major(const float* rew, const float* val, const uint8_t* st,
                             const float* last_val, float gamma, float lam, float* adv,
                             float* ret, int T, int64_t N, cudaStream_t s) {
    // CTA size: the columns are independent (no shared memory), so small CTAs cost nothing and spread evenly --
    // with 256-thread CTAs N = 65,536 is 256 CTAs on 148 SMs (2 on some, 1 on others: the kernel runs as long
    // as the 2-CTA SMs); aim for >= 6 CTAs per SM
    const int64_t sms = b2g_sm_count();
    const int threads = (N >= sms * 256 * 6) ? 256 : (N >= sms * 128 * 6 ? 128 : (N >= sms * 64 * 6 ? 64 : 32));
    const unsigned blocks = (unsigned)((N + threads - 1) / threads);
    const float gl = gamma * lam;
#define B2G_GAE_LAUNCH(ST, A, R)                                                          \
    gae_time_major_kernel<U, ST, A, R><<<blocks, threads, 0, s>>>(rew, val, st, last_val, \
                                                                   gamma, gl, adv, ret, T, N)
    const bool hs = st != nullptr, ha = adv != nullptr, hr = ret != nullptr;
    if (hs && ha && hr) B2G_GAE_LAUNCH(true, true, true);
    else if (hs && ha && !hr) B2G_GAE_LAUNCH(true, true, false);
    else if (hs && !ha && hr) B2G_GAE_LAUNCH(true, false, true);
    else if (!hs && ha && hr) B2G_GAE_LAUNCH(false, true, true);
    else if (!hs && ha && !hr) B2G_GAE_LAUNCH(false, true, false);
    else B2G_GAE_LAUNCH(false, false, true);
#undef B2G_GAE_LAUNCH
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

extern "C" int b2g_gae_time_major(const float* rew, const float* val,
                                  const uint8_t* step_type, const float* last_val,
                                  float gamma, float lam, float* adv_out, float* ret_out,
                                  int T, int64_t N, void* stream) {
    B2G_CHECK_ARG(T >= 0 && N >= 0, "negative shape");
    B2G_CHECK_ARG(adv_out || ret_out, "adv_out and ret_out are both NULL");
    B2G_CHECK_ARG(rew != nullptr || T * N == 0, "rew is NULL");
    B2G_CHECK_ARG(!adv_out || val != nullptr || T * N == 0, "val is NULL but adv_out given");
    if (T == 0 || N == 0) return B2G_OK;
    cudaStream_t s = (cudaStream_t)stream;
    // few envs, long enough in time: split time over the warps of a CTA (env B2G_GAE_SPLIT=0 turns it off)
    if (T >= 64 && gae_split_enabled()) {
        // experiment hook: B2G_GAE_FORCE="L,W" forces one split configuration for every shape
        static int force = -1;
        if (force < 0) {
            const char* e = getenv("B2G_GAE_FORCE");
            force = 0;
            if (e) {
                int l = 0, w = 0;
                if (sscanf(e, "%d,%d", &l, &w) == 2) force = l * 100 + w;
            }
        }
#define B2G_SPLIT(LL, WW) return launch_time_split<LL, WW>(rew, val, step_type, last_val, gamma, lam, adv_out, ret_out, T, N, s)
        if (force == 3216) B2G_SPLIT(32, 16);
        if (force == 1608) B2G_SPLIT(16, 8);
        if (force == 1616) B2G_SPLIT(16, 16);
        if (force == 816) B2G_SPLIT(8, 16);
        if (force == 832) B2G_SPLIT(8, 32);
        if (force == 1632) B2G_SPLIT(16, 32);
        if (force == 432) B2G_SPLIT(4, 32);
        if (force == 1) goto sequential;
        if (N <= 4096) B2G_SPLIT(32, 16);
        static int64_t split_n = -1;
        if (split_n < 0) {
            const char* e = getenv("B2G_GAE_SPLIT_N");
            split_n = e ? atoll(e) : (1 << 15);
        }
        if (N < split_n) B2G_SPLIT(16, 8);
#undef B2G_SPLIT
    }
sequential:
    // bytes in flight ~ N * U * 9: aim for >= ~6 MB (HBM latency x bandwidth)
    if (N >= (1 << 17)) return launch_time_major<4>(rew, val, step_type, last_val, gamma, lam, adv_out, ret_out, T, N, s);
    if (N >= (1 << 16)) return launch_time_major<8>(rew, val, step_type, last_val, gamma, lam, adv_out, ret_out, T, N, s);
    if (N >= (1 << 15)) return launch_time_major<16>(rew, val, step_type, last_val, gamma, lam, adv_out, ret_out, T, N, s);
    return launch_time_major<32>(rew, val, step_type, last_val, gamma, lam, adv_out, ret_out, T, N, s);
}

// --------------------------------------------------- episode-major kernel --
// garage's own layout: rows = episodes [N,P], zero padded.  One warp per row,
// lanes cover 32 consecutive time steps (one 128 B line), segments walked from
// the end of the row.  Inside a segment the recurrence x_t += c * x_{t+1} is a
// Kogge-Stone scan over affine maps with warp shuffles (5 steps, powers
// c, c^2, c^4, c^8, c^16); across segments a scalar carry is applied with
// c^(32-lane).
template <typename F>
__device__ __forceinline__ F warp_reverse_decay_scan(F x, const F* cpow, int lane) {
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const int d = 1 << k;
        const F y = __shfl_down_sync(0xffffffffu, x, d);
        if (lane + d < 32) x += cpow[k] * y;
    }
    return x;
}

template <bool HAS_RET>
__global__ void __launch_bounds__(256)
gae_episode_major_kernel(const float* __restrict__ base, const float* __restrict__ rew,
                         float gamma, float gl, float* __restrict__ adv,
                         float* __restrict__ ret, int64_t N, int P) {
    const int lane = threadIdx.x & 31;
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= N) return;
    float pa[5], pr[5];
    pa[0] = gl; pr[0] = gamma;
#pragma unroll
    for (int k = 1; k < 5; ++k) { pa[k] = pa[k - 1] * pa[k - 1]; pr[k] = pr[k - 1] * pr[k - 1]; }
    // c^(32-lane): lane 31 -> c, lane 0 -> c^32
    float ta = 1.0f, tr = 1.0f;
    {
        const int e = 32 - lane;
#pragma unroll
        for (int k = 0; k < 5; ++k)
            if (e & (1 << k)) { ta *= pa[k]; tr *= pr[k]; }
        if (e & 32) { ta = pa[4] * pa[4]; tr = pr[4] * pr[4]; }
    }
    const float* b = base + row * (int64_t)P;
    const float* r = rew + row * (int64_t)P;
    double prd[5];
    prd[0] = (double)gamma;
#pragma unroll
    for (int k = 1; k < 5; ++k) prd[k] = prd[k - 1] * prd[k - 1];
    double trd = 1.0;
    {
        const int e = 32 - lane;
#pragma unroll
        for (int k = 0; k < 5; ++k)
            if (e & (1 << k)) trd *= prd[k];
        if (e & 32) trd = prd[4] * prd[4];
    }
    float carry_a = 0.0f, v_next_seg = 0.0f;  // V_P := 0
    double carry_r = 0.0;
    const int nseg = (P + 31) >> 5;
    for (int sgm = nseg - 1; sgm >= 0; --sgm) {
        const int t = (sgm << 5) + lane;
        const bool in = t < P;
        const float vt = in ? __ldcs(b + t) : 0.0f;
        const float rt = in ? __ldcs(r + t) : 0.0f;
        float vn = __shfl_down_sync(0xffffffffu, vt, 1);
        if (lane == 31) vn = v_next_seg;
        if (t + 1 >= P) vn = 0.0f;
        float xa = in ? (rt + gamma * vn - vt) : 0.0f;
        xa = warp_reverse_decay_scan(xa, pa, lane) + ta * carry_a;
        if (in) __stcs(adv + row * (int64_t)P + t, xa);
        carry_a = __shfl_sync(0xffffffffu, xa, 0);
        v_next_seg = __shfl_sync(0xffffffffu, vt, 0);
        if (HAS_RET) {
            // fp64 like the reference: discount_cumsum on float64 rewards, cast after
            double xr = warp_reverse_decay_scan((double)rt, prd, lane) + trd * carry_r;
            if (in) __stcs(ret + row * (int64_t)P + t, (float)xr);
            carry_r = __shfl_sync(0xffffffffu, xr, 0);
        }
    }
}

extern "C" int b2g_gae_episode_major(const float* baselines, const float* rewards,
                                     float gamma, float lam, float* adv_out,
                                     float* ret_out, int64_t N, int P, void* stream) {
    B2G_CHECK_ARG(N >= 0 && P >= 0, "negative shape");
    if (N == 0 || P == 0) return B2G_OK;
    B2G_CHECK_ARG(baselines && rewards && adv_out, "NULL pointer");
    const int threads = 256;
    const unsigned blocks = (unsigned)((N * 32 + threads - 1) / threads);
    cudaStream_t s = (cudaStream_t)stream;
    if (ret_out)
        gae_episode_major_kernel<true><<<blocks, threads, 0, s>>>(baselines, rewards, gamma, gamma * lam, adv_out, ret_out, N, P);
    else
        gae_episode_major_kernel<false><<<blocks, threads, 0, s>>>(baselines, rewards, gamma, gamma * lam, adv_out, ret_out, N, P);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

// discount_cumsum on float64 rows (np/_functions.py:111-128; garage's rewards
// are float64 python floats, vpg.py:149-153 casts the result to float32).
__global__ void __launch_bounds__(256)
discount_cumsum_f64_kernel(const double* __restrict__ x, double gamma,
                           double* __restrict__ out, int64_t N, int P) {
    const int lane = threadIdx.x & 31;
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (row >= N) return;
    double pr[5];
    pr[0] = gamma;
#pragma unroll
    for (int k = 1; k < 5; ++k) pr[k] = pr[k - 1] * pr[k - 1];
    double tr = 1.0;
    {
        const int e = 32 - lane;
#pragma unroll
        for (int k = 0; k < 5; ++k)
            if (e & (1 << k)) tr *= pr[k];
        if (e & 32) tr = pr[4] * pr[4];
    }
    double carry = 0.0;
    const int nseg = (P + 31) >> 5;
    for (int sgm = nseg - 1; sgm >= 0; --sgm) {
        const int t = (sgm << 5) + lane;
        const bool in = t < P;
        double xr = in ? x[row * (int64_t)P + t] : 0.0;
        xr = warp_reverse_decay_scan(xr, pr, lane) + tr * carry;
        if (in) out[row * (int64_t)P + t] = xr;
        carry = __shfl_sync(0xffffffffu, xr, 0);
    }
}

extern "C" int b2g_discount_cumsum_f64(const double* x, double gamma, double* out,
                                       int64_t N, int P, void* stream) {
    B2G_CHECK_ARG(N >= 0 && P >= 0, "negative shape");
    if (N == 0 || P == 0) return B2G_OK;
    B2G_CHECK_ARG(x && out, "NULL pointer");
    const int threads = 256;
    const unsigned blocks = (unsigned)((N * 32 + threads - 1) / threads);
    discount_cumsum_f64_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(x, gamma, out, N, P);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
