// Shared device/host helpers for libb2g (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b2g.h"

// ---------------------------------------------------------------- errors --
void b2g_set_error(const char* fmt, ...);

#define B2G_CHECK_ARG(cond, msg)                                   \
    do {                                                           \
        if (!(cond)) {                                             \
            b2g_set_error("%s: %s", __func__, msg);                \
            return B2G_ERR_INVALID;                                \
        }                                                          \
    } while (0)

#define B2G_CHECK_LAUNCH()                                         \
    do {                                                           \
        cudaError_t e_ = cudaGetLastError();                       \
        if (e_ != cudaSuccess) {                                   \
            b2g_set_error("%s: launch failed: %s", __func__,       \
                          cudaGetErrorString(e_));                 \
            return (int)e_;                                        \
        }                                                          \
    } while (0)

#define B2G_CUDA(call)                                             \
    do {                                                           \
        cudaError_t e_ = (call);                                   \
        if (e_ != cudaSuccess) {                                   \
            b2g_set_error("%s: %s: %s", __func__, #call,           \
                          cudaGetErrorString(e_));                 \
            return (int)e_;                                        \
        }                                                          \
    } while (0)

int b2g_sm_count();  // cached cudaDevAttrMultiProcessorCount of the current device

// StepType codes, garage `_dtypes.py:15-42`.
enum : uint8_t { ST_FIRST = 0, ST_MID = 1, ST_TERMINAL = 2, ST_TIMEOUT = 3 };

// garage `StepType.get_step_type`, `_dtypes.py:44-68` (timeout wins over done;
// a 1-step terminal episode is TERMINAL, not FIRST).  max_len <= 0 == None/inf.
__host__ __device__ __forceinline__ uint8_t b2g_step_type(int step_cnt, int max_len,
                                                          bool done) {
    if (max_len > 0 && step_cnt >= max_len) return ST_TIMEOUT;
    if (done) return ST_TERMINAL;
    if (step_cnt == 1) return ST_FIRST;
    return ST_MID;
}

// ------------------------------------------------------------- Philox RNG --
// Philox4x32-10 (Salmon et al. 2011), counter based: no state in HBM.
struct Philox4 {
    uint32_t x, y, z, w;
};

__host__ __device__ __forceinline__ uint32_t b2g_mulhi(uint32_t a, uint32_t b) {
#ifdef __CUDA_ARCH__
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}

__host__ __device__ __forceinline__ Philox4 b2g_philox(uint32_t k0, uint32_t k1,
                                                       uint32_t c0, uint32_t c1,
                                                       uint32_t c2, uint32_t c3) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    const uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = b2g_mulhi(M0, c0), lo0 = M0 * c0;
        uint32_t hi1 = b2g_mulhi(M1, c2), lo1 = M1 * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += W0; k1 += W1;
    }
    return Philox4{c0, c1, c2, c3};
}

// uint32 -> (0,1] float / double
__host__ __device__ __forceinline__ float b2g_u01f(uint32_t u) {
    return ((float)(u >> 8) + 1.0f) * (1.0f / 16777216.0f);
}
__host__ __device__ __forceinline__ double b2g_u01d(uint32_t hi, uint32_t lo) {
    uint64_t v = (((uint64_t)hi << 32) | lo) >> 11;  // 53 bits
    return ((double)v + 0.5) * (1.0 / 9007199254740992.0);
}

// Philox stream ids (counter word c2)
enum : uint32_t { B2G_STREAM_ACTION = 0, B2G_STREAM_RESET = 1, B2G_STREAM_PERM = 2 };

// ------------------------------------------------------------ permutation --
// Minibatch shuffle without a sort: a keyed balanced Feistel network over the smallest
// even-width bit domain >= n, cycle-walked back into [0,n).  Counter based (keys =
// Philox(seed; counter)): the same (seed, counter) gives the same permutation on any number of
// ranks, and any kernel can evaluate single elements on the fly (b2g_perm_apply).
// rounds of the balanced Feistel network (4 = Luby-Rackoff's strong pseudo-random permutation with
// independent round functions; the id computation sits on the critical path of a minibatch step)
constexpr int B2G_PERM_ROUNDS = 4;
__host__ __device__ __forceinline__ uint32_t b2g_mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
    return x;
}
__host__ __device__ __forceinline__ void b2g_perm_keys(uint32_t seed_lo, uint32_t seed_hi, uint64_t c, uint32_t* keys) {
    const Philox4 k0 = b2g_philox(seed_lo, seed_hi, (uint32_t)c, (uint32_t)(c >> 32), B2G_STREAM_PERM, 0);
    const Philox4 k1 = b2g_philox(seed_lo, seed_hi, (uint32_t)c, (uint32_t)(c >> 32), B2G_STREAM_PERM, 1);
    keys[0] = k0.x; keys[1] = k0.y; keys[2] = k0.z; keys[3] = k0.w;
    keys[4] = k1.x; keys[5] = k1.y; keys[6] = k1.z; keys[7] = k1.w;
}
__host__ __device__ __forceinline__ int b2g_perm_half_bits(int64_t n) {
    int bits = 1;
    while (((int64_t)1 << bits) < n) ++bits;
    return (bits + 1) / 2 < 1 ? 1 : (bits + 1) / 2;
}
__host__ __device__ __forceinline__ uint64_t b2g_perm_apply(const uint32_t* keys, int half_bits, uint64_t n, uint64_t i) {
    const uint32_t mask = (1u << half_bits) - 1u;
    uint64_t x = i;
    do {
        uint32_t L = (uint32_t)(x >> half_bits) & mask, R = (uint32_t)x & mask;
#pragma unroll
        for (int r = 0; r < B2G_PERM_ROUNDS; ++r) {
            const uint32_t f = b2g_mix32(R ^ keys[r]) & mask;
            const uint32_t nl = R;
            R = L ^ f;
            L = nl;
        }
        x = ((uint64_t)L << half_bits) | R;
    } while (x >= n);
    return x;
}
