// Flat parameter layout of garage's GaussianMLPModule with two equal hidden
// layers (torch/modules/gaussian_mlp_module.py:195-305, state_dict order of
// torch/policies/gaussian_mlp_policy.py / value_functions/gaussian_mlp_value_function.py):
//   [ log_std (1) | W1 [H,O] | b1 [H] | W2 [H,H] | b2 [H] | W3 [A,H] | b3 [A] ]
// all float32, row-major [out,in] exactly as nn.Linear.weight.
#pragma once
#include "common.cuh"

struct MlpLayout {
    int O, H, A;
    __host__ __device__ int off_logstd() const { return 0; }
    __host__ __device__ int off_w1() const { return 1; }
    __host__ __device__ int off_b1() const { return 1 + H * O; }
    __host__ __device__ int off_w2() const { return off_b1() + H; }
    __host__ __device__ int off_b2() const { return off_w2() + H * H; }
    __host__ __device__ int off_w3() const { return off_b2() + H; }
    __host__ __device__ int off_b3() const { return off_w3() + A * H; }
    __host__ __device__ int total() const { return off_b3() + A; }
};

// torch.distributions.Normal.log_prob summed over the action dims (Independent):
//   -((a - mu)^2) / (2 var) - log(std) - log(sqrt(2 pi))
__device__ __forceinline__ float normal_logp_term(float a, float mu, float var, float log_std) {
    const float d = a - mu;
    return -(d * d) / (2.0f * var) - log_std - 0.9189385332046727f;
}

// two standard normals from two uint32 (Box-Muller)
__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float* n0, float* n1) {
    const float u1 = b2g_u01f(a), u2 = b2g_u01f(b);
    const float r = sqrtf(-2.0f * logf(u1));
    float s, c;
    sincospif(2.0f * u2, &s, &c);
    *n0 = r * c;
    *n1 = r * s;
}

// eps[A] for (global env id, global step) from the Philox action stream
template <int A>
__device__ __forceinline__ void action_noise(uint32_t seed_lo, uint32_t seed_hi, uint32_t gid,
                                             uint32_t t, float* eps) {
#pragma unroll
    for (int q = 0; q < (A + 3) / 4; ++q) {
        const Philox4 r = b2g_philox(seed_lo, seed_hi, gid, t, B2G_STREAM_ACTION, (uint32_t)q);
        float n[4];
        box_muller(r.x, r.y, &n[0], &n[1]);
        box_muller(r.z, r.w, &n[2], &n[3]);
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (q * 4 + i < A) eps[q * 4 + i] = n[i];
    }
}
