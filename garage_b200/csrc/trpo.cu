// Vector algebra of the reference's ConjugateGradientOptimizer
// (torch/optimizers/conjugate_gradient_optimizer.py) on flat float32 parameter-sized vectors:
//   _conjugate_gradient  :69-110   b2g_cg_init + cg_iters x (b2g_mlp_fvp, b2g_cg_step)
//   step                 :141-178  b2g_cg_sanitize (nan -> 0), b2g_mlp_fvp, b2g_cg_descent
//   _backtracking_line_search :218-260  b2g_line_search_point (params = prev - ratio * step)
// The vectors are a few thousand floats (one policy), so every kernel is ONE block: dot products
// are accumulated in float64 per thread over a fixed stride and folded in a fixed tree -- the same
// bits on every run -- and the scalars (r.r, the "converged" flag, the step size) stay in device
// memory, so a whole CG solve is enqueued without a host synchronisation.  Like the reference's
// tensors the vectors themselves are float32 and every scalar that multiplies them is rounded to
// float32 first.
#include "common.cuh"

constexpr int CG_NT = 1024;

__device__ __forceinline__ double block_sum(double v, double* sh) {
    const int tid = threadIdx.x;
    sh[tid] = v;
    __syncthreads();
#pragma unroll
    for (int s = CG_NT / 2; s > 0; s >>= 1) {
        if (tid < s) sh[tid] += sh[tid + s];
        __syncthreads();
    }
    const double out = sh[0];
    __syncthreads();
    return out;
}

// scal: [0] r.r  [1] converged flag  [2] step size  [3] x.Ax  [4] iterations run
__global__ void __launch_bounds__(CG_NT) cg_init_kernel(const float* __restrict__ b, float* __restrict__ x,
                                                        float* __restrict__ r, float* __restrict__ p,
                                                        double* __restrict__ scal, int n) {
    __shared__ double sh[CG_NT];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += CG_NT) {
        const float v = b[i];
        x[i] = 0.0f; r[i] = v; p[i] = v;
        acc += (double)v * (double)v;
    }
    const double rr = block_sum(acc, sh);
    if (threadIdx.x == 0) { scal[0] = (double)(float)rr; scal[1] = 0.0; scal[2] = 0.0; scal[3] = 0.0; scal[4] = 0.0; }
}

// one iteration of conjugate_gradient_optimizer.py:93-108 given z = A p; a no-op once converged
__global__ void __launch_bounds__(CG_NT) cg_step_kernel(const float* __restrict__ z, float* __restrict__ x,
                                                        float* __restrict__ r, float* __restrict__ p,
                                                        double* __restrict__ scal, int n, double tol) {
    __shared__ double sh[CG_NT];
    if (scal[1] != 0.0) return;  // `break` of the reference (uniform: every thread reads the same flag)
    const float rdotr = (float)scal[0];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += CG_NT) acc += (double)p[i] * (double)z[i];
    const float pz = (float)block_sum(acc, sh);
    const float v = rdotr / pz;
    acc = 0.0;
    for (int i = threadIdx.x; i < n; i += CG_NT) {
        x[i] = __fadd_rn(x[i], __fmul_rn(v, p[i]));       // x += v * p
        const float ri = __fsub_rn(r[i], __fmul_rn(v, z[i]));  // r -= v * z
        r[i] = ri;
        acc += (double)ri * (double)ri;
    }
    const float newrdotr = (float)block_sum(acc, sh);
    const float mu = newrdotr / rdotr;
    for (int i = threadIdx.x; i < n; i += CG_NT) p[i] = __fadd_rn(r[i], __fmul_rn(mu, p[i]));  // p = r + mu * p
    if (threadIdx.x == 0) {
        scal[0] = (double)newrdotr;
        scal[4] += 1.0;
        if ((double)newrdotr < tol) scal[1] = 1.0;
    }
}

__global__ void __launch_bounds__(CG_NT) cg_sanitize_kernel(float* __restrict__ x, int n) {
    for (int i = threadIdx.x; i < n; i += CG_NT)
        if (x[i] != x[i]) x[i] = 0.0f;  // step_dir[step_dir.ne(step_dir)] = 0.
}

// step_size = sqrt(2 delta / (x.Ax + 1e-8)) (1 if nan), descent = step_size * x   (:163-172)
__global__ void __launch_bounds__(CG_NT) cg_descent_kernel(const float* __restrict__ x, const float* __restrict__ Ax,
                                                           double delta, float* __restrict__ descent,
                                                           double* __restrict__ scal, int n) {
    __shared__ double sh[CG_NT];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += CG_NT) acc += (double)x[i] * (double)Ax[i];
    const float xAx = (float)block_sum(acc, sh);
    double step = sqrt(2.0 * delta * (1.0 / ((double)__fadd_rn(xAx, 1e-8f))));
    if (step != step) step = 1.0;
    const float sf = (float)step;
    for (int i = threadIdx.x; i < n; i += CG_NT) descent[i] = __fmul_rn(sf, x[i]);
    if (threadIdx.x == 0) { scal[2] = step; scal[3] = (double)xAx; }
}

// params = prev - ratio * descent   (:231-236; ratio is rounded to float32 like a Python scalar times a tensor)
__global__ void line_search_point_kernel(const float* __restrict__ prev, const float* __restrict__ descent,
                                         float ratio, float* __restrict__ params, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) params[i] = __fsub_rn(prev[i], __fmul_rn(ratio, descent[i]));
}

extern "C" int b2g_cg_init(const float* b, float* x, float* r, float* p, double* scal, int n, void* stream) {
    B2G_CHECK_ARG(b && x && r && p && scal && n > 0, "bad arguments");
    cg_init_kernel<<<1, CG_NT, 0, (cudaStream_t)stream>>>(b, x, r, p, scal, n);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
extern "C" int b2g_cg_step(const float* z, float* x, float* r, float* p, double* scal, int n, double residual_tol,
                           void* stream) {
    B2G_CHECK_ARG(z && x && r && p && scal && n > 0, "bad arguments");
    cg_step_kernel<<<1, CG_NT, 0, (cudaStream_t)stream>>>(z, x, r, p, scal, n, residual_tol);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
extern "C" int b2g_cg_sanitize(float* x, int n, void* stream) {
    B2G_CHECK_ARG(x && n > 0, "bad arguments");
    cg_sanitize_kernel<<<1, CG_NT, 0, (cudaStream_t)stream>>>(x, n);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
extern "C" int b2g_cg_descent(const float* x, const float* Ax, double max_constraint_value, float* descent,
                              double* scal, int n, void* stream) {
    B2G_CHECK_ARG(x && Ax && descent && scal && n > 0, "bad arguments");
    cg_descent_kernel<<<1, CG_NT, 0, (cudaStream_t)stream>>>(x, Ax, max_constraint_value, descent, scal, n);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
extern "C" int b2g_line_search_point(const float* prev, const float* descent, float ratio, float* params, int n,
                                     void* stream) {
    B2G_CHECK_ARG(prev && descent && params && n > 0, "bad arguments");
    line_search_point_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(prev, descent, ratio, params, n);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
