// Host launchers for the fused MLP minibatch kernel, partial-gradient reduction,
// Adam, and the forward-only evaluation paths.
#include "mlp_tc.cuh"
#include "mlp_general.cuh"
#include <stdlib.h>

// CTAs per launch: one per 128-row tile, capped at SMs x resident CTAs per SM
// (persistent grid-stride over tiles beyond that).
template <int H, int O, int A, bool BWD>
static int mlp_occupancy() {
    static int occ = 0;
    if (occ == 0) {
        constexpr size_t smem = mlp_train_smem_bytes<H, O, A, BWD>();
        cudaFuncSetAttribute(mlp_train_kernel<H, O, A, BWD>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, mlp_train_kernel<H, O, A, BWD>,
                                                          MLP_NT, smem) != cudaSuccess || n < 1)
            n = 1;
        occ = n;
    }
    return occ;
}

template <int H, int O, int A, bool BWD>
static int mlp_grid(int64_t rows) {
    const int64_t tiles = (rows + MLP_R - 1) / MLP_R;
    const int64_t cap = (int64_t)b2g_sm_count() * mlp_occupancy<H, O, A, BWD>();
    return (int)(tiles < cap ? (tiles > 0 ? tiles : 1) : cap);
}

template <int H, int O, int A, bool BWD>
static int launch_mlp(const MlpArgs& a, cudaStream_t s) {
    constexpr size_t smem = mlp_train_smem_bytes<H, O, A, BWD>();
    const int grid = mlp_grid<H, O, A, BWD>(a.mb);
    mlp_train_kernel<H, O, A, BWD><<<grid, MLP_NT, smem, s>>>(a);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

// ---- tensor-core path (hidden 64): one CTA per SM, 128-row tiles
static int g_mlp_path = -1;  // 0 auto (tcgen05 for hidden 64 backward), 1 CUDA-core fp32 only, 2 = 0, 3 general-shape kernels
static int mlp_path() {
    if (g_mlp_path < 0) {
        const char* e = getenv("B2G_MLP_PATH");
        g_mlp_path = (e && (e[0] == 'f' || e[0] == '1')) ? 1 : ((e && (e[0] == 'g' || e[0] == '3')) ? 3 : 0);
    }
    return g_mlp_path;
}
extern "C" int b2g_set_mlp_path(int path) {
    B2G_CHECK_ARG(path >= 0 && path <= 3, "path must be 0 (auto), 1 (CUDA-core fp32), 2 (tcgen05) or 3 (general-shape kernels)");
    g_mlp_path = path == 2 ? 0 : path;
    return B2G_OK;
}
extern "C" int b2g_get_mlp_path(void) { return mlp_path(); }

// debug hook (tools/profile_mlp.py --timing): device buffer [grid][32] of clock64 stamps
static long long* g_mlp_timing = nullptr;
extern "C" int b2g_mlp_debug_timing(long long* buf) { g_mlp_timing = buf; return B2G_OK; }

static bool force_general() { return mlp_path() == 3; }

static int mlp_tc_grid(int64_t rows) {
    const int64_t tiles = (rows + tc::R - 1) / tc::R;
    const int64_t cap = 2 * (int64_t)b2g_sm_count();  // two CTAs per SM (smem and TMEM halves)
    return (int)(tiles < cap ? (tiles > 0 ? tiles : 1) : cap);
}

template <int O, int A>
static int launch_mlp_tc(const MlpArgs& a, cudaStream_t s) {
    constexpr size_t smem = tc::smem_bytes<O, A>();
    static bool configured = false;
    if (!configured) {
        B2G_CUDA(cudaFuncSetAttribute(tc::mlp_train_tc_kernel<O, A>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem));
        B2G_CUDA(cudaFuncSetAttribute(tc::mlp_train_tc_kernel<O, A>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                      cudaSharedmemCarveoutMaxShared));
        configured = true;
    }
    tc::mlp_train_tc_kernel<O, A><<<mlp_tc_grid(a.mb), tc::NT, smem, s>>>(a);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

template <int O, int A>
static int launch_mlp_eval_tc(const MlpArgs& a, cudaStream_t s) {
    constexpr size_t smem = tc::smem_bytes<O, A>();
    static bool configured = false;
    if (!configured) {
        B2G_CUDA(cudaFuncSetAttribute(tc::mlp_eval_tc_kernel<O, A>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)smem));
        B2G_CUDA(cudaFuncSetAttribute(tc::mlp_eval_tc_kernel<O, A>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                      cudaSharedmemCarveoutMaxShared));
        configured = true;
    }
    tc::mlp_eval_tc_kernel<O, A><<<mlp_tc_grid(a.mb), tc::NT, smem, s>>>(a);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

#define B2G_MLP_SHAPES(X) X(64, 3, 1) X(64, 3, 2) X(64, 4, 1) X(32, 3, 1) X(32, 3, 2) X(32, 4, 1)

static int mlp_unsupported(int O, int H, int A) {
    b2g_set_error("MLP shape obs=%d hidden=(%d,%d) out=%d is not built into libb2g "
                  "(built: obs 1..8, hidden in {32,64,128}, out 1..4)", O, H, H, A);
    return B2G_ERR_UNSUPPORTED;
}

// ---- general shapes (mlp_general.cuh): capacities the kernels are compiled for
int b2g_gen_launch_h32(int mode, int op, int ap, const MlpArgs& a, cudaStream_t s);
int b2g_gen_launch_h64(int mode, int op, int ap, const MlpArgs& a, cudaStream_t s);
int b2g_gen_launch_h128(int mode, int op, int ap, const MlpArgs& a, cudaStream_t s);
int b2g_gen_grid_h32(int mode, int op, int ap, int64_t rows);
int b2g_gen_grid_h64(int mode, int op, int ap, int64_t rows);
int b2g_gen_grid_h128(int mode, int op, int ap, int64_t rows);
static bool gen_supported(int O, int H, int A) {
    return (H == 32 || H == 64 || H == 128) && O >= 1 && O <= 8 && A >= 1 && A <= 4;
}
static int gen_op(int O) { return O <= 4 ? 4 : 8; }
static int gen_ap(int A) { return A <= 1 ? 1 : (A == 2 ? 2 : 4); }
static int gen_launch(int mode, int O, int H, int A, MlpArgs a, cudaStream_t s) {
    a.O_rt = O; a.A_rt = A;
    if (H == 32) return b2g_gen_launch_h32(mode, gen_op(O), gen_ap(A), a, s);
    if (H == 64) return b2g_gen_launch_h64(mode, gen_op(O), gen_ap(A), a, s);
    return b2g_gen_launch_h128(mode, gen_op(O), gen_ap(A), a, s);
}
static int gen_grid(int mode, int O, int H, int A, int64_t rows) {
    if (H == 32) return b2g_gen_grid_h32(mode, gen_op(O), gen_ap(A), rows);
    if (H == 64) return b2g_gen_grid_h64(mode, gen_op(O), gen_ap(A), rows);
    return b2g_gen_grid_h128(mode, gen_op(O), gen_ap(A), rows);
}
// path 3 (B2G_MLP_PATH=general / b2g_set_mlp_path(3)): every shape through the general kernels (tests)
static bool force_general();

template <bool BWD>
static int dispatch_mlp(int O, int H, int A, const MlpArgs& a, cudaStream_t s) {
    if (force_general() && gen_supported(O, H, A)) return gen_launch(BWD ? 1 : 0, O, H, A, a, s);
#define B2G_CASE(HH, OO, AA)                                                         \
    if (H == HH && O == OO && A == AA) {                                             \
        if constexpr (HH == 64) {                                                    \
            if (mlp_path() == 0) return BWD ? launch_mlp_tc<OO, AA>(a, s) : launch_mlp_eval_tc<OO, AA>(a, s); \
        }                                                                            \
        return launch_mlp<HH, OO, AA, BWD>(a, s);                                    \
    }
    B2G_MLP_SHAPES(B2G_CASE)
#undef B2G_CASE
    if (gen_supported(O, H, A)) return gen_launch(BWD ? 1 : 0, O, H, A, a, s);
    return mlp_unsupported(O, H, A);
}

// number of per-CTA partial rows a launch over `rows` rows writes (<0: unsupported shape)
extern "C" int b2g_mlp_num_partials(int obs_dim, int hidden, int out_dim, int backward,
                                    int64_t rows) {
    if (force_general() && gen_supported(obs_dim, hidden, out_dim))
        return gen_grid(backward ? 1 : 0, obs_dim, hidden, out_dim, rows);
#define B2G_CASE(HH, OO, AA)                                                                   \
    if (hidden == HH && obs_dim == OO && out_dim == AA) {                                      \
        if (HH == 64 && mlp_path() == 0) return mlp_tc_grid(rows);                             \
        return backward ? mlp_grid<HH, OO, AA, true>(rows) : mlp_grid<HH, OO, AA, false>(rows); \
    }
    B2G_MLP_SHAPES(B2G_CASE)
#undef B2G_CASE
    if (gen_supported(obs_dim, hidden, out_dim)) return gen_grid(backward ? 1 : 0, obs_dim, hidden, out_dim, rows);
    return mlp_unsupported(obs_dim, hidden, out_dim);
}

// floats of workspace that any launch for this shape can need
extern "C" int64_t b2g_mlp_ws_floats(int obs_dim, int hidden, int out_dim) {
    const int P = MlpLayout{obs_dim, hidden, out_dim}.total();
    return (int64_t)b2g_sm_count() * 4 * (P + MLP_NSTATS);
}

extern "C" int b2g_mlp_minibatch_grad(int obs_dim, int hidden, int out_dim, const float* params,
                                      const int32_t* idx, int64_t start, int64_t mb,
                                      const float* obs, const float* act, const float* adv,
                                      const float* old_logp, const float* ret, int loss_kind,
                                      int flags, float clip, float ent_coeff, float min_log_std,
                                      float max_log_std, float* ws_partial,
                                      unsigned long long* step_counter, void* stream) {
    B2G_CHECK_ARG(mb > 0, "empty minibatch");
    B2G_CHECK_ARG(params && obs && ws_partial, "NULL pointer");
    B2G_CHECK_ARG(loss_kind >= LOSS_PPO && loss_kind <= LOSS_VF, "bad loss_kind");
    if (loss_kind == LOSS_VF) B2G_CHECK_ARG(ret != nullptr && out_dim == 1, "value loss needs ret and out_dim 1");
    else B2G_CHECK_ARG(act && adv && (loss_kind != LOSS_PPO || old_logp), "policy loss needs act, adv (and old_logp for PPO)");
    MlpArgs a{};
    a.params = params; a.idx = idx; a.start = start; a.mb = mb; a.obs = obs; a.act = act;
    a.adv = adv; a.old_logp = old_logp; a.ret = ret; a.old_mean = nullptr; a.old_log_std = nullptr;
    a.loss_kind = loss_kind; a.flags = flags; a.clip = clip; a.ent_coeff = ent_coeff;
    a.min_log_std = min_log_std; a.max_log_std = max_log_std; a.partial = ws_partial;
    a.mean_out = nullptr; a.step_counter = step_counter; a.timing = g_mlp_timing;
    return dispatch_mlp<true>(obs_dim, hidden, out_dim, a, (cudaStream_t)stream);
}

extern "C" int b2g_mlp_eval(int obs_dim, int hidden, int out_dim, const float* params,
                            const int32_t* idx, int64_t start, int64_t n, const float* obs,
                            const float* act, const float* adv, const float* old_logp,
                            const float* ret, const float* old_mean, const float* old_log_std,
                            int loss_kind, int flags, float clip, float ent_coeff,
                            float min_log_std, float max_log_std, float* mean_out,
                            float* ws_partial, void* stream) {
    B2G_CHECK_ARG(n > 0, "empty batch");
    B2G_CHECK_ARG(params && obs && ws_partial, "NULL pointer");
    B2G_CHECK_ARG(loss_kind >= LOSS_PPO && loss_kind <= LOSS_NONE, "bad loss_kind");
    if (loss_kind == LOSS_VF) B2G_CHECK_ARG(ret != nullptr && out_dim == 1, "value loss needs ret and out_dim 1");
    else if (loss_kind != LOSS_NONE) B2G_CHECK_ARG(act && adv && (loss_kind != LOSS_PPO || old_logp), "policy loss needs act, adv (and old_logp for PPO)");
    MlpArgs a{};
    a.params = params; a.idx = idx; a.start = start; a.mb = n; a.obs = obs; a.act = act;
    a.adv = adv; a.old_logp = old_logp; a.ret = ret; a.old_mean = old_mean; a.old_log_std = old_log_std;
    a.loss_kind = loss_kind; a.flags = flags; a.clip = clip; a.ent_coeff = ent_coeff;
    a.min_log_std = min_log_std; a.max_log_std = max_log_std; a.partial = ws_partial;
    a.mean_out = mean_out; a.step_counter = nullptr;
    return dispatch_mlp<false>(obs_dim, hidden, out_dim, a, (cudaStream_t)stream);
}

// ------------------------------------------------ partial reduction + Adam --
// grad[p] = scale * sum_b partial[b][p] (fixed order -> deterministic);
// stats[s] = sum_b partial[b][P+s] in fp64.  With `params` given, applies
// torch.optim.Adam's single-tensor update (make_optimizer, _functions.py:25-65;
// amsgrad=False, weight_decay=0) in the same pass.
constexpr int RED_PX = 32, RED_SL = 8;  // 32 parameters x 8 partial-row slices per block
__global__ void __launch_bounds__(RED_PX * RED_SL)
reduce_adam_kernel(const float* __restrict__ partial, int nb, int P, float scale,
                   float* __restrict__ grad_out, double* __restrict__ stats_out,
                   float* __restrict__ params, float* __restrict__ m, float* __restrict__ v,
                   const unsigned long long* __restrict__ step, double lr, double beta1,
                   double beta2, double eps) {
    __shared__ double sh[RED_SL][RED_PX];
    const int px = threadIdx.x & (RED_PX - 1), sl = threadIdx.x / RED_PX;
    const int p = blockIdx.x * RED_PX + px;
    const int W = P + MLP_NSTATS;
    const bool is_stat = p >= P;
    // slice sl sums rows sl, sl+8, ...: a warp reads 32 consecutive floats of one row
    float s = 0.0f;
    double sd = 0.0;
    if (p < W) {
        if (is_stat) {
            for (int b = sl; b < nb; b += RED_SL) sd += (double)partial[(int64_t)b * W + p];
        } else {
#pragma unroll 4
            for (int b = sl; b < nb; b += RED_SL) s += partial[(int64_t)b * W + p];
            sd = (double)s;
        }
    }
    sh[sl][px] = sd;
    __syncthreads();
    if (sl != 0 || p >= W) return;
    if (is_stat) {
        double t = 0.0;
#pragma unroll
        for (int k = 0; k < RED_SL; ++k) t += sh[k][px];
        if (stats_out) stats_out[p - P] = t;
        return;
    }
    float sum = 0.0f;  // fixed order -> deterministic
#pragma unroll
    for (int k = 0; k < RED_SL; ++k) sum += (float)sh[k][px];
    const float g = sum * scale;
    if (grad_out) grad_out[p] = g;
    if (params) {
        const double t = (double)(*step);
        const double bc1 = 1.0 - pow(beta1, t), bc2 = 1.0 - pow(beta2, t);
        const float step_size = (float)(lr / bc1);
        const float bc2_sqrt = (float)sqrt(bc2);
        float mm = m[p], vv = v[p];
        mm = mm + (float)(1.0 - beta1) * (g - mm);                  // exp_avg.lerp_(grad, 1-beta1)
        vv = vv * (float)beta2 + ((float)(1.0 - beta2) * g) * g;     // mul_(beta2).addcmul_(g, g, 1-beta2)
        const float denom = sqrtf(vv) / bc2_sqrt + (float)eps;
        params[p] = params[p] - step_size * (mm / denom);           // addcdiv_(exp_avg, denom, -step_size)
        m[p] = mm; v[p] = vv;
    }
}

extern "C" int b2g_mlp_reduce(const float* ws_partial, int n_partials, int n_params, float scale,
                              float* grad_out, double* stats_out, float* params, float* adam_m,
                              float* adam_v, const unsigned long long* step_counter, double lr,
                              double beta1, double beta2, double eps, void* stream) {
    B2G_CHECK_ARG(ws_partial && n_partials > 0 && n_params > 0, "bad arguments");
    B2G_CHECK_ARG(!params || (adam_m && adam_v && step_counter), "Adam needs m, v and the step counter");
    const int W = n_params + MLP_NSTATS;
    reduce_adam_kernel<<<(W + RED_PX - 1) / RED_PX, RED_PX * RED_SL, 0, (cudaStream_t)stream>>>(
        ws_partial, n_partials, n_params, scale, grad_out, stats_out, params, adam_m, adam_v,
        step_counter, lr, beta1, beta2, eps);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

// Adam alone, on an (all-reduced) gradient vector.
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ params, const float* __restrict__ grad, float gscale,
            float* __restrict__ m, float* __restrict__ v,
            const unsigned long long* __restrict__ step, double lr, double beta1, double beta2,
            double eps, int n) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const double t = (double)(*step);
    const double bc1 = 1.0 - pow(beta1, t), bc2 = 1.0 - pow(beta2, t);
    const float step_size = (float)(lr / bc1);
    const float bc2_sqrt = (float)sqrt(bc2);
    const float g = grad[p] * gscale;
    float mm = m[p], vv = v[p];
    mm = mm + (float)(1.0 - beta1) * (g - mm);
    vv = vv * (float)beta2 + ((float)(1.0 - beta2) * g) * g;
    const float denom = sqrtf(vv) / bc2_sqrt + (float)eps;
    params[p] = params[p] - step_size * (mm / denom);
    m[p] = mm; v[p] = vv;
}

extern "C" int b2g_adam_step(float* params, const float* grad, float grad_scale, float* adam_m,
                             float* adam_v, const unsigned long long* step_counter, double lr,
                             double beta1, double beta2, double eps, int n, void* stream) {
    B2G_CHECK_ARG(params && grad && adam_m && adam_v && step_counter && n > 0, "bad arguments");
    adam_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(params, grad, grad_scale, adam_m, adam_v,
                                                                   step_counter, lr, beta1, beta2, eps, n);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

// ------------------------------------------------ Fisher-vector product --
// out = (1/n) sum_rows J_r^T diag(1/var) J_r vec  (+ 2 A vec[log_std] on the log_std entry when the
// clamp passes) + reg_coeff * vec: the Hessian of mean KL(old || new) at new == old for the
// state-independent-std Gaussian head, i.e. what `_build_hessian_vector_product(f_constraint, ...)`
// (torch/optimizers/conjugate_gradient_optimizer.py:18-66) returns with `f_constraint` =
// `VPG._compute_kl_constraint` (torch/algos/vpg.py:381-406) when the step starts at the old policy
// (TRPO's default full-batch single-epoch update, trpo.py:119-144).  The second-order term of the
// network vanishes there because it multiplies (mu_old - mu_new) = 0, so the Gauss-Newton product is exact.
__global__ void fvp_finish_kernel(float* __restrict__ out, const float* __restrict__ vec,
                                  const float* __restrict__ params, int P, int A, float min_ls, float max_ls,
                                  float reg) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    float v = out[p] + reg * vec[p];
    if (p == 0) {
        const float raw = params[0];
        out[0] = reg * vec[0] + ((raw >= min_ls && raw <= max_ls) ? 2.0f * (float)A * vec[0] : 0.0f);
        return;
    }
    out[p] = v;
}

extern "C" int b2g_mlp_fvp(int obs_dim, int hidden, int out_dim, const float* params, const float* vec,
                           const int32_t* idx, int64_t start, int64_t n, const float* obs,
                           float min_log_std, float max_log_std, float reg_coeff, float* u_scratch,
                           float* ws_partial, float* out, void* stream) {
    B2G_CHECK_ARG(n > 0, "empty batch");
    B2G_CHECK_ARG(params && vec && obs && u_scratch && ws_partial && out, "NULL pointer");
    if (!gen_supported(obs_dim, hidden, out_dim)) return mlp_unsupported(obs_dim, hidden, out_dim);
    cudaStream_t s = (cudaStream_t)stream;
    MlpArgs a{};
    a.params = params; a.idx = idx; a.start = start; a.mb = n; a.obs = obs;
    a.loss_kind = LOSS_NONE; a.min_log_std = min_log_std; a.max_log_std = max_log_std;
    a.partial = ws_partial; a.vparams = vec; a.mean_out = u_scratch;
    int rc = gen_launch(gen::M_JVP, obs_dim, hidden, out_dim, a, s);
    if (rc != B2G_OK) return rc;
    a.vparams = nullptr; a.mean_out = nullptr; a.dmu_in = u_scratch; a.loss_kind = LOSS_GIVEN;
    rc = gen_launch(gen::M_BWD, obs_dim, hidden, out_dim, a, s);
    if (rc != B2G_OK) return rc;
    const int P = MlpLayout{obs_dim, hidden, out_dim}.total();
    rc = b2g_mlp_reduce(ws_partial, gen_grid(gen::M_BWD, obs_dim, hidden, out_dim, n), P, 1.0f / (float)n, out, nullptr,
                        nullptr, nullptr, nullptr, nullptr, 0.0, 0.0, 0.0, 0.0, stream);
    if (rc != B2G_OK) return rc;
    fvp_finish_kernel<<<(P + 255) / 256, 256, 0, s>>>(out, vec, params, P, out_dim, min_log_std, max_log_std, reg_coeff);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
