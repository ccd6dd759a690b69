// Fused minibatch kernel for garage's two-hidden-layer Gaussian MLPs:
//   gather rows -> MLP forward -> loss (PPO clipped surrogate / VPG / Gaussian-NLL
//   value loss, + entropy) -> full backward -> per-CTA gradient partials.
// replaces, per minibatch, torch/algos/vpg.py:250-293 (_train_policy /
// _train_value_function: zero_grad, loss, backward), torch/algos/ppo.py:96-132,
// vpg.py:324-347,408-455, value_functions/gaussian_mlp_value_function.py:81-97
// and the nn.Linear / tanh autograd graph of
// torch/modules/multi_headed_mlp_module.py:136-151.
//
// fp32 CUDA-core path (parity mode, 1e-5): 128-row tiles, 256 threads, register
// tiles of (R/TYN rows x 4 cols).  Activations live TRANSPOSED in shared memory
// ([feature][row], row stride 132 floats) so that every inner-loop operand is an
// LDS.128: a-operand rows are contiguous, b-operand weights are stored column-
// permuted so that a thread's 4 interleaved columns (c = tx + TXN*i) are adjacent;
// the interleaving makes the transposed epilogue stores bank-conflict free.
// Three H x H x 128 contractions per tile: Z2 = H1 W2^T, dW2 += dZ2^T H1,
// dH1 = dZ2 W2.  dW accumulators stay in registers across all tiles of a CTA.
#pragma once
#include "mlp.cuh"

enum : int { LOSS_PPO = 0, LOSS_VPG = 1, LOSS_VF = 2, LOSS_NONE = 3,
              LOSS_GIVEN = 4 };  // d loss / d mean supplied per row (dmu_in): the backward half of a Hessian-vector product
enum : int {
    FLAG_ENT_REG = 1,        // entropy_method == 'regularized' (vpg.py:341-343)
    FLAG_SOFTPLUS_ENT = 2,   // use_softplus_entropy (vpg.py:429-430)
    FLAG_STOP_ENT_GRAD = 4,  // stop_entropy_gradient (vpg.py:423-427)
};
constexpr int MLP_NSTATS = 8;  // [0] sum(-objective) [1] sum KL [2] sum entropy [3] #clipped [4] rows
constexpr int MLP_R = 128, MLP_NT = 256, MLP_RS = MLP_R + 4;

struct MlpArgs {
    const float* params;
    const int32_t* idx;  // minibatch row ids or NULL (rows start .. start+mb)
    int64_t start, mb;
    const float* obs;       // [B,O]
    const float* act;       // [B,A]    policy losses
    const float* adv;       // [B]
    const float* old_logp;  // [B]      PPO
    const float* ret;       // [B]      value loss
    const float* old_mean;  // [B,A] or NULL: KL(old || new) statistic
    const float* old_log_std;  // device scalar (clamped log-std of the old policy) or NULL
    int loss_kind, flags;
    float clip, ent_coeff, min_log_std, max_log_std;
    float* partial;         // [grid][P + MLP_NSTATS]
    float* mean_out;        // [mb,A] or NULL (forward output, minibatch order)
    unsigned long long* step_counter;  // Adam step, ticked once per backward launch (or NULL)
    long long* timing;      // debug: [grid][32] clock64 stamps of the tensor-core kernel (or NULL)
    // on-the-fly minibatch shuffle (tensor-core paths): row id = perm(start + m) over [0, perm_n)
    int64_t perm_n;         // 0: off
    int perm_half_bits;
    uint32_t perm_keys[8];
    // general-shape kernels (mlp_general.cuh): runtime dimensions inside the compiled capacities
    int O_rt, A_rt;
    const float* dmu_in;    // [mb,A] LOSS_GIVEN: d loss / d mean per minibatch row
    const float* vparams;   // tangent vector in the parameter layout (forward-mode pass)
};


// ---- per-row loss and d loss / d mean, shared by the CUDA-core and the tensor-core kernels
struct MlpDist {  // the (state-independent) scale of the Gaussian head, gaussian_mlp_module.py:171-176
    float stdv, var, log_scale;
    bool ls_pass;  // clamp backward mask
    // row-invariant pieces of the loss, evaluated once per weight load
    float inv_var, inv_2var;  // 1/var, 1/(2 var): the per-row divisions become multiplications
    float ent, dent_dls;      // entropy of Independent(Normal) (after the optional softplus) and d/d log_std
};
struct MlpRowAcc {  // row-owned sums: d loss / d log_std and the logging statistics
    float gLs, st_loss, st_kl, st_ent, st_clip, st_rows;
};

__device__ __forceinline__ MlpDist mlp_dist(const MlpArgs& a, float raw_ls) {
    MlpDist d;
    const float log_std = fminf(fmaxf(raw_ls, a.min_log_std), a.max_log_std);
    d.ls_pass = (raw_ls >= a.min_log_std) && (raw_ls <= a.max_log_std);
    d.stdv = expf(log_std);
    d.var = d.stdv * d.stdv;
    d.log_scale = logf(d.stdv);
    d.inv_var = 1.0f / d.var;
    d.inv_2var = 1.0f / (2.0f * d.var);
    d.ent = 0.0f;
    d.dent_dls = 0.0f;
    return d;
}
// entropy terms for an action dimension A (sum_a 0.5 + 0.5 log(2 pi) + log(std); F.softplus of it when
// use_softplus_entropy, vpg.py:429-430)
template <int A>
__device__ __forceinline__ void mlp_dist_entropy(const MlpArgs& a, MlpDist& d, int A_rt = A) {
    float ent = (float)A_rt * (1.4189385332046727f + d.log_scale);
    float dent_dls = (float)A_rt;
    if (a.flags & FLAG_SOFTPLUS_ENT) {
        const float sp = (ent > 20.0f) ? ent : log1pf(expf(ent));
        dent_dls *= (ent > 20.0f) ? 1.0f : 1.0f / (1.0f + expf(-ent));
        ent = sp;
    }
    d.ent = ent;
    d.dent_dls = dent_dls;
}

// per-row inputs of the loss (gathered at the minibatch row id)
template <int A>
struct MlpRowIn {
    float act[A], old_mean[A], adv, old_logp, ret;
};
template <int A>
__device__ __forceinline__ MlpRowIn<A> mlp_row_in(const MlpArgs& a, int64_t src) {
    MlpRowIn<A> in;
    in.adv = 0.0f; in.old_logp = 0.0f; in.ret = 0.0f;
#pragma unroll
    for (int q = 0; q < A; ++q) { in.act[q] = 0.0f; in.old_mean[q] = 0.0f; }
    if (a.loss_kind == LOSS_VF) {
        in.ret = a.ret[src];
    } else if (a.loss_kind != LOSS_NONE) {
#pragma unroll
        for (int q = 0; q < A; ++q) in.act[q] = a.act[src * A + q];
        in.adv = a.adv[src];
        if (a.loss_kind == LOSS_PPO) in.old_logp = a.old_logp[src];
        if (a.old_mean) {
#pragma unroll
            for (int q = 0; q < A; ++q) in.old_mean[q] = a.old_mean[src * A + q];
        }
    }
    return in;
}

// loss of one valid row given the network output mu: writes dmu (x 1/mb later), accumulates acc.
// STATS = false drops the logging sums (the fused training loop does not report them);
// FAST uses ex2.approx for the probability ratio (2 ulp; the tensor-core paths).
template <int A, bool STATS = true, bool FAST = false>
__device__ __forceinline__ void mlp_row_loss(const MlpArgs& a, const MlpRowIn<A>& in, const float (&mu)[A],
                                             const MlpDist& ds, float (&dmu)[A], MlpRowAcc& acc, int A_rt = A) {
    // A_rt <= A: action dimensions in use (general-shape kernels compile for a capacity A and keep the
    // unused means, actions and W3 rows at zero)
    const float fA = (float)A_rt;
    const float stdv = ds.stdv, log_scale = ds.log_scale, inv_var = ds.inv_var;
    const bool ls_pass = ds.ls_pass;
    const float ent = ds.ent, dent_dls = ds.dent_dls;  // mlp_dist_entropy<A>() ran after mlp_dist()
    if (STATS) {
        acc.st_rows += 1.0f;
        acc.st_ent += ent;
    }
    if (a.loss_kind == LOSS_NONE || a.loss_kind == LOSS_GIVEN) {
        // forward only (baselines / means), or d loss / d mean supplied by the caller
    } else if (a.loss_kind == LOSS_VF) {
        // -log N(ret; mu, std): gaussian_mlp_value_function.py:81-97
        const float d = in.ret - mu[0];
        if (STATS) acc.st_loss += (d * d) * ds.inv_2var + log_scale + 0.9189385332046727f;
        dmu[0] = -d * inv_var;                                  // d loss / d mu (x 1/mb later)
        if (ls_pass) acc.gLs += 1.0f - (d * d) * inv_var;       // d loss / d log_std
    } else {
        float ll = 0.0f, dsq = 0.0f, dact[A];
#pragma unroll
        for (int q = 0; q < A; ++q) {
            dact[q] = in.act[q] - mu[q];
            dsq += dact[q] * dact[q];
        }
        // Independent(Normal).log_prob: sum_a -(a - mu)^2 / (2 var) - log(std) - log(sqrt(2 pi))
        ll = -dsq * ds.inv_2var - fA * (log_scale + 0.9189385332046727f);
        const float adv = in.adv;
        float obj, dobj_dll;
        if (a.loss_kind == LOSS_PPO) {  // ppo.py:113-132
            float ratio;
            if (FAST) {
                asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ratio) : "f"((ll - in.old_logp) * 1.4426950408889634f));
            } else {
                ratio = expf(ll - in.old_logp);
            }
            const float lo = 1.0f - a.clip, hi = 1.0f + a.clip;
            const float rc = fminf(fmaxf(ratio, lo), hi);
            const float surr = ratio * adv, surr_c = rc * adv;
            obj = fminf(surr, surr_c);
            const bool inside = (ratio >= lo) && (ratio <= hi);
            // torch.min ties split the gradient, clamp passes it inside [lo,hi]
            const bool pass = inside || (surr < surr_c);
            dobj_dll = pass ? ratio * adv : 0.0f;
            if (STATS && !inside) acc.st_clip += 1.0f;
        } else {  // vpg.py:452-455
            obj = ll * adv;
            dobj_dll = adv;
        }
        float dobj_dls = 0.0f;
        if (a.flags & FLAG_ENT_REG) {  // vpg.py:341-343
            obj += a.ent_coeff * ent;
            if (!(a.flags & FLAG_STOP_ENT_GRAD)) dobj_dls = a.ent_coeff * dent_dls;
        }
        if (STATS) acc.st_loss -= obj;
        const float s = -dobj_dll * inv_var;
#pragma unroll
        for (int q = 0; q < A; ++q) dmu[q] = s * dact[q];
        if (ls_pass) acc.gLs -= dobj_dll * (dsq * inv_var - fA) + dobj_dls;
        if (STATS && a.old_mean) {  // torch kl_normal_normal(old, new), vpg.py:381-406
            const float so = expf(a.old_log_std ? __ldg(a.old_log_std) : 0.0f);
            const float vr = (so / stdv) * (so / stdv);
#pragma unroll
            for (int q = 0; q < A; ++q) {
                const float t1 = (in.old_mean[q] - mu[q]) / stdv;
                if (q < A_rt) acc.st_kl += 0.5f * (vr + t1 * t1 - 1.0f - logf(vr));
            }
        }
    }
}

template <int H>
struct MlpSmem {
    static constexpr int TXN = H / 4, TYN = MLP_NT / TXN, RPT = MLP_R / TYN, KPT = H / TYN;
};

__device__ __forceinline__ float4 lds4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void sts4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

template <int H, int O, int A, bool BWD>
__global__ void __launch_bounds__(MLP_NT, 2) mlp_train_kernel(const MlpArgs a) {
    constexpr int R = MLP_R, RS = MLP_RS;
    constexpr int TXN = MlpSmem<H>::TXN, TYN = MlpSmem<H>::TYN, RPT = MlpSmem<H>::RPT,
                  KPT = MlpSmem<H>::KPT;
    constexpr int WX = TXN / 8;
    const MlpLayout L{O, H, A};
    const int P = L.total();

    extern __shared__ __align__(16) float smem[];
    float* sX = smem;                       // [O][RS]
    float* sH1 = sX + O * RS;               // [H][RS]
    float* sH2 = sH1 + H * RS;              // [H][RS]  (h2, then dz2 in place)
    float* sDmu = sH2 + H * RS;             // [A][RS]
    float* sW2T = sDmu + A * RS;            // [k][tx*4+i] = W2[tx+TXN*i][k]
    float* sW2P = sW2T + H * H;             // [c][tx*4+i] = W2[c][tx+TXN*i]   (BWD)
    float* sW1 = sW2P + (BWD ? H * H : 0);  // [H][O]
    float* sB1 = sW1 + H * O;
    float* sB2 = sB1 + H;
    float* sW3 = sB2 + H;                   // [A][H]
    float* sB3 = sW3 + A * H;               // [A] (+pad)

    const int tid = threadIdx.x, w = tid >> 5, l = tid & 31;
    const int tx = (w % WX) * 8 + (l & 7);
    const int ty = (w / WX) * 4 + (l >> 3);
    const int r0 = ty * RPT;

    // ---- stage weights
    for (int i = tid; i < H * H; i += MLP_NT) {
        const int c = i / H, k = i % H;
        const float v = a.params[L.off_w2() + i];
        sW2T[k * H + (c % TXN) * 4 + (c / TXN)] = v;
        if (BWD) sW2P[c * H + (k % TXN) * 4 + (k / TXN)] = v;
    }
    for (int i = tid; i < H * O; i += MLP_NT) sW1[i] = a.params[L.off_w1() + i];
    for (int i = tid; i < H; i += MLP_NT) {
        sB1[i] = a.params[L.off_b1() + i];
        sB2[i] = a.params[L.off_b2() + i];
    }
    for (int i = tid; i < A * H; i += MLP_NT) sW3[i] = a.params[L.off_w3() + i];
    if (tid < A) sB3[tid] = a.params[L.off_b3() + tid];
    MlpDist dist = mlp_dist(a, a.params[L.off_logstd()]);
    mlp_dist_entropy<A>(a, dist);
    if (BWD && a.step_counter && blockIdx.x == 0 && tid == 0) *a.step_counter += 1ull;
    __syncthreads();

    // ---- persistent gradient accumulators
    float gW2[4][KPT], gW3[A][4], gB2[4], gW1[4][O], gB1[4];
    float gB3[A];
    MlpRowAcc racc{};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int j = 0; j < KPT; ++j) gW2[i][j] = 0.0f;
#pragma unroll
        for (int q = 0; q < A; ++q) gW3[q][i] = 0.0f;
#pragma unroll
        for (int o = 0; o < O; ++o) gW1[i][o] = 0.0f;
        gB2[i] = 0.0f; gB1[i] = 0.0f;
    }
#pragma unroll
    for (int q = 0; q < A; ++q) gB3[q] = 0.0f;

    const int64_t ntiles = (a.mb + R - 1) / R;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t m0 = tile * R;
        const int valid = (int)((a.mb - m0) < R ? (a.mb - m0) : R);
        // ---- 1. gather rows (threads 0..R-1 own one row each for the loss stage)
        int64_t src = -1;
        if (tid < R) {
            if (tid < valid) src = a.idx ? (int64_t)a.idx[m0 + tid] : (a.start + m0 + tid);
#pragma unroll
            for (int o = 0; o < O; ++o) sX[o * RS + tid] = (src >= 0) ? a.obs[src * O + o] : 0.0f;
        }
        __syncthreads();
        // ---- 2. layer 1: h1 = tanh(W1 x + b1) -> sH1[c][r]
        {
            float xr[O][RPT];
#pragma unroll
            for (int o = 0; o < O; ++o)
#pragma unroll
                for (int q = 0; q < RPT / 4; ++q) {
                    const float4 v = lds4(sX + o * RS + r0 + 4 * q);
                    xr[o][4 * q] = v.x; xr[o][4 * q + 1] = v.y; xr[o][4 * q + 2] = v.z; xr[o][4 * q + 3] = v.w;
                }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int c = tx + TXN * i;
                float h[RPT];
#pragma unroll
                for (int r = 0; r < RPT; ++r) {
                    float z = sB1[c];
#pragma unroll
                    for (int o = 0; o < O; ++o) z = fmaf(sW1[c * O + o], xr[o][r], z);
                    h[r] = tanhf(z);
                }
#pragma unroll
                for (int q = 0; q < RPT / 4; ++q)
                    sts4(sH1 + c * RS + r0 + 4 * q, make_float4(h[4 * q], h[4 * q + 1], h[4 * q + 2], h[4 * q + 3]));
            }
        }
        __syncthreads();
        // ---- 3. GEMM1: z2[r][c] = sum_k h1[r][k] W2[c][k]; h2 = tanh(z2 + b2) -> sH2[c][r]
        {
            float acc[RPT][4];
#pragma unroll
            for (int r = 0; r < RPT; ++r)
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[r][i] = 0.0f;
#pragma unroll 8
            for (int k = 0; k < H; ++k) {
                float av[RPT];
#pragma unroll
                for (int q = 0; q < RPT / 4; ++q) {
                    const float4 v = lds4(sH1 + k * RS + r0 + 4 * q);
                    av[4 * q] = v.x; av[4 * q + 1] = v.y; av[4 * q + 2] = v.z; av[4 * q + 3] = v.w;
                }
                const float4 b = lds4(sW2T + k * H + tx * 4);
#pragma unroll
                for (int r = 0; r < RPT; ++r) {
                    acc[r][0] = fmaf(av[r], b.x, acc[r][0]);
                    acc[r][1] = fmaf(av[r], b.y, acc[r][1]);
                    acc[r][2] = fmaf(av[r], b.z, acc[r][2]);
                    acc[r][3] = fmaf(av[r], b.w, acc[r][3]);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int c = tx + TXN * i;
                const float bb = sB2[c];
#pragma unroll
                for (int q = 0; q < RPT / 4; ++q)
                    sts4(sH2 + c * RS + r0 + 4 * q,
                         make_float4(tanhf(acc[4 * q][i] + bb), tanhf(acc[4 * q + 1][i] + bb),
                                     tanhf(acc[4 * q + 2][i] + bb), tanhf(acc[4 * q + 3][i] + bb)));
            }
        }
        __syncthreads();
        // ---- 4. layer 3 + loss, one thread per row: mu = W3 h2 + b3
        if (tid < R) {
            float mu[A];
#pragma unroll
            for (int q = 0; q < A; ++q) mu[q] = sB3[q];
#pragma unroll 8
            for (int c = 0; c < H; ++c) {
                const float h = sH2[c * RS + tid];
#pragma unroll
                for (int q = 0; q < A; ++q) mu[q] = fmaf(sW3[q * H + c], h, mu[q]);
            }
            float dmu[A];
#pragma unroll
            for (int q = 0; q < A; ++q) dmu[q] = 0.0f;
            if (src >= 0) {
                if (a.mean_out) {
#pragma unroll
                    for (int q = 0; q < A; ++q) a.mean_out[(m0 + tid) * A + q] = mu[q];
                }
                mlp_row_loss<A>(a, mlp_row_in<A>(a, src), mu, dist, dmu, racc);
            }
            if (BWD) {
#pragma unroll
                for (int q = 0; q < A; ++q) {
                    sDmu[q * RS + tid] = dmu[q];
                    gB3[q] += dmu[q];
                }
            }
        }
        if (!BWD) {
            __syncthreads();
            continue;
        }
        __syncthreads();
        // ---- 5. dz2 = (W3^T dmu) * (1 - h2^2), in place over sH2; dW3, db2 partials
        {
            float dm[A][RPT];
#pragma unroll
            for (int q = 0; q < A; ++q)
#pragma unroll
                for (int p = 0; p < RPT / 4; ++p) {
                    const float4 v = lds4(sDmu + q * RS + r0 + 4 * p);
                    dm[q][4 * p] = v.x; dm[q][4 * p + 1] = v.y; dm[q][4 * p + 2] = v.z; dm[q][4 * p + 3] = v.w;
                }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int c = tx + TXN * i;
                float w3c[A];
#pragma unroll
                for (int q = 0; q < A; ++q) w3c[q] = sW3[q * H + c];
#pragma unroll
                for (int p = 0; p < RPT / 4; ++p) {
                    const float4 hv = lds4(sH2 + c * RS + r0 + 4 * p);
                    const float h[4] = {hv.x, hv.y, hv.z, hv.w};
                    float dz[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float dh = 0.0f;
#pragma unroll
                        for (int q = 0; q < A; ++q) {
                            dh = fmaf(w3c[q], dm[q][4 * p + e], dh);
                            gW3[q][i] = fmaf(dm[q][4 * p + e], h[e], gW3[q][i]);
                        }
                        dz[e] = dh * (1.0f - h[e] * h[e]);
                        gB2[i] += dz[e];
                    }
                    sts4(sH2 + c * RS + r0 + 4 * p, make_float4(dz[0], dz[1], dz[2], dz[3]));
                }
            }
        }
        __syncthreads();
        // ---- 6a. GEMM3: dW2[c][k] += sum_r dz2[r][c] h1[r][k]   (c = tx+TXN*i, k = ty+TYN*j)
#pragma unroll 2
        for (int r = 0; r < R; r += 4) {
            float4 d[4], hh[KPT];
#pragma unroll
            for (int i = 0; i < 4; ++i) d[i] = lds4(sH2 + (tx + TXN * i) * RS + r);
#pragma unroll
            for (int j = 0; j < KPT; ++j) hh[j] = lds4(sH1 + (ty + TYN * j) * RS + r);
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < KPT; ++j) {
                    gW2[i][j] = fmaf(d[i].x, hh[j].x, gW2[i][j]);
                    gW2[i][j] = fmaf(d[i].y, hh[j].y, gW2[i][j]);
                    gW2[i][j] = fmaf(d[i].z, hh[j].z, gW2[i][j]);
                    gW2[i][j] = fmaf(d[i].w, hh[j].w, gW2[i][j]);
                }
        }
        // ---- 6b. GEMM2: dh1[r][k] = sum_c dz2[r][c] W2[c][k]; dz1 = dh1 (1 - h1^2); dW1, db1
        {
            float acc[RPT][4];
#pragma unroll
            for (int r = 0; r < RPT; ++r)
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[r][i] = 0.0f;
#pragma unroll 8
            for (int c = 0; c < H; ++c) {
                float av[RPT];
#pragma unroll
                for (int p = 0; p < RPT / 4; ++p) {
                    const float4 v = lds4(sH2 + c * RS + r0 + 4 * p);
                    av[4 * p] = v.x; av[4 * p + 1] = v.y; av[4 * p + 2] = v.z; av[4 * p + 3] = v.w;
                }
                const float4 b = lds4(sW2P + c * H + tx * 4);
#pragma unroll
                for (int r = 0; r < RPT; ++r) {
                    acc[r][0] = fmaf(av[r], b.x, acc[r][0]);
                    acc[r][1] = fmaf(av[r], b.y, acc[r][1]);
                    acc[r][2] = fmaf(av[r], b.z, acc[r][2]);
                    acc[r][3] = fmaf(av[r], b.w, acc[r][3]);
                }
            }
            float xr[O][RPT];
#pragma unroll
            for (int o = 0; o < O; ++o)
#pragma unroll
                for (int p = 0; p < RPT / 4; ++p) {
                    const float4 v = lds4(sX + o * RS + r0 + 4 * p);
                    xr[o][4 * p] = v.x; xr[o][4 * p + 1] = v.y; xr[o][4 * p + 2] = v.z; xr[o][4 * p + 3] = v.w;
                }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int k = tx + TXN * i;
#pragma unroll
                for (int p = 0; p < RPT / 4; ++p) {
                    const float4 hv = lds4(sH1 + k * RS + r0 + 4 * p);
                    const float h[4] = {hv.x, hv.y, hv.z, hv.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float dz1 = acc[4 * p + e][i] * (1.0f - h[e] * h[e]);
                        gB1[i] += dz1;
#pragma unroll
                        for (int o = 0; o < O; ++o) gW1[i][o] = fmaf(dz1, xr[o][4 * p + e], gW1[i][o]);
                    }
                }
            }
        }
        __syncthreads();
    }  // tiles

    // ---- write this CTA's partial: [P gradients | MLP_NSTATS statistics]
    float* out = a.partial + (int64_t)blockIdx.x * (P + MLP_NSTATS);
    float* red = sH1;  // scratch (sH1 and sH2 are contiguous, 2*H*RS floats)
    if (BWD) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < KPT; ++j)
                out[L.off_w2() + (tx + TXN * i) * H + (ty + TYN * j)] = gW2[i][j];
        // column-owned sums are replicated over the TYN row groups: reduce in fixed order
        constexpr int VV = A + O + 2, V = 4 * VV;
        static_assert(MLP_NT * V <= 2 * H * MLP_RS, "reduction scratch too small");
        float* mine = red + (ty * TXN + tx) * V;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
            for (int q = 0; q < A; ++q) mine[i * VV + q] = gW3[q][i];
            mine[i * VV + A] = gB2[i];
#pragma unroll
            for (int o = 0; o < O; ++o) mine[i * VV + A + 1 + o] = gW1[i][o];
            mine[i * VV + A + O + 1] = gB1[i];
        }
        __syncthreads();
        for (int e = tid; e < TXN * V; e += MLP_NT) {
            const int txp = e / V, v = e % V, i = v / VV, s = v % VV;
            float sum = 0.0f;
            for (int y = 0; y < TYN; ++y) sum += red[(y * TXN + txp) * V + v];
            const int c = txp + TXN * i;
            int dst;
            if (s < A) dst = L.off_w3() + s * H + c;
            else if (s == A) dst = L.off_b2() + c;
            else if (s <= A + O) dst = L.off_w1() + c * O + (s - A - 1);
            else dst = L.off_b1() + c;
            out[dst] = sum;
        }
        __syncthreads();
    }
    // row-owned scalars (threads 0..R-1): db3, dlog_std, statistics
    {
        constexpr int NS = A + 1 + 5;
        float vals[NS];
#pragma unroll
        for (int q = 0; q < A; ++q) vals[q] = gB3[q];
        vals[A] = racc.gLs; vals[A + 1] = racc.st_loss; vals[A + 2] = racc.st_kl; vals[A + 3] = racc.st_ent;
        vals[A + 4] = racc.st_clip; vals[A + 5] = racc.st_rows;
#pragma unroll
        for (int s = 0; s < NS; ++s) {
            float v = vals[s];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
            if (l == 0) red[s * 8 + w] = v;
        }
        __syncthreads();
        if (tid < NS) {
            float sum = 0.0f;
            for (int y = 0; y < MLP_NT / 32; ++y) sum += red[tid * 8 + y];
            if (tid < A) { if (BWD) out[L.off_b3() + tid] = sum; }
            else if (tid == A) { if (BWD) out[L.off_logstd()] = sum; }
            else out[P + (tid - A - 1)] = sum;
        }
        if (tid >= NS && tid < NS + (MLP_NSTATS - 5)) out[P + 5 + (tid - NS)] = 0.0f;
    }
}

template <int H, int O, int A, bool BWD>
constexpr size_t mlp_train_smem_bytes() {
    return sizeof(float) * (size_t)(O * MLP_RS + 2 * H * MLP_RS + A * MLP_RS + H * H +
                                    (BWD ? H * H : 0) + H * O + 2 * H + A * H + 4);
}
