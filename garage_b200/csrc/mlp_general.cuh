// General-shape form of the fused minibatch kernel (mlp_train.cuh): any observation dimension
// O <= OP, any action dimension A <= AP, hidden in {32, 64, 128} -- what the reference accepts
// through `hidden_sizes`, `env_spec.observation_space.flat_dim` and `action_space.flat_dim`
// (torch/policies/gaussian_mlp_policy.py:53-87, torch/modules/multi_headed_mlp_module.py:49-106)
// for two equal tanh hidden layers.  OP / AP are compile-time CAPACITIES (4 or 8 / 1, 2 or 4); the
// real dimensions are runtime values: unused observation columns, means, actions and W3 rows are
// kept at zero, so they drop out of every sum.  fp32 CUDA-core arithmetic, same tiling as
// mlp_train.cuh (register tiles of RPT rows x 4 columns, transposed activations in shared memory,
// LDS.128 operands); 128-row tiles for hidden <= 64, 64-row tiles for hidden 128 (two 64 KB weight
// copies leave room for nothing larger).
//
// Three passes share the tile code (one `Cta` struct, used by the one-launch kernels below and by
// the persistent training loop in train_loop_general.cu):
//   FWD  rows -> mu -> loss statistics (baselines, logging passes)
//   BWD  + backward -> per-CTA gradient partial; with LOSS_GIVEN the per-row d loss / d mu is read
//        from `dmu_in` instead of a loss (the J^T half of a Gauss-Newton / Fisher-vector product)
//   JVP  forward-mode pass: d mu / d theta . v for a tangent v in the parameter layout, scaled by
//        1 / var -> `mean_out` (the J half of the Fisher-vector product of
//        torch/optimizers/conjugate_gradient_optimizer.py:18-66 for a diagonal Gaussian)
#pragma once
#include "mlp_train.cuh"

namespace gen {
constexpr int NT = 256;
enum : int { M_FWD = 0, M_BWD = 1, M_JVP = 2 };

template <int H>
struct Cfg {
    static constexpr int R = H <= 64 ? 128 : 64, RS = R + 4;
    static constexpr int TXN = H / 4, TYN = NT / TXN, RPT = R / TYN, KPT = H / TYN, WX = TXN / 8;
    static_assert(RPT % 4 == 0 && KPT >= 1 && WX >= 1, "tile shape");
};

// floats of shared memory: activations + weights of the pass
template <int H, int OP, int AP, int MODE>
constexpr size_t smem_floats() {
    using C = Cfg<H>;
    return (size_t)OP * C::RS + 2 * (size_t)H * C::RS + (size_t)AP * C::RS + (size_t)H * H +
           (MODE != M_FWD ? (size_t)H * H : 0) +                       // BWD: W2 column-permuted; JVP: tangent of W2
           (size_t)H * OP + 2 * H + (size_t)AP * H + AP +              // W1 b1 b2 W3 b3
           (MODE == M_JVP ? (size_t)H * OP + 2 * H + (size_t)AP * H + AP : 0) + 8;
}
template <int H, int OP, int AP, int MODE>
constexpr size_t smem_bytes() { return sizeof(float) * smem_floats<H, OP, AP, MODE>(); }

template <int H, int OP, int AP, int MODE>
struct Cta {
    using C = Cfg<H>;
    static constexpr int R = C::R, RS = C::RS, TXN = C::TXN, TYN = C::TYN, RPT = C::RPT, KPT = C::KPT, WX = C::WX;
    static constexpr bool BWD = MODE == M_BWD, JVP = MODE == M_JVP;
    float *sX, *sH1, *sH2, *sDmu, *sW2T, *sW2P, *sW1, *sB1, *sB2, *sW3, *sB3;
    float *sV1, *sVB1, *sVB2, *sV3, *sVB3;  // JVP: tangent of the small layers (sW2P holds the tangent of W2)
    int tid, w, l, tx, ty, r0, O, A;
    MlpDist dist;
    float tan_ls;  // JVP: tangent of log_std
    // gradient accumulators (BWD), persistent over the tiles of a CTA
    float gW2[4][KPT], gW3[AP][4], gB2[4], gW1[4][OP], gB1[4], gB3[AP];
    MlpRowAcc racc;

    __device__ __forceinline__ void init(float* smem, const MlpArgs& a) {
        sX = smem;
        sH1 = sX + OP * RS;
        sH2 = sH1 + H * RS;
        sDmu = sH2 + H * RS;
        sW2T = sDmu + AP * RS;
        sW2P = sW2T + H * H;
        sW1 = sW2P + (MODE != M_FWD ? H * H : 0);
        sB1 = sW1 + H * OP;
        sB2 = sB1 + H;
        sW3 = sB2 + H;
        sB3 = sW3 + AP * H;
        sV1 = sB3 + AP;
        sVB1 = sV1 + H * OP;
        sVB2 = sVB1 + H;
        sV3 = sVB2 + H;
        sVB3 = sV3 + AP * H;
        tid = threadIdx.x; w = tid >> 5; l = tid & 31;
        tx = (w % WX) * 8 + (l & 7);
        ty = (w / WX) * 4 + (l >> 3);
        r0 = ty * RPT;
        O = a.O_rt; A = a.A_rt;
        tan_ls = 0.0f;
    }

    // stage the parameters (and, for JVP, the tangent): W2 transposed and column-permuted so that a thread's
    // four interleaved columns c = tx + TXN i are adjacent; small layers padded with zeros up to OP / AP.
    // `ld` loads one parameter (plain or L2-coherent: the training loop rewrites them between steps)
    template <class Ld>
    __device__ __forceinline__ void load_weights(const MlpArgs& a, const float* params, Ld ld) {
        const MlpLayout L{O, H, A};
        for (int i = tid; i < H * H; i += NT) {
            const int c = i / H, k = i % H;
            const float v = ld(params + L.off_w2() + i);
            sW2T[k * H + (c % TXN) * 4 + (c / TXN)] = v;
            if (BWD) sW2P[c * H + (k % TXN) * 4 + (k / TXN)] = v;
            if (JVP) sW2P[k * H + (c % TXN) * 4 + (c / TXN)] = ld(a.vparams + L.off_w2() + i);
        }
        for (int i = tid; i < H * OP; i += NT) {
            const int c = i / OP, o = i % OP;
            sW1[i] = o < O ? ld(params + L.off_w1() + c * O + o) : 0.0f;
            if (JVP) sV1[i] = o < O ? ld(a.vparams + L.off_w1() + c * O + o) : 0.0f;
        }
        for (int i = tid; i < H; i += NT) {
            sB1[i] = ld(params + L.off_b1() + i);
            sB2[i] = ld(params + L.off_b2() + i);
            if (JVP) { sVB1[i] = ld(a.vparams + L.off_b1() + i); sVB2[i] = ld(a.vparams + L.off_b2() + i); }
        }
        for (int i = tid; i < AP * H; i += NT) {
            sW3[i] = i < A * H ? ld(params + L.off_w3() + i) : 0.0f;
            if (JVP) sV3[i] = i < A * H ? ld(a.vparams + L.off_w3() + i) : 0.0f;
        }
        if (tid < AP) {
            sB3[tid] = tid < A ? ld(params + L.off_b3() + tid) : 0.0f;
            if (JVP) sVB3[tid] = tid < A ? ld(a.vparams + L.off_b3() + tid) : 0.0f;
        }
        dist = mlp_dist(a, ld(params + L.off_logstd()));
        mlp_dist_entropy<AP>(a, dist, A);
        if (JVP) tan_ls = ld(a.vparams + L.off_logstd());
        __syncthreads();
    }

    __device__ __forceinline__ void zero_acc() {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
            for (int j = 0; j < KPT; ++j) gW2[i][j] = 0.0f;
#pragma unroll
            for (int q = 0; q < AP; ++q) gW3[q][i] = 0.0f;
#pragma unroll
            for (int o = 0; o < OP; ++o) gW1[i][o] = 0.0f;
            gB2[i] = 0.0f; gB1[i] = 0.0f;
        }
#pragma unroll
        for (int q = 0; q < AP; ++q) gB3[q] = 0.0f;
        racc = MlpRowAcc{};
    }

    // minibatch row id of row m, or -1 past the end: index list, on-the-fly Feistel permutation, or sequential
    static __device__ __forceinline__ int64_t mb_src(const MlpArgs& a, int64_t m) {
        if (m >= a.mb) return -1;
        if (a.idx) return (int64_t)a.idx[m];
        if (a.perm_n > 0) return (int64_t)b2g_perm_apply(a.perm_keys, a.perm_half_bits, (uint64_t)a.perm_n, (uint64_t)(a.start + m));
        return a.start + m;
    }

    // the R-row tiles tile0, tile0 + stride, ... of minibatch `a`
    __device__ __forceinline__ void run_tiles(const MlpArgs& a, int64_t tile0, int64_t stride) {
        const int64_t ntiles = (a.mb + R - 1) / R;
        for (int64_t tile = tile0; tile < ntiles; tile += stride) {
            const int64_t m0 = tile * R;
            // ---- 1. gather rows (threads 0..R-1 own one row each for the loss stage)
            int64_t src = -1;
            if (tid < R) {
                src = mb_src(a, m0 + tid);
#pragma unroll
                for (int o = 0; o < OP; ++o) sX[o * RS + tid] = (src >= 0 && o < O) ? a.obs[src * O + o] : 0.0f;
            }
            __syncthreads();
            // ---- 2. layer 1: h1 = tanh(W1 x + b1) -> sH1[c][r]   (JVP: dh1 = (1 - h1^2)(V1 x + vb1) -> sH2)
            {
                float xr[OP][RPT];
#pragma unroll
                for (int o = 0; o < OP; ++o)
#pragma unroll
                    for (int q = 0; q < RPT / 4; ++q) {
                        const float4 v = lds4(sX + o * RS + r0 + 4 * q);
                        xr[o][4 * q] = v.x; xr[o][4 * q + 1] = v.y; xr[o][4 * q + 2] = v.z; xr[o][4 * q + 3] = v.w;
                    }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int c = tx + TXN * i;
                    float h[RPT], dh[RPT];
#pragma unroll
                    for (int r = 0; r < RPT; ++r) {
                        float z = sB1[c], dz = JVP ? sVB1[c] : 0.0f;
#pragma unroll
                        for (int o = 0; o < OP; ++o) {
                            z = fmaf(sW1[c * OP + o], xr[o][r], z);
                            if (JVP) dz = fmaf(sV1[c * OP + o], xr[o][r], dz);
                        }
                        h[r] = tanhf(z);
                        dh[r] = (1.0f - h[r] * h[r]) * dz;
                    }
#pragma unroll
                    for (int q = 0; q < RPT / 4; ++q) {
                        sts4(sH1 + c * RS + r0 + 4 * q, make_float4(h[4 * q], h[4 * q + 1], h[4 * q + 2], h[4 * q + 3]));
                        if (JVP) sts4(sH2 + c * RS + r0 + 4 * q, make_float4(dh[4 * q], dh[4 * q + 1], dh[4 * q + 2], dh[4 * q + 3]));
                    }
                }
            }
            __syncthreads();
            // ---- 3. GEMM1: z2[r][c] = sum_k h1[r][k] W2[c][k]; h2 = tanh(z2 + b2) -> sH2[c][r]
            //      JVP: dz2 = sum_k dh1[r][k] W2[c][k] + h1[r][k] V2[c][k] + vb2; dh2 = (1 - h2^2) dz2 -> sH1
            {
                float acc[RPT][4], acc2[JVP ? RPT : 1][4];
#pragma unroll
                for (int r = 0; r < RPT; ++r)
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        acc[r][i] = 0.0f;
                        if (JVP) acc2[r][i] = 0.0f;
                    }
#pragma unroll 4
                for (int k = 0; k < H; ++k) {
                    float av[RPT], dv[JVP ? RPT : 1];
#pragma unroll
                    for (int q = 0; q < RPT / 4; ++q) {
                        const float4 v = lds4(sH1 + k * RS + r0 + 4 * q);
                        av[4 * q] = v.x; av[4 * q + 1] = v.y; av[4 * q + 2] = v.z; av[4 * q + 3] = v.w;
                        if (JVP) {
                            const float4 d = lds4(sH2 + k * RS + r0 + 4 * q);
                            dv[4 * q] = d.x; dv[4 * q + 1] = d.y; dv[4 * q + 2] = d.z; dv[4 * q + 3] = d.w;
                        }
                    }
                    const float4 b = lds4(sW2T + k * H + tx * 4);
                    const float bb[4] = {b.x, b.y, b.z, b.w};
                    float vv[4] = {0.0f, 0.0f, 0.0f, 0.0f};
                    if (JVP) {
                        const float4 t = lds4(sW2P + k * H + tx * 4);
                        vv[0] = t.x; vv[1] = t.y; vv[2] = t.z; vv[3] = t.w;
                    }
#pragma unroll
                    for (int r = 0; r < RPT; ++r)
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            acc[r][i] = fmaf(av[r], bb[i], acc[r][i]);
                            if (JVP) acc2[r][i] = fmaf(dv[r], bb[i], fmaf(av[r], vv[i], acc2[r][i]));
                        }
                }
                if (JVP) __syncthreads();  // every thread has read h1 / dh1: both buffers are rewritten below
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int c = tx + TXN * i;
                    const float bb = sB2[c], vb = JVP ? sVB2[c] : 0.0f;
                    float h[RPT], dh[RPT];
#pragma unroll
                    for (int r = 0; r < RPT; ++r) {
                        h[r] = tanhf(acc[r][i] + bb);
                        dh[r] = JVP ? (1.0f - h[r] * h[r]) * (acc2[r][i] + vb) : 0.0f;
                    }
#pragma unroll
                    for (int q = 0; q < RPT / 4; ++q) {
                        sts4(sH2 + c * RS + r0 + 4 * q, make_float4(h[4 * q], h[4 * q + 1], h[4 * q + 2], h[4 * q + 3]));
                        if (JVP) sts4(sH1 + c * RS + r0 + 4 * q, make_float4(dh[4 * q], dh[4 * q + 1], dh[4 * q + 2], dh[4 * q + 3]));
                    }
                }
            }
            __syncthreads();
            // ---- 4. layer 3 + loss, one thread per row: mu = W3 h2 + b3
            if (tid < R) {
                float mu[AP], dmu[AP];
#pragma unroll
                for (int q = 0; q < AP; ++q) { mu[q] = JVP ? sVB3[q] : sB3[q]; dmu[q] = 0.0f; }
#pragma unroll 8
                for (int c = 0; c < H; ++c) {
                    const float h = sH2[c * RS + tid];
                    if (JVP) {  // d mu = W3 dh2 + V3 h2 + vb3
                        const float dh = sH1[c * RS + tid];
#pragma unroll
                        for (int q = 0; q < AP; ++q) mu[q] = fmaf(sW3[q * H + c], dh, fmaf(sV3[q * H + c], h, mu[q]));
                    } else {
#pragma unroll
                        for (int q = 0; q < AP; ++q) mu[q] = fmaf(sW3[q * H + c], h, mu[q]);
                    }
                }
                if (src >= 0) {
                    if (JVP) {
#pragma unroll
                        for (int q = 0; q < AP; ++q)
                            if (q < A) a.mean_out[(m0 + tid) * A + q] = mu[q] * dist.inv_var;
                    } else {
                        if (a.mean_out) {
#pragma unroll
                            for (int q = 0; q < AP; ++q)
                                if (q < A) a.mean_out[(m0 + tid) * A + q] = mu[q];
                        }
                        MlpRowIn<AP> in;
                        in.adv = 0.0f; in.old_logp = 0.0f; in.ret = 0.0f;
#pragma unroll
                        for (int q = 0; q < AP; ++q) { in.act[q] = 0.0f; in.old_mean[q] = 0.0f; }
                        if (a.loss_kind == LOSS_VF) {
                            in.ret = a.ret[src];
                        } else if (a.loss_kind == LOSS_GIVEN) {
#pragma unroll
                            for (int q = 0; q < AP; ++q)
                                if (q < A) dmu[q] = a.dmu_in[(m0 + tid) * A + q];
                        } else if (a.loss_kind != LOSS_NONE) {
#pragma unroll
                            for (int q = 0; q < AP; ++q)
                                if (q < A) {
                                    in.act[q] = a.act[src * A + q];
                                    if (a.old_mean) in.old_mean[q] = a.old_mean[src * A + q];
                                }
                            in.adv = a.adv[src];
                            if (a.loss_kind == LOSS_PPO) in.old_logp = a.old_logp[src];
                        }
                        mlp_row_loss<AP>(a, in, mu, dist, dmu, racc, A);
                    }
                }
                if (BWD) {
#pragma unroll
                    for (int q = 0; q < AP; ++q) {
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
            backward_tile();
            __syncthreads();
        }
    }

    // ---- 5./6. of one tile: dz2, dW3, db2; dW2 += dz2^T h1; dh1 = dz2 W2; dz1, dW1, db1
    __device__ __forceinline__ void backward_tile() {
        {
            float dm[AP][RPT];
#pragma unroll
            for (int q = 0; q < AP; ++q)
#pragma unroll
                for (int p = 0; p < RPT / 4; ++p) {
                    const float4 v = lds4(sDmu + q * RS + r0 + 4 * p);
                    dm[q][4 * p] = v.x; dm[q][4 * p + 1] = v.y; dm[q][4 * p + 2] = v.z; dm[q][4 * p + 3] = v.w;
                }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int c = tx + TXN * i;
                float w3c[AP];
#pragma unroll
                for (int q = 0; q < AP; ++q) w3c[q] = sW3[q * H + c];
#pragma unroll
                for (int p = 0; p < RPT / 4; ++p) {
                    const float4 hv = lds4(sH2 + c * RS + r0 + 4 * p);
                    const float h[4] = {hv.x, hv.y, hv.z, hv.w};
                    float dz[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float dh = 0.0f;
#pragma unroll
                        for (int q = 0; q < AP; ++q) {
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
        // dW2[c][k] += sum_r dz2[r][c] h1[r][k]   (c = tx + TXN i, k = ty + TYN j)
#pragma unroll 2
        for (int r = 0; r < R; r += 4) {
            float4 d[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) d[i] = lds4(sH2 + (tx + TXN * i) * RS + r);
#pragma unroll
            for (int j = 0; j < KPT; ++j) {
                const float4 hh = lds4(sH1 + (ty + TYN * j) * RS + r);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    gW2[i][j] = fmaf(d[i].x, hh.x, gW2[i][j]);
                    gW2[i][j] = fmaf(d[i].y, hh.y, gW2[i][j]);
                    gW2[i][j] = fmaf(d[i].z, hh.z, gW2[i][j]);
                    gW2[i][j] = fmaf(d[i].w, hh.w, gW2[i][j]);
                }
            }
        }
        // dh1[r][k] = sum_c dz2[r][c] W2[c][k]; dz1 = dh1 (1 - h1^2); dW1, db1
        {
            float acc[RPT][4];
#pragma unroll
            for (int r = 0; r < RPT; ++r)
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[r][i] = 0.0f;
#pragma unroll 4
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
            float xr[OP][RPT];
#pragma unroll
            for (int o = 0; o < OP; ++o)
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
                        for (int o = 0; o < OP; ++o) gW1[i][o] = fmaf(dz1, xr[o][4 * p + e], gW1[i][o]);
                    }
                }
            }
        }
    }

    // this CTA's partial row [P gradients | MLP_NSTATS statistics]; afterwards the accumulators are zero again.
    // hvp_ls: (JVP / LOSS_GIVEN callers) nothing special -- the log_std entry of a Fisher-vector product is added
    // by the host-side finishing kernel.
    __device__ __forceinline__ void write_partial(float* out) {
        const MlpLayout L{O, H, A};
        const int P = L.total();
        float* red = sH1;  // scratch (sH1 and sH2 are contiguous, 2*H*RS floats)
        if (BWD) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < KPT; ++j)
                    out[L.off_w2() + (tx + TXN * i) * H + (ty + TYN * j)] = gW2[i][j];
            // column-owned sums are replicated over the TYN row groups: the four groups inside a warp
            // (lanes l, l^8, l^16, l^24 share tx) fold with shuffles, the TYN / 4 warps per column
            // group through shared memory, both in a fixed order
            constexpr int VV = AP + OP + 2, V = 4 * VV, YW = TYN / 4;
            static_assert((size_t)(NT / 4) * V <= 2 * (size_t)H * RS, "reduction scratch too small");
            float vals[V];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
#pragma unroll
                for (int q = 0; q < AP; ++q) vals[i * VV + q] = gW3[q][i];
                vals[i * VV + AP] = gB2[i];
#pragma unroll
                for (int o = 0; o < OP; ++o) vals[i * VV + AP + 1 + o] = gW1[i][o];
                vals[i * VV + AP + OP + 1] = gB1[i];
            }
            __syncthreads();  // the tile loop's last reads of sH1 / sH2 are done (also when this CTA had no tile)
#pragma unroll
            for (int v = 0; v < V; ++v) {
                float x = vals[v];
                x += __shfl_xor_sync(0xffffffffu, x, 8);
                x += __shfl_xor_sync(0xffffffffu, x, 16);
                if ((l >> 3) == 0) red[((w / WX) * TXN + tx) * V + v] = x;
            }
            __syncthreads();
            for (int e = tid; e < TXN * V; e += NT) {
                const int txp = e / V, v = e % V, i = v / VV, s = v % VV;
                float sum = 0.0f;
                for (int y = 0; y < YW; ++y) sum += red[(y * TXN + txp) * V + v];
                const int c = txp + TXN * i;
                int dst = -1;
                if (s < AP) { if (s < A) dst = L.off_w3() + s * H + c; }
                else if (s == AP) dst = L.off_b2() + c;
                else if (s <= AP + OP) { if (s - AP - 1 < O) dst = L.off_w1() + c * O + (s - AP - 1); }
                else dst = L.off_b1() + c;
                if (dst >= 0) out[dst] = sum;
            }
            __syncthreads();
        }
        // row-owned scalars (threads 0..R-1): db3, dlog_std, statistics
        {
            constexpr int NS = AP + 1 + 5;
            float vals[NS];
#pragma unroll
            for (int q = 0; q < AP; ++q) vals[q] = gB3[q];
            vals[AP] = racc.gLs; vals[AP + 1] = racc.st_loss; vals[AP + 2] = racc.st_kl; vals[AP + 3] = racc.st_ent;
            vals[AP + 4] = racc.st_clip; vals[AP + 5] = racc.st_rows;
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
                for (int y = 0; y < NT / 32; ++y) sum += red[tid * 8 + y];
                if (tid < AP) { if (BWD && tid < A) out[L.off_b3() + tid] = sum; }
                else if (tid == AP) { if (BWD) out[L.off_logstd()] = sum; }
                else out[P + (tid - AP - 1)] = sum;
            }
            if (tid >= NS && tid < NS + (MLP_NSTATS - 5)) out[P + 5 + (tid - NS)] = 0.0f;
            __syncthreads();
        }
        zero_acc();
    }
};

// one launch over one minibatch: grid-stride over the tiles, one partial row per CTA
template <int H, int OP, int AP, int MODE>
__global__ void __launch_bounds__(NT, (H <= 64 && MODE != M_JVP) ? 2 : 1) mlp_general_kernel(const MlpArgs a) {
    extern __shared__ __align__(16) float gen_smem[];
    Cta<H, OP, AP, MODE> cta;
    cta.init(gen_smem, a);
    if (MODE == M_BWD && a.step_counter && blockIdx.x == 0 && threadIdx.x == 0) *a.step_counter += 1ull;
    cta.load_weights(a, a.params, [](const float* q) { return __ldg(q); });
    cta.zero_acc();
    cta.run_tiles(a, blockIdx.x, gridDim.x);
    if (MODE != M_JVP) {
        const int W = MlpLayout{a.O_rt, H, a.A_rt}.total() + MLP_NSTATS;
        cta.write_partial(a.partial + (int64_t)blockIdx.x * W);
    }
}
}  // namespace gen
