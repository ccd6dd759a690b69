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

enum : int { LOSS_PPO = 0, LOSS_VPG = 1, LOSS_VF = 2 };
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
    float old_log_std;
    int loss_kind, flags;
    float clip, ent_coeff, min_log_std, max_log_std;
    float* partial;         // [grid][P + MLP_NSTATS]
    float* mean_out;        // [mb,A] or NULL (forward output, minibatch order)
    unsigned long long* step_counter;  // Adam step, ticked once per backward launch (or NULL)
};

template <int H>
struct MlpSmem {
    static constexpr int TXN = H / 4, TYN = MLP_NT / TXN, RPT = MLP_R / TYN, KPT = H / TYN;
};

__device__ __forceinline__ float4 lds4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void sts4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

template <int H, int O, int A, bool BWD>
__global__ void __launch_bounds__(MLP_NT, 1) mlp_train_kernel(const MlpArgs a) {
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
    const float raw_ls = a.params[L.off_logstd()];
    const float log_std = fminf(fmaxf(raw_ls, a.min_log_std), a.max_log_std);
    const bool ls_pass = (raw_ls >= a.min_log_std) && (raw_ls <= a.max_log_std);  // clamp backward mask
    const float stdv = expf(log_std), var = stdv * stdv, log_scale = logf(stdv);
    if (BWD && a.step_counter && blockIdx.x == 0 && tid == 0) *a.step_counter += 1ull;
    __syncthreads();

    // ---- persistent gradient accumulators
    float gW2[4][KPT], gW3[A][4], gB2[4], gW1[4][O], gB1[4];
    float gB3[A], gLs = 0.0f, st_loss = 0.0f, st_kl = 0.0f, st_ent = 0.0f, st_clip = 0.0f, st_rows = 0.0f;
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
