// Persistent training loop on the general-shape CUDA-core tiles (mlp_general.cuh) for ONE hidden
// size (GEN_H, defined by the including .cu): every minibatch step of `_train`
// (torch/algos/vpg.py:230-293 -- `for epoch: for minibatch: zero_grad; loss; backward;
// optimizer.step()`), for the policy and for the value function, in one launch.
//
// Same skeleton as train_loop.cu (tcgen05 tiles, hidden 64): G co-resident CTAs per network; per step
// each CTA runs its R-row tiles, writes its partial gradient row, meets the others at a device-wide
// barrier, reduces its slice of the parameters over the G partial rows in a fixed order, applies
// torch.optim.Adam's update, and after a second barrier reloads the parameters.  G follows the
// minibatch: a 64-row minibatch (the reference's default minibatch_size, ppo.py:65-76) is ONE tile, so
// G = 1 and both barriers are trivial -- C1's 160 + 160 minibatch steps per iteration cost one launch
// instead of two launches each.
#include "mlp_general.cuh"
#include "train_loop_common.cuh"

namespace {
template <int OP, int AP>
__device__ __forceinline__ void gen_net_loop(const NetLoop& n, int j, int G, float* smem) {
    constexpr int H = GEN_H, NT = gen::NT;
    MlpArgs a = n.a;
    a.params = n.params;
    a.step_counter = nullptr;
    gen::Cta<H, OP, AP, gen::M_BWD> cta;
    cta.init(smem, a);
    const MlpLayout L{a.O_rt, H, a.A_rt};
    const int P = L.total(), W = P + MLP_NSTATS;
    const int tid = threadIdx.x;
    auto ld = [](const float* q) { return __ldcg(q); };  // parameters are rewritten between steps: L2 only
    unsigned long long t = *reinterpret_cast<volatile unsigned long long*>(n.step);
    const unsigned long long t_begin = t;
    unsigned int nbar = 0;
    float* out = n.partial + (int64_t)j * W;
    __shared__ float adam_c[2];
    __shared__ float red[8][32];
    const int64_t steps = (n.B + n.mbs - 1) / n.mbs;
    const int64_t total = steps * n.epochs;
    const int perm_half_bits = n.shuffle == 1 ? b2g_perm_half_bits(n.B) : 0;
    double b1t = 1.0, b2t = 1.0;  // beta1^t, beta2^t (thread 64 only)
    if (tid == 64) { b1t = pow(n.beta1, (double)t); b2t = pow(n.beta2, (double)t); }
    const int ppc = (P + G - 1) / G;  // parameters reduced + updated by this CTA
    const int px = tid & 31, sl = tid >> 5;
    cta.load_weights(a, n.params, ld);
    cta.zero_acc();
    int64_t e = 0, s = 0;
    for (int64_t g = 0; g < total; ++g) {
        // minibatch g: its rows (index list, on-the-fly permutation keyed by the Adam step at the epoch's
        // start -- what b2g_permutation(seed, counter_dev = step) reads --, or sequential)
        a.start = s * n.mbs;
        a.mb = (n.B - a.start) < n.mbs ? (n.B - a.start) : n.mbs;
        a.idx = (n.shuffle == 2) ? n.idx_all + e * n.B + a.start : nullptr;
        a.perm_n = 0;
        if (n.shuffle == 1) {
            a.perm_n = n.B;
            a.perm_half_bits = perm_half_bits;
            if (s == 0) b2g_perm_keys(n.seed_lo, n.seed_hi, (uint64_t)(t_begin + e * steps), a.perm_keys);
        }
        if (++s == steps) { s = 0; ++e; }
        ++t;
        if (tid == 64) {  // Adam bias corrections of this step (running products: one pow per launch)
            b1t *= n.beta1;
            b2t *= n.beta2;
            adam_c[0] = (float)(n.lr / (1.0 - b1t));
            adam_c[1] = (float)sqrt(1.0 - b2t);
        }
        const float inv_mb = 1.0f / (float)a.mb;
        cta.run_tiles(a, j, G);
        cta.write_partial(out);
        if (G > 1) net_barrier(n.sync, ++nbar * (unsigned int)G); else __syncthreads();
        // ---- fixed-order reduction of this CTA's parameter slice over the G partial rows + Adam
        auto adam = [&](int pidx, float tot) {
            const float gr = tot * inv_mb;
            const float step_size = adam_c[0], bc2_sqrt = adam_c[1];
            float mm = n.adam_m[pidx], vv = n.adam_v[pidx];
            const float pp = __ldcg(n.params + pidx);
            mm = mm + (float)(1.0 - n.beta1) * (gr - mm);                   // exp_avg.lerp_(grad, 1-beta1)
            vv = vv * (float)n.beta2 + ((float)(1.0 - n.beta2) * gr) * gr;  // mul_(beta2).addcmul_(g, g, 1-beta2)
            const float denom = sqrtf(vv) / bc2_sqrt + (float)n.eps;
            n.params[pidx] = pp - step_size * (mm / denom);
            n.adam_m[pidx] = mm;
            n.adam_v[pidx] = vv;
        };
        if (G <= 8) {
            // few partial rows (small minibatches: G = 1 for the reference's 64-row default): every thread owns
            // parameters and adds the rows itself, no staging
            for (int q = tid; q < ppc; q += NT) {
                const int pidx = j * ppc + q;
                if (pidx < P) {
                    float tot = 0.0f;
                    for (int b = 0; b < G; ++b) tot += __ldcg(n.partial + (int64_t)b * W + pidx);
                    adam(pidx, tot);
                }
            }
        } else {
            // many rows: 8 warps split the rows of 32 parameters at a time, partial sums meet in shared memory
            for (int c0 = 0; c0 < ppc; c0 += 32) {
                const int pidx = j * ppc + c0 + px;
                const bool own = c0 + px < ppc && pidx < P;
                float sum = 0.0f;
                if (own)
                    for (int b = sl; b < G; b += 8) sum += __ldcg(n.partial + (int64_t)b * W + pidx);
                red[sl][px] = sum;
                __syncthreads();
                if (sl == 0 && own) {
                    float tot = 0.0f;
#pragma unroll
                    for (int k = 0; k < 8; ++k) tot += red[k][px];
                    adam(pidx, tot);
                }
                __syncthreads();
            }
        }
        if (G > 1) net_barrier(n.sync, ++nbar * (unsigned int)G); else __syncthreads();
        cta.load_weights(a, n.params, ld);
    }
    if (j == 0 && tid == 0) *n.step = t;
}

template <int OP, int AP>
__global__ void __launch_bounds__(gen::NT, GEN_H <= 64 ? 2 : 1) train_loop_general_kernel(const __grid_constant__ LoopArgs args) {
    extern __shared__ __align__(16) float gen_loop_smem[];
    const int net = blockIdx.x / args.ctas_per_net, j = blockIdx.x % args.ctas_per_net;
    // network 0 with a policy loss has up to AP action dimensions, every other network is a value function
    if (AP > 1 && net == 0 && args.net[0].a.loss_kind != LOSS_VF) gen_net_loop<OP, AP>(args.net[0], j, args.ctas_per_net, gen_loop_smem);
    else gen_net_loop<OP, 1>(args.net[net], j, args.ctas_per_net, gen_loop_smem);
}

template <int OP, int AP>
int gen_loop_max_ctas() {
    static int cap = 0;
    if (cap == 0) {
        constexpr size_t smem = gen::smem_bytes<GEN_H, OP, AP, gen::M_BWD>();
        cudaFuncSetAttribute(train_loop_general_kernel<OP, AP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, train_loop_general_kernel<OP, AP>, gen::NT, smem) != cudaSuccess || n < 1)
            n = 1;
        cap = n * b2g_sm_count();
    }
    return cap;
}
template <int OP, int AP>
int gen_loop_launch(LoopArgs& args, int64_t max_mb, cudaStream_t s) {
    constexpr size_t smem = gen::smem_bytes<GEN_H, OP, AP, gen::M_BWD>();
    constexpr int R = gen::Cfg<GEN_H>::R;
    const int cap = gen_loop_max_ctas<OP, AP>() / args.n_nets;  // every CTA must be resident: they wait for each other
    const int64_t tiles = (max_mb + R - 1) / R;
    const int G = (int)(tiles < cap ? (tiles > 0 ? tiles : 1) : cap);
    args.ctas_per_net = G;
    float* base = args.net[0].partial;
    int64_t off = 0;
    for (int i = 0; i < args.n_nets; ++i) {  // [G][P_i + MLP_NSTATS] partial rows per network
        args.net[i].partial = base + off;
        off += (int64_t)G * (MlpLayout{args.net[i].a.O_rt, GEN_H, args.net[i].a.A_rt}.total() + MLP_NSTATS);
    }
    train_loop_general_kernel<OP, AP><<<G * args.n_nets, gen::NT, smem, s>>>(args);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
}  // namespace

#define GEN_CAT2(a, b) a##b
#define GEN_CAT(a, b) GEN_CAT2(a, b)
// op / ap: capacities of network 0 (the caller checked the shapes); max_mb: largest minibatch of the networks
int GEN_CAT(b2g_gen_loop_h, GEN_H)(int op, int ap, LoopArgs& args, int64_t max_mb, cudaStream_t s) {
#define GEN_CASE(OO, AA) if (op == OO && ap == AA) return gen_loop_launch<OO, AA>(args, max_mb, s);
    GEN_CASE(4, 1) GEN_CASE(4, 2) GEN_CASE(4, 4) GEN_CASE(8, 1) GEN_CASE(8, 2) GEN_CASE(8, 4)
#undef GEN_CASE
    b2g_set_error("general training loop: capacity (%d, %d) is not built", op, ap);
    return B2G_ERR_UNSUPPORTED;
}
