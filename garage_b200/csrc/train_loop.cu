// Persistent training loop: ALL minibatch steps of `_train` (torch/algos/vpg.py:230-293 --
// `for epoch: for minibatch: zero_grad; loss; backward; optimizer.step()` for the policy and
// for the value function) in ONE cooperative launch.
//
// Grid = 2 CTAs per SM.  With two networks, CTAs [0, n) train the policy and CTAs [n, 2n) the
// value function (one CTA of each per SM: while one waits for its MMAs the other runs its
// CUDA-core stage); the two loops are independent, as in the reference, and never synchronise
// with each other.  Per minibatch step every CTA of a network
//   1. runs its 64-row tiles through the tcgen05 forward/loss/backward (mlp_tc.cuh) and writes
//      its partial gradient vector,
//   2. meets the other CTAs of the network at a device-wide barrier (atomic counter),
//   3. reduces its slice of the parameters over all partials in a fixed order, applies
//      torch.optim.Adam's update (same arithmetic as reduce_adam_kernel, mlp.cu) and
//   4. after a second barrier reloads the updated parameters (W2 re-split into tf32 panels).
// The minibatch shuffle is evaluated on the fly (b2g_perm_apply) or read from the supplied
// index lists; weights, Adam moments and the step counter are updated in place.
#include "mlp_tc.cuh"
#include "train_loop_common.cuh"
#include <stdlib.h>
#include <string.h>

constexpr int COMM_P = 4608;  // per-network capacity of the exchange area (parameters)

// One exchange slot = {value bits, step sequence number} packed into ONE 64-bit word and moved with a
// single scalar 8-byte store / load (the PTX memory model treats the halves of a .v2 access as two
// independent accesses in unspecified order; a scalar b64 access is single-copy atomic), so the
// sequence number doubles as the "data ready" flag and no system-scope fence (a multi-microsecond
// drain of every outstanding NVLink write) is needed.
__device__ __forceinline__ void st_slot_bits(unsigned long long* p, unsigned int bits, unsigned int seq) {
    const unsigned long long word = ((unsigned long long)seq << 32) | (unsigned long long)bits;
    asm volatile("st.relaxed.sys.global.b64 [%0], %1;" ::"l"(p), "l"(word) : "memory");
}
__device__ __forceinline__ void st_slot(unsigned long long* p, float v, unsigned int seq) {
    st_slot_bits(p, __float_as_uint(v), seq);
}
__device__ __forceinline__ unsigned long long ld_slot(const unsigned long long* p) {
    unsigned long long word;
    asm volatile("ld.relaxed.sys.global.b64 %0, [%1];" : "=l"(word) : "l"(p) : "memory");
    return word;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// A peer that never answers must surface as an error, not as a hung GPU -- but ordinary host skew
// between ranks (a snapshot on rank 0, first-call IPC mapping, a slow log disk) must not kill the
// context: the limit is wall time (globaltimer), 120 s unless B2G_PEER_TIMEOUT_S says otherwise.
__device__ unsigned long long g_peer_timeout_ns = 120ull * 1000000000ull;

// All-reduce of one CTA's parameter slice over the ranks, inside the kernel (warp 0 of the CTA):
// every rank stores {slice sum, seq} into every other rank's receive area over NVLink (peer-mapped
// memory), polls its own area until the slots of all source ranks carry this step's seq, and adds
// the contributions in rank order -- every rank obtains the same bits.  All polls are issued
// together (one load per source rank in flight, re-issued only for the ranks still missing), so the
// wait is one NVLink latency plus the slowest rank's skew instead of a chain of serial round trips.
// The receive area is double-buffered on the step parity: a slot of step s is rewritten at step
// s+2, which its writer can only reach after the reader's whole network passed the barrier of step
// s+1, i.e. after the reader consumed step s.
// Two halves, so that work which does not depend on the peers (the row ids and the gather of the next
// minibatch) runs while the stores cross NVLink.
__device__ __forceinline__ void slice_send(const NetLoop& n, int pidx, float tot, unsigned int seq) {
    const int par = (int)(seq & 1u);
    for (int r = 0; r < n.world; ++r)
        if (r != n.rank) st_slot(n.peer_recv[r] + ((size_t)(n.rank * 2 + par) * COMM_P + pidx), tot, seq);
}
__device__ __forceinline__ float slice_recv(const NetLoop& n, int pidx, float tot, unsigned int seq) {
    const int par = (int)(seq & 1u);
    float v[B2G_MAX_RANKS];
    unsigned int missing = 0;
#pragma unroll
    for (int r = 0; r < B2G_MAX_RANKS; ++r) {
        v[r] = (r == n.rank) ? tot : 0.0f;
        if (r < n.world && r != n.rank) missing |= 1u << r;
    }
    const unsigned long long* base = n.peer_recv[n.rank] + ((size_t)par * COMM_P + pidx);
    unsigned long long t0 = 0;
    for (unsigned int spin = 0; missing; ++spin) {
        unsigned long long wd[B2G_MAX_RANKS];
#pragma unroll
        for (int r = 0; r < B2G_MAX_RANKS; ++r)  // every outstanding poll in flight before any is examined
            wd[r] = (missing & (1u << r)) ? ld_slot(base + (size_t)r * 2 * COMM_P) : 0ull;
#pragma unroll
        for (int r = 0; r < B2G_MAX_RANKS; ++r)
            if ((missing & (1u << r)) && (unsigned int)(wd[r] >> 32) == seq) {
                v[r] = __uint_as_float((unsigned int)wd[r]);
                missing &= ~(1u << r);
            }
        if ((spin & 0xfffu) == 0xfffu) {
            const unsigned long long now = global_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > g_peer_timeout_ns) __trap();
        }
    }
    float sum = 0.0f;
#pragma unroll
    for (int r = 0; r < B2G_MAX_RANKS; ++r)
        if (r < n.world) sum += v[r];
    return sum;
}

template <int O, int A>
__device__ __forceinline__ void net_loop(const NetLoop& n, int j, int ctas, uint8_t* smem) {
    using namespace tc;
    const MlpLayout L{O, H, A};
    const int P = L.total(), W = P + MLP_NSTATS;
    Cta<O, A> cta;
    cta.init(smem);
    const int tid = threadIdx.x;
    MlpArgs a = n.a;
    a.params = n.params;
    a.step_counter = nullptr;
    a.timing = nullptr;
    auto ld = [](const float* q) { return __ldcg(q); };  // parameters are rewritten by other CTAs: L2 only
    unsigned long long t = *reinterpret_cast<volatile unsigned long long*>(n.step);
    unsigned int nbar = 0;
    // flags only ever grow: the sequence number continues across launches (kept on the device so that a
    // captured CUDA graph can be replayed)
    const unsigned int seq_base = n.world > 1 ? *reinterpret_cast<volatile unsigned int*>(n.seq_ptr) : 0u;
    const int ppc = (P + ctas - 1) / ctas;  // parameters reduced + updated by each CTA (<= 32)
    float* scr = reinterpret_cast<float*>(smem + OFF_ACT);
    float* out = n.partial + (int64_t)j * W;
    cta.zero_acc();
    const int64_t steps = (n.B + n.mbs - 1) / n.mbs;
    int64_t gstep = 0;
#define LOOP_STAMP(i) do { if (n.timing && tid == 0 && gstep == n.timing_step) n.timing[j * 16 + (i)] = clock64(); } while (0)
    // minibatch g (0-based, over all epochs) -> its arguments; the shuffle of an epoch is keyed by
    // the Adam step at the epoch's start (what b2g_permutation(seed, counter_dev = step) reads)
    const unsigned long long t_begin = t;
    // (epoch, step in epoch) of the minibatch prepared next: running counters, no 64-bit divisions
    const int perm_half_bits = n.shuffle == 1 ? b2g_perm_half_bits(n.B) : 0;
    int64_t ns_e = 0, ns_s = 0;
    auto set_next_step = [&](MlpArgs& m) {
        m.start = ns_s * n.mbs;
        m.mb = (n.B - m.start) < n.mbs ? (n.B - m.start) : n.mbs;
        m.idx = (n.shuffle == 2) ? n.idx_all + ns_e * n.B + m.start : nullptr;
        m.perm_n = 0;
        if (n.shuffle == 1) {
            m.perm_n = n.B;
            m.perm_half_bits = perm_half_bits;
            if (ns_s == 0) b2g_perm_keys(n.seed_lo, n.seed_hi, (uint64_t)(t_begin + ns_e * steps), m.perm_keys);
        }
        if (++ns_s == steps) { ns_s = 0; ++ns_e; }
    };
    const int64_t total = steps * n.epochs;
    double b1t = 1.0, b2t = 1.0;  // beta1^t, beta2^t (thread 64 only)
    if (tid == 64) { b1t = pow(n.beta1, (double)t); b2t = pow(n.beta2, (double)t); }
    const int px = tid & 31, sl = tid >> 5;
    const int pidx = j * ppc + px;
    const bool own = px < ppc && pidx < P;
    // gstep = -1 is the prologue (prime the first minibatch, load the weights): one call site each
    // for prime / run_tiles / write_partial / load_weights keeps the loop body small
    for (gstep = -1; gstep < total; ++gstep) {
        const bool live = gstep >= 0;
        if (live) {
            ++t;
            if (tid == 64) {  // Adam bias corrections of this step: beta^t as a running product (one pow per launch)
                b1t *= n.beta1;
                b2t *= n.beta2;
                const double bc1 = 1.0 - b1t, bc2 = 1.0 - b2t;
                float* adam_c = reinterpret_cast<float*>(smem + OFF_SMALL) + Small<O, A>::END;
                adam_c[0] = (float)(n.lr / bc1);
                adam_c[1] = (float)sqrt(bc2);
            }
            LOOP_STAMP(0);
            cta.stamps = (n.timing && gstep == n.timing_step) ? n.tile_timing + (int64_t)j * 32 : nullptr;
            cta.stamp(0);
            cta.template run_tiles<false>(a, j, ctas);
        }
        LOOP_STAMP(1);
        const float inv_mb = 1.0f / ((float)a.mb * (float)(n.world > 1 ? n.world : 1));  // the minibatch spans all ranks
        if (live) {  // publish this CTA's partial row first: the others wait for it, not for what follows
            cta.write_partial(out, false);
            cta.stamps = nullptr;
            LOOP_STAMP(2);
            net_barrier_arrive(n.sync);
        }
        const bool more = gstep + 1 < total;
        if (more) {  // rows of the next minibatch: independent of the update, gathered while the other CTAs
            set_next_step(a);  // finish their tiles (one GPU) or while the slice sums cross NVLink (several)
            LOOP_STAMP(8);
            if (n.world <= 1 || !live) cta.prime(a, j, ctas);  // (the prologue primes the first minibatch either way)
        }
        LOOP_STAMP(9);
        if (live) {
        // this CTA's slice of the optimiser state, fetched before the wait
        float mm = 0.0f, vv = 0.0f, pp = 0.0f;
        if (sl == 0 && own) {
            mm = n.adam_m[pidx];
            vv = n.adam_v[pidx];
            pp = __ldcg(n.params + pidx);
        }
        net_barrier_wait(n.sync, ++nbar * (unsigned int)ctas);
        LOOP_STAMP(3);
        // ---- fixed-order reduction of this CTA's parameter slice + Adam (reduce_adam_kernel's arithmetic)
        {
            // all loads in flight at once (the partials sit in the L2 of either die), summed in a fixed order
            constexpr int MAXB = 37;  // ceil(2 * 148 / 8)
            float pv[MAXB];
#pragma unroll
            for (int i = 0; i < MAXB; ++i) {
                const int b = sl + 8 * i;
                pv[i] = (own && b < ctas) ? __ldcg(n.partial + (int64_t)b * W + pidx) : 0.0f;
            }
            float sum = 0.0f;
#pragma unroll
            for (int i = 0; i < MAXB; ++i) sum += pv[i];
            scr[sl * 32 + px] = sum;
            __syncthreads();
            float tot = 0.0f;
            if (sl == 0 && own) {
#pragma unroll
                for (int k = 0; k < 8; ++k) tot += scr[k * 32 + px];
                if (n.world > 1) slice_send(n, pidx, tot, seq_base + (unsigned int)gstep + 1u);
            }
            if (n.world > 1 && more) cta.prime(a, j, ctas);
            if (sl == 0 && own) {
                if (n.world > 1) tot = slice_recv(n, pidx, tot, seq_base + (unsigned int)gstep + 1u);
                {
                    const float* adam_c = reinterpret_cast<const float*>(smem + OFF_SMALL) + Small<O, A>::END;
                    const float g = tot * inv_mb;
                    const float step_size = adam_c[0], bc2_sqrt = adam_c[1];
                    mm = mm + (float)(1.0 - n.beta1) * (g - mm);                 // exp_avg.lerp_(grad, 1-beta1)
                    vv = vv * (float)n.beta2 + ((float)(1.0 - n.beta2) * g) * g;  // mul_(beta2).addcmul_(g, g, 1-beta2)
                    const float denom = sqrtf(vv) / bc2_sqrt + (float)n.eps;
                    n.params[pidx] = pp - step_size * (mm / denom);
                    n.adam_m[pidx] = mm;
                    n.adam_v[pidx] = vv;
                }
            }
        }
        LOOP_STAMP(4);
        net_barrier(n.sync, ++nbar * (unsigned int)ctas);
        LOOP_STAMP(5);
        }
        cta.load_weights(a, n.params, ld);
        LOOP_STAMP(6);
        if (n.timing && tid == 0 && gstep == n.timing_step) {
            unsigned int smid;
            asm("mov.u32 %0, %%smid;" : "=r"(smid));
            n.timing[j * 16 + 7] = smid;
        }
    }
    if (j == 0 && tid == 0) {
        *n.step = t;
        if (n.world > 1) *n.seq_ptr = seq_base + (unsigned int)total;
    }
    cta.fini();
}

template <int O, int AP>
__global__ void __maxnreg__(120) train_loop_kernel(const __grid_constant__ LoopArgs args) {
    extern __shared__ __align__(1024) uint8_t loop_smem_raw[];
    uint8_t* smem = loop_smem_raw;
    const int net = blockIdx.x / args.ctas_per_net, j = blockIdx.x % args.ctas_per_net;
    if (net == 1 && args.stagger_cycles > 0) {
        const long long t0 = clock64();
        while (clock64() - t0 < args.stagger_cycles) __nanosleep(200);
    }
    if constexpr (AP == 1) {  // both networks run the SAME code: one copy in the instruction cache
        net_loop<O, 1>(args.net[net], j, args.ctas_per_net, smem);
    } else {
        if (net == 0 && args.net[0].a.loss_kind != LOSS_VF) net_loop<O, AP>(args.net[0], j, args.ctas_per_net, smem);
        else net_loop<O, 1>(args.net[net], j, args.ctas_per_net, smem);
    }
}

// ------------------------------------------------------------------ host --
template <int O, int AP>
static size_t loop_smem() {
    const size_t a = tc::smem_bytes<O, AP>(), b = tc::smem_bytes<O, 1>();
    return (a > b ? a : b) + 16;  // + Adam bias corrections
}

static int check_net(const b2g_train_net* t) {
    B2G_CHECK_ARG(t->hidden == 32 || t->hidden == 64 || t->hidden == 128, "the fused training loop is built for hidden in {32, 64, 128}");
    B2G_CHECK_ARG(t->obs_dim >= 1 && t->obs_dim <= 8 && t->out_dim >= 1 && t->out_dim <= 4, "obs_dim 1..8, out_dim 1..4");
    B2G_CHECK_ARG(t->params && t->adam_m && t->adam_v && t->step_counter && t->obs, "NULL pointer");
    B2G_CHECK_ARG(t->n_rows > 0 && t->epochs >= 0, "empty batch");
    B2G_CHECK_ARG(t->loss_kind >= LOSS_PPO && t->loss_kind <= LOSS_VF, "bad loss_kind");
    if (t->loss_kind == LOSS_VF) B2G_CHECK_ARG(t->ret && t->out_dim == 1, "value loss needs ret and out_dim 1");
    else B2G_CHECK_ARG(t->act && t->adv && (t->loss_kind != LOSS_PPO || t->old_logp), "policy loss needs act, adv (and old_logp for PPO)");
    B2G_CHECK_ARG(t->shuffle >= 0 && t->shuffle <= 2 && (t->shuffle != 2 || t->idx), "bad shuffle mode");
    B2G_CHECK_ARG(t->n_rows <= 0x7fffffff, "n_rows out of range for int32 ids");
    return B2G_OK;
}

static long long* g_loop_timing = nullptr;
static int g_loop_timing_step = 0;
// debug hook: device buffer [n_nets][ctas][16] of clock64 stamps of one minibatch step
extern "C" int b2g_mlp_train_loop_debug_timing(long long* buf, int step) {
    g_loop_timing = buf;
    g_loop_timing_step = step;
    return B2G_OK;
}

static void fill_net(NetLoop& n, const b2g_train_net* t, int ctas, float* partial, unsigned int* sync, int net_index) {
    n = NetLoop{};
    n.a.obs = t->obs; n.a.act = t->act; n.a.adv = t->adv; n.a.old_logp = t->old_logp; n.a.ret = t->ret;
    n.a.loss_kind = t->loss_kind; n.a.flags = t->flags; n.a.clip = t->clip; n.a.ent_coeff = t->ent_coeff;
    n.a.min_log_std = t->min_log_std; n.a.max_log_std = t->max_log_std;
    n.params = t->params; n.adam_m = t->adam_m; n.adam_v = t->adam_v; n.step = t->step_counter;
    n.lr = t->lr; n.beta1 = t->beta1; n.beta2 = t->beta2; n.eps = t->eps;
    n.B = t->n_rows;
    n.mbs = (t->minibatch > 0 && t->minibatch < t->n_rows) ? t->minibatch : t->n_rows;
    n.epochs = t->epochs; n.shuffle = t->shuffle; n.idx_all = t->idx;
    n.seed_lo = (uint32_t)t->shuffle_seed; n.seed_hi = (uint32_t)(t->shuffle_seed >> 32);
    n.partial = partial; n.sync = sync;
    (void)ctas;
    n.world = 1; n.rank = 0; n.seq_ptr = nullptr;
    n.timing = g_loop_timing ? g_loop_timing + (int64_t)net_index * ctas * 16 : nullptr;
    n.timing_step = g_loop_timing_step;
    n.tile_timing = g_loop_timing ? g_loop_timing + (int64_t)2 * 296 * 16 + (int64_t)net_index * ctas * 32 : nullptr;
}

extern "C" int64_t b2g_mlp_train_loop_ws_floats(void) {
    // (2 CTAs per SM) x (largest built parameter vector + statistics) + barrier counters
    const int64_t P = MlpLayout{8, 128, 4}.total() + MLP_NSTATS;
    return 2 * (int64_t)b2g_sm_count() * P + 64;
}

// general-shape loops (train_loop_general_h*.cu)
int b2g_gen_loop_h32(int op, int ap, LoopArgs& args, int64_t max_mb, cudaStream_t s);
int b2g_gen_loop_h64(int op, int ap, LoopArgs& args, int64_t max_mb, cudaStream_t s);
int b2g_gen_loop_h128(int op, int ap, LoopArgs& args, int64_t max_mb, cudaStream_t s);
extern "C" int b2g_get_mlp_path(void);
static bool tc_loop_built(int O, int A) { return (O == 3 && A == 1) || (O == 3 && A == 2) || (O == 4 && A == 1); }

// which loop runs these networks: 1 tcgen05 (hidden 64, built (obs, out) pairs, default MLP path), 2 general
// CUDA-core loop (one hidden size for all networks, single GPU), 0 neither
extern "C" int b2g_mlp_train_loop_kind(const b2g_train_net* nets, int n_nets, int world) {
    if (!nets || n_nets < 1 || n_nets > 2) return 0;
    bool tc = b2g_get_mlp_path() == 0;
    for (int i = 0; i < n_nets; ++i) tc = tc && nets[i].hidden == 64;
    tc = tc && tc_loop_built(nets[0].obs_dim, nets[0].out_dim) && (n_nets == 1 || nets[1].obs_dim == nets[0].obs_dim);
    if (tc) return 1;
    if (world > 1) return 0;
    for (int i = 0; i < n_nets; ++i) {
        const b2g_train_net& t = nets[i];
        if (!(t.hidden == 32 || t.hidden == 64 || t.hidden == 128) || t.hidden != nets[0].hidden) return 0;
        if (t.obs_dim < 1 || t.obs_dim > 8 || t.out_dim < 1 || t.out_dim > 4) return 0;
    }
    return 2;
}

template <int O, int AP>
static int launch_loop(const LoopArgs& args, int grid, cudaStream_t s) {
    const size_t smem = loop_smem<O, AP>();
    static bool configured = false;
    if (!configured) {
        B2G_CUDA(cudaFuncSetAttribute(train_loop_kernel<O, AP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        B2G_CUDA(cudaFuncSetAttribute(train_loop_kernel<O, AP>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                      cudaSharedmemCarveoutMaxShared));
        // The grid (2 CTAs per SM) must be co-resident: its CTAs wait for each other.  The runtime's
        // occupancy calculator answers 1 for every kernel that allocates tensor memory (so a
        // cooperative launch of this grid is refused), although two CTAs with 256 TMEM columns each
        // do share an SM; check the two budgets that decide it here instead.
        cudaFuncAttributes fa{};
        B2G_CUDA(cudaFuncGetAttributes(&fa, train_loop_kernel<O, AP>));
        int dev = 0, regs_sm = 0, smem_sm = 0;
        B2G_CUDA(cudaGetDevice(&dev));
        B2G_CUDA(cudaDeviceGetAttribute(&regs_sm, cudaDevAttrMaxRegistersPerMultiprocessor, dev));
        B2G_CUDA(cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev));
        const int regs_cta = ((fa.numRegs + 7) / 8 * 8) * tc::NT;
        if (2 * regs_cta > regs_sm - 4096 || 2 * (smem + 1024 + 128) > (size_t)smem_sm) {
            b2g_set_error("b2g_mlp_train_loop: two CTAs do not fit one SM (registers %d x 2 of %d, shared %zu x 2 of %d)",
                          regs_cta, regs_sm, smem + 1024, smem_sm);
            return B2G_ERR_UNSUPPORTED;
        }
        configured = true;
    }
    train_loop_kernel<O, AP><<<grid, tc::NT, smem, s>>>(args);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

constexpr size_t recv_bytes_c = (size_t)B2G_MAX_RANKS * 2 * COMM_P * 8;
constexpr int SMALL_N = 64;                                   // doubles per rank of b2g_peer_reduce_f64
constexpr size_t small_off_c = 2 * recv_bytes_c + 256;        // [src rank][parity][2 * SMALL_N] slots of 8 bytes
constexpr size_t small_bytes_c = (size_t)B2G_MAX_RANKS * 2 * (2 * SMALL_N) * 8;
extern "C" int64_t b2g_train_comm_bytes(void) {
    // per network: receive area [ranks][2][COMM_P] slots of 8 bytes; then 256 bytes of sequence counters
    // of this rank (two networks, small exchange); then the receive area of b2g_peer_reduce_f64
    return (int64_t)(small_off_c + small_bytes_c);
}

// ---- small all-gather + combine over the same peer-mapped buffers (advantage moments, logging scalars) ----
// One block per rank.  Thread i < 2n owns one 32-bit half of the rank's n doubles: it stores
// {half, seq} into every peer's receive area, polls its own area for the same half of every other
// rank (all polls in flight together) and leaves the gathered values in shared memory in rank order;
// thread 0 then combines them -- in rank order, so every rank computes identical bits.
struct PeerSmall {
    unsigned long long* recv[B2G_MAX_RANKS];
    unsigned int* seq_ptr;
    int world, rank, n, op;
    const double* in;
    double* out;
    double scalar0, scalar1;  // op 1: local element count; op 2: local sample count B, local episode count
};
__global__ void __launch_bounds__(2 * SMALL_N) peer_small_kernel(const PeerSmall q) {
    __shared__ double send[SMALL_N];
    __shared__ __align__(8) unsigned int halves[B2G_MAX_RANKS][2 * SMALL_N];
    const int i = threadIdx.x;
    if (i == 0) {
        if (q.op == B2G_PEER_MOMENTS) {  // [count, mean, sum of squared deviations, min]
            send[0] = q.scalar0; send[1] = q.in[0]; send[2] = q.in[1] * (q.scalar0 - 1.0); send[3] = q.in[2];
        } else if (q.op == B2G_PEER_LOGS) {
            const double ne = q.scalar1, mean = q.in[33], sd = q.in[34];
            for (int k = 0; k < 32; ++k) send[k] = q.in[k];
            send[32] = ne; send[33] = ne * q.in[32]; send[34] = ne * mean; send[35] = ne * (sd * sd + mean * mean);
            send[36] = ne * q.in[37]; send[37] = ne * q.in[38]; send[38] = q.in[39]; send[39] = q.scalar0;
            send[40] = q.in[35]; send[41] = -q.in[36];
        } else {
            for (int k = 0; k < q.n; ++k) send[k] = q.in[k];
        }
    }
    __syncthreads();
    const unsigned int seq = *reinterpret_cast<volatile unsigned int*>(q.seq_ptr) + 1u;
    const int par = (int)(seq & 1u);
    if (i < 2 * q.n) {
        const unsigned int mine = reinterpret_cast<const unsigned int*>(send)[i];
        halves[q.rank][i] = mine;
        for (int r = 0; r < q.world; ++r)
            if (r != q.rank) st_slot_bits(q.recv[r] + ((size_t)(q.rank * 2 + par) * (2 * SMALL_N) + i), mine, seq);
        unsigned int missing = 0;
        for (int r = 0; r < q.world; ++r)
            if (r != q.rank) missing |= 1u << r;
        const unsigned long long* base = q.recv[q.rank] + ((size_t)par * (2 * SMALL_N) + i);
        unsigned long long t0 = 0;
        for (unsigned int spin = 0; missing; ++spin) {
            unsigned long long wd[B2G_MAX_RANKS];
#pragma unroll
            for (int r = 0; r < B2G_MAX_RANKS; ++r)
                wd[r] = (missing & (1u << r)) ? ld_slot(base + (size_t)r * 2 * (2 * SMALL_N)) : 0ull;
#pragma unroll
            for (int r = 0; r < B2G_MAX_RANKS; ++r)
                if ((missing & (1u << r)) && (unsigned int)(wd[r] >> 32) == seq) {
                    halves[r][i] = (unsigned int)wd[r];
                    missing &= ~(1u << r);
                }
            if ((spin & 0xfffu) == 0xfffu) {
                const unsigned long long now = global_ns();
                if (t0 == 0) t0 = now;
                else if (now - t0 > g_peer_timeout_ns) __trap();
            }
        }
    }
    __syncthreads();
    if (i == 0) {
        *q.seq_ptr = seq;
        auto val = [&](int r, int k) { return reinterpret_cast<const double*>(halves[r])[k]; };
        if (q.op == B2G_PEER_MOMENTS) {
            // Chan et al.: mean and sum of squared deviations of the union from per-rank (n, mean, M2)
            double ntot = 0.0, wsum = 0.0, mn = val(0, 3);
            for (int r = 0; r < q.world; ++r) { ntot += val(r, 0); wsum += val(r, 0) * val(r, 1); mn = fmin(mn, val(r, 3)); }
            const double mean = wsum / ntot;
            double m2 = 0.0;
            for (int r = 0; r < q.world; ++r) { const double d = val(r, 1) - mean; m2 += val(r, 2) + val(r, 0) * d * d; }
            q.out[0] = mean; q.out[1] = m2 / (ntot - 1.0); q.out[2] = mn;
        } else if (q.op == B2G_PEER_LOGS) {
            double t[40], mx = val(0, 40), mnn = val(0, 41);
            for (int k = 0; k < 40; ++k) {
                double a = 0.0;
                for (int r = 0; r < q.world; ++r) a += val(r, k);
                t[k] = a;
            }
            for (int r = 1; r < q.world; ++r) { mx = fmax(mx, val(r, 40)); mnn = fmax(mnn, val(r, 41)); }
            const double nt = t[32], gmean = t[34] / nt;
            for (int k = 0; k < 32; ++k) q.out[k] = t[k];
            q.out[32] = t[33] / nt; q.out[33] = gmean; q.out[34] = sqrt(fmax(t[35] / nt - gmean * gmean, 0.0));
            q.out[35] = mx; q.out[36] = -mnn; q.out[37] = t[36] / nt; q.out[38] = t[37] / nt; q.out[39] = t[38];
            q.out[48] = t[39]; q.out[49] = nt; q.out[50] = 1.0;  // [40:48] (this rank's pad-KL statistics) stay
        } else {
            for (int r = 0; r < q.world; ++r)
                for (int k = 0; k < q.n; ++k) q.out[r * q.n + k] = val(r, k);
        }
    }
}

extern "C" int b2g_peer_reduce_f64(const b2g_train_comm* comm, int op, const double* in, int n, double scalar0,
                                   double scalar1, double* out, void* stream) {
    B2G_CHECK_ARG(comm && comm->world >= 1 && comm->world <= B2G_MAX_RANKS && comm->rank >= 0 && comm->rank < comm->world,
                  "bad world/rank");
    B2G_CHECK_ARG(in && out, "NULL pointer");
    B2G_CHECK_ARG(op >= B2G_PEER_GATHER && op <= B2G_PEER_LOGS, "bad op");
    PeerSmall q{};
    q.n = op == B2G_PEER_MOMENTS ? 4 : (op == B2G_PEER_LOGS ? 42 : n);
    B2G_CHECK_ARG(q.n > 0 && q.n <= SMALL_N, "1..64 doubles per rank");
    q.world = comm->world; q.rank = comm->rank; q.op = op; q.in = in; q.out = out;
    q.scalar0 = scalar0; q.scalar1 = scalar1;
    for (int r = 0; r < comm->world; ++r) {
        B2G_CHECK_ARG(comm->peer_base[r] != nullptr, "NULL peer buffer");
        q.recv[r] = reinterpret_cast<unsigned long long*>(reinterpret_cast<uint8_t*>(comm->peer_base[r]) + small_off_c);
    }
    q.seq_ptr = reinterpret_cast<unsigned int*>(reinterpret_cast<uint8_t*>(comm->peer_base[comm->rank]) + 2 * recv_bytes_c) + 32;
    peer_small_kernel<<<1, 2 * SMALL_N, 0, (cudaStream_t)stream>>>(q);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

extern "C" int b2g_mlp_train_loop(const b2g_train_net* nets, int n_nets, float* workspace,
                                  int64_t workspace_floats, void* stream) {
    return b2g_mlp_train_loop_dist(nets, n_nets, nullptr, workspace, workspace_floats, stream);
}

extern "C" int b2g_mlp_train_loop_dist(const b2g_train_net* nets, int n_nets, const b2g_train_comm* comm,
                                       float* workspace, int64_t workspace_floats, void* stream) {
    B2G_CHECK_ARG(nets && (n_nets == 1 || n_nets == 2), "one or two networks");
    B2G_CHECK_ARG(workspace && workspace_floats >= b2g_mlp_train_loop_ws_floats(), "workspace too small");
    for (int i = 0; i < n_nets; ++i) {
        const int rc = check_net(&nets[i]);
        if (rc != B2G_OK) return rc;
    }
    B2G_CHECK_ARG(n_nets == 1 || nets[0].obs_dim == nets[1].obs_dim, "both networks read the same observations");
    B2G_CHECK_ARG(n_nets == 1 || nets[1].out_dim == 1, "the second network is the value function (out_dim 1)");
    if (comm && comm->world > 1) {
        B2G_CHECK_ARG(comm->world <= B2G_MAX_RANKS && comm->rank >= 0 && comm->rank < comm->world, "bad world/rank");
        for (int r = 0; r < comm->world; ++r) B2G_CHECK_ARG(comm->peer_base[r] != nullptr, "NULL peer buffer");
    }
    cudaStream_t s = (cudaStream_t)stream;
    const int kind = b2g_mlp_train_loop_kind(nets, n_nets, comm ? comm->world : 1);
    if (kind == 0) {
        b2g_set_error("b2g_mlp_train_loop: no fused loop for these networks (tcgen05: hidden 64, (obs, out) in {(3,1), (3,2), "
                      "(4,1)}; general: one hidden size in {32, 64, 128}, obs 1..8, out 1..4, one GPU)");
        return B2G_ERR_UNSUPPORTED;
    }
    if (kind == 2) {
        LoopArgs args{};
        args.n_nets = n_nets;
        unsigned int* sync = reinterpret_cast<unsigned int*>(workspace);
        B2G_CUDA(cudaMemsetAsync(sync, 0, 256, s));
        int64_t max_mb = 1;
        for (int i = 0; i < n_nets; ++i) {
            fill_net(args.net[i], &nets[i], 0, workspace + 64, sync + 16 * i, i);
            args.net[i].a.O_rt = nets[i].obs_dim;
            args.net[i].a.A_rt = nets[i].out_dim;
            args.net[i].timing = nullptr;
            if (args.net[i].mbs > max_mb) max_mb = args.net[i].mbs;
        }
        const int op = nets[0].obs_dim <= 4 ? 4 : 8;
        int amax = 1;
        for (int i = 0; i < n_nets; ++i) amax = nets[i].out_dim > amax ? nets[i].out_dim : amax;
        const int ap = amax <= 1 ? 1 : (amax == 2 ? 2 : 4);
        B2G_CHECK_ARG(n_nets == 1 || nets[1].obs_dim <= op, "both networks read the same observations");
        if (nets[0].hidden == 32) return b2g_gen_loop_h32(op, ap, args, max_mb, s);
        if (nets[0].hidden == 64) return b2g_gen_loop_h64(op, ap, args, max_mb, s);
        return b2g_gen_loop_h128(op, ap, args, max_mb, s);
    }
    if (comm && comm->world > 1) {
        static bool timeout_set = false;
        if (!timeout_set) {
            const char* e = getenv("B2G_PEER_TIMEOUT_S");
            const double sec = e ? atof(e) : 120.0;
            const unsigned long long ns = (unsigned long long)((sec > 0.001 ? sec : 0.001) * 1e9);
            B2G_CUDA(cudaMemcpyToSymbol(g_peer_timeout_ns, &ns, sizeof(ns)));
            timeout_set = true;
        }
    }
    const int grid = 2 * b2g_sm_count();
    LoopArgs args{};
    args.n_nets = n_nets;
    args.ctas_per_net = grid / n_nets;
    {
        static int stagger = -1;
        if (stagger < 0) {
            const char* e = getenv("B2G_LOOP_STAGGER");
            stagger = e ? atoi(e) : 0;
        }
        args.stagger_cycles = stagger;
    }
    unsigned int* sync = reinterpret_cast<unsigned int*>(workspace);
    float* partial = workspace + 64;
    B2G_CUDA(cudaMemsetAsync(sync, 0, 256, s));
    int64_t off = 0;
    for (int i = 0; i < n_nets; ++i) {
        fill_net(args.net[i], &nets[i], args.ctas_per_net, partial + off, sync + 16 * i, i);
        if (comm && comm->world > 1) {
            NetLoop& n = args.net[i];
            {
                const MlpLayout lay{nets[i].obs_dim, 64, nets[i].out_dim};
                B2G_CHECK_ARG(lay.total() <= COMM_P, "network larger than the exchange area");
            }
            n.world = comm->world; n.rank = comm->rank;
            n.seq_ptr = reinterpret_cast<unsigned int*>(reinterpret_cast<uint8_t*>(comm->peer_base[comm->rank]) +
                                                        2 * recv_bytes_c) + 16 * i;
            for (int r = 0; r < comm->world; ++r)
                n.peer_recv[r] = reinterpret_cast<unsigned long long*>(reinterpret_cast<uint8_t*>(comm->peer_base[r]) +
                                                                       (size_t)i * recv_bytes_c);
        }
        off += (int64_t)args.ctas_per_net *
               (MlpLayout{nets[i].obs_dim, 64, nets[i].out_dim}.total() + MLP_NSTATS);
    }
    const int O = nets[0].obs_dim, AP = nets[0].out_dim;
#define B2G_LOOP_CASE(OO, AA) \
    if (O == OO && AP == AA) return launch_loop<OO, AA>(args, grid, s);
    B2G_LOOP_CASE(3, 1) B2G_LOOP_CASE(3, 2) B2G_LOOP_CASE(4, 1)
#undef B2G_LOOP_CASE
    b2g_set_error("b2g_mlp_train_loop: obs=%d out=%d is not built (built: (3,1), (3,2), (4,1))", O, AP);
    return B2G_ERR_UNSUPPORTED;
}

// debug: registers / occupancy of the tensor-core kernels as the driver sees them
template <class K>
static void print_occ(const char* name, K kernel, size_t smem) {
    cudaFuncAttributes fa{};
    cudaFuncGetAttributes(&fa, kernel);
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    int occ = 0, occ0 = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, tc::NT, smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ0, kernel, tc::NT, 0);
    printf("%-28s regs %3d  maxThreads %4d  smem %zu  occupancy %d (without smem %d)\n", name, fa.numRegs,
           fa.maxThreadsPerBlock, smem, occ, occ0);
}
extern "C" int b2g_debug_occupancy(void) {
    print_occ("mlp_train_tc<3,1>", tc::mlp_train_tc_kernel<3, 1>, tc::smem_bytes<3, 1>());
    print_occ("mlp_train_tc<3,2>", tc::mlp_train_tc_kernel<3, 2>, tc::smem_bytes<3, 2>());
    print_occ("mlp_train_tc<4,1>", tc::mlp_train_tc_kernel<4, 1>, tc::smem_bytes<4, 1>());
    print_occ("train_loop<3,1>", train_loop_kernel<3, 1>, loop_smem<3, 1>());
    print_occ("train_loop<3,2>", train_loop_kernel<3, 2>, loop_smem<3, 2>());
    print_occ("train_loop<4,1>", train_loop_kernel<4, 1>, loop_smem<4, 1>());
    int dev = 0, regs = 0, smem_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&regs, cudaDevAttrMaxRegistersPerMultiprocessor, dev);
    cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
    printf("device: %d registers / SM, %d B shared memory / SM\n", regs, smem_sm);
    return B2G_OK;
}

// ---- peer-mapped exchange buffers (CUDA IPC: one process per GPU on one node) ----
extern "C" int b2g_comm_alloc(int64_t bytes, void** ptr_out, void* handle_out_64) {
    B2G_CHECK_ARG(bytes > 0 && ptr_out && handle_out_64, "bad arguments");
    void* p = nullptr;
    B2G_CUDA(cudaMalloc(&p, (size_t)bytes));
    B2G_CUDA(cudaMemset(p, 0, (size_t)bytes));
    cudaIpcMemHandle_t h;
    B2G_CUDA(cudaIpcGetMemHandle(&h, p));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle_out_64, &h, 64);
    *ptr_out = p;
    return B2G_OK;
}
extern "C" int b2g_comm_open(const void* handle_64, void** ptr_out) {
    B2G_CHECK_ARG(handle_64 && ptr_out, "bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle_64, 64);
    B2G_CUDA(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
    return B2G_OK;
}
extern "C" int b2g_comm_close(void* ptr) {
    B2G_CUDA(cudaIpcCloseMemHandle(ptr));
    return B2G_OK;
}
extern "C" int b2g_comm_free(void* ptr) {
    B2G_CUDA(cudaFree(ptr));
    return B2G_OK;
}
