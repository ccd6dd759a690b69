// Tensor-core (tcgen05, kind::tf32, 3xTF32 split) form of the fused minibatch kernel of
// mlp_train.cuh for hidden = (64, 64): same inputs, same per-CTA partial output, same
// per-row loss code (mlp_row_loss) -- only the three H x H contractions, the bias/W1
// gradient column sums and their staging differ.  Replaces the same reference code:
// torch/algos/vpg.py:250-293, ppo.py:96-132, vpg.py:324-347,408-455,
// value_functions/gaussian_mlp_value_function.py:81-97 and the autograd graph of
// torch/modules/multi_headed_mlp_module.py:136-151.
//
// Shape of the launch: 64-row tiles, 256 threads, ~110 KB of shared memory and 256 TMEM
// columns per CTA, so TWO CTAs are resident per SM: while one waits for its MMAs the other
// runs its CUDA-core stage (tanh, loss, splits), which is how the serial per-tile chain
// L1 -> GEMM1 -> EPI1 -> GEMM2|GEMM3 -> EPI2 -> GEMM4 keeps both pipes busy.
//
// Data layout in shared memory: every operand is a stack of "panels"
//   panel = [rows][32 floats = 128 B], the four 32-byte chunks of a row XOR-swizzled by row%4
// (UMMA descriptor layout_type 1: 128-byte swizzle with 32-byte atoms).  The SAME bytes are a
// K-major operand (M/N = row, K = feature; SBO = 512, +32 B per K=8 step) and an MN-major
// operand (M/N = feature, K = row; LBO = panel stride, SBO = 512, +1024 B per K=8 step), so the
// activations are written once and feed both the forward/input-gradient GEMMs (contraction over
// features) and the weight-gradient GEMMs (contraction over rows).  Verified on B200 by
// tools/umma_probe.cu (`panel` cases G1-G4).
//
// Per 64-row tile (fp32 = hi + lo, hi = tf32 round-to-nearest, lo = x - hi; products hi*hi +
// lo*hi + hi*lo accumulate in fp32 in TMEM):
//   L1 (CUDA cores)  h1 = tanh(W1 x + b1)                       -> ACT panels hi0 hi1 | aug | lo0 lo1
//   GEMM1  D[r][c]  = sum_k h1[r][k] W2[c][k]        M64 N64    A=ACT (K-major)  B=W (K-major)
//   EPI1   h2 = tanh(D + b2); mu = W3 h2 + b3; loss; dz2        -> DZ panels hi0 hi1 lo0 lo1
//   GEMM2  D[r][k]  = sum_c dz2[r][c] W2[c][k]       M64 N64    A=DZ (K-major)   B=W (MN-major)
//   GEMM3  G[m][n] += sum_r dz2{hi|lo}[r][m] [h1hi|1 x|h1lo][r][n]  M128(hi;lo stacked) N160, MN/MN
//   EPI2   dz1 = D (1 - h1^2)                                   -> DZ panels (after GEMM3 drained)
//   GEMM4  F[m][n] += sum_r dz1{hi|lo}[r][m] [1 xhi|0 xlo][r][n]  M128 N16, MN/MN
// dW2 = G[:, 0:64] + G[:, 96:160] (lanes m and m+64 added), db2 = G[:, 64], db1 = F[:, 0],
// dW1 = F[:, 1:1+O] + F[:, 9:9+O];
// dW3 accumulates in registers, db3 / dlog_std / statistics are row sums as in mlp_train.cuh.
// The M=64 accumulator D lives in TMEM lanes (m%16) + 32*(m/16); the epilogues read it with
// tcgen05.ld.16x256b (thread t of warp q: rows 16q + t/4 and +8, column pairs 8i + 2(t%4)),
// so all 32 lanes of every warp carry data.
#pragma once
#include "mlp_train.cuh"

namespace tc {
constexpr int R = 64, NT = 256, H = 64;
constexpr uint32_t PANEL = R * 128, WPANEL = H * 128;
constexpr uint32_t OFF_ACT = 0;                      // 5 panels: hi0 hi1 aug lo0 lo1
constexpr uint32_t OFF_DZ = OFF_ACT + 5 * PANEL;     // 4 panels: hi0 hi1 lo0 lo1
constexpr uint32_t OFF_W = OFF_DZ + 4 * PANEL;       // 4 weight panels: hi0 hi1 lo0 lo1
constexpr uint32_t OFF_SMALL = OFF_W + 4 * WPANEL;   // 106,496
constexpr uint32_t TM_D = 0, TM_G = 64, TM_F = 224;  // D 64 | G 160 (h1hi 64, aug 32, h1lo 64) | F 16 columns

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// shared-memory matrix descriptor: layout_type 1 (128B swizzle, 32B atoms), version 1
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46) | (1ull << 61);
}
// instruction descriptor, kind::tf32, fp32 accumulate
__host__ __device__ constexpr uint32_t umma_idesc(int M, int N, int a_mn, int b_mn) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),
        "l"(da), "l"(db), "r"(idesc), "r"(acc)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// one lane of a converged warp
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(pred));
    return pred != 0;
}

// bounded spin: a lost MMA completion must surface as an error, not as a hung GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    for (uint32_t spin = 0; !done; ++spin) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}\n"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (spin > (1u << 24)) __trap();
    }
}

// M=64 accumulator fragment: 16 lanes x 32 columns -> v[4i+0,1] = (row g, cols 8i+2p,+1),
// v[4i+2,3] = (row g+8, same cols); g = lane/4, p = lane%4
__device__ __forceinline__ void tmem_ld_frag(uint32_t taddr, float (&v)[16]) {
    uint32_t u[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x4.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]),
          "=r"(u[8]), "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(u[i]);
}
// 16 consecutive fp32 columns of this thread's TMEM lane (warp w reads lanes 32*(w%4)..+31)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t u[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]),
          "=r"(u[8]), "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(u[i]);
}

__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, float (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]),
          "=f"(v[8]), "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// pins the uses of registers filled by tmem_ld16_nowait behind the wait (volatile asms keep their order)
__device__ __forceinline__ void tmem_ld_use(float (&v)[16]) {
    asm volatile("" : "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3]), "+f"(v[4]), "+f"(v[5]), "+f"(v[6]), "+f"(v[7]),
                      "+f"(v[8]), "+f"(v[9]), "+f"(v[10]), "+f"(v[11]), "+f"(v[12]), "+f"(v[13]), "+f"(v[14]), "+f"(v[15]));
}

// tanh x = 1 - 2 / (1 + e^{2x}) with the SFU's ex2 / rcp: absolute error ~2e-7 (|tanh| <= 1),
// a quarter of tanhf's instruction count; saturates correctly (e -> inf: 1, e -> 0: -1)
__device__ __forceinline__ float tanh_sfu(float x) {
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * 2.8853900817779268f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.0f));
    return fmaf(-2.0f, r, 1.0f);
}
#ifdef B2G_TC_PRECISE_TANH
#define TC_TANH(x) tanhf(x)
#else
#define TC_TANH(x) tanh_sfu(x)
#endif

// fp32 -> (tf32 round-to-nearest, exact remainder)
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
    uint32_t u;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
    hi = __uint_as_float(u);
    lo = x - hi;
}
// byte offset of the 16-byte chunk j (features 4j..4j+3 of the panel) of `row`
__device__ __forceinline__ uint32_t chunk_off(int row, int j) {
    return (uint32_t)row * 128u + ((((uint32_t)(j >> 1) ^ (uint32_t)row) & 3u) << 5) + ((uint32_t)(j & 1) << 4);
}
// byte offset of the feature pair (8i + 2p, +1) of `row` (fragment layout of tmem_ld_frag)
__device__ __forceinline__ uint32_t pair_off(int row, int i, int p) {
    return (uint32_t)row * 128u + ((((uint32_t)i ^ (uint32_t)row) & 3u) << 5) + (uint32_t)p * 8u;
}
__device__ __forceinline__ void store_split4(uint8_t* hi_panel, uint8_t* lo_panel, uint32_t off, const float* v) {
    float4 h, l;
    split_tf32(v[0], h.x, l.x); split_tf32(v[1], h.y, l.y);
    split_tf32(v[2], h.z, l.z); split_tf32(v[3], h.w, l.w);
    *reinterpret_cast<float4*>(hi_panel + off) = h;
    *reinterpret_cast<float4*>(lo_panel + off) = l;
}
__device__ __forceinline__ void store_split2(uint8_t* hi_panel, uint8_t* lo_panel, uint32_t off, float v0, float v1) {
    float2 h, l;
    split_tf32(v0, h.x, l.x); split_tf32(v1, h.y, l.y);
    *reinterpret_cast<float2*>(hi_panel + off) = h;
    *reinterpret_cast<float2*>(lo_panel + off) = l;
}

// ---- MMA issue sequences (one thread) ----
__device__ __forceinline__ void issue_gemm1(uint32_t sb, uint32_t tmem) {
    constexpr uint32_t ID = umma_idesc(64, 64, 0, 0);
#pragma unroll
    for (int pass = 0; pass < 3; ++pass) {
        const uint32_t ao = OFF_ACT + (pass == 1 ? 3 * PANEL : 0), bo = OFF_W + (pass == 2 ? 2 * WPANEL : 0);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
            umma_tf32(tmem + TM_D, umma_desc(sb + ao + (ks >> 2) * PANEL + (ks & 3) * 32, 1024, 512),
                      umma_desc(sb + bo + (ks >> 2) * WPANEL + (ks & 3) * 32, 1024, 512), ID, (pass | ks) != 0);
    }
}
// MMAs [lo, hi) of the 24 (pass-major: hi*hi, lo*hi, hi*lo; 8 K steps each): the issuing lane is held
// back while the tensor pipe's queue is full, so the caller interleaves chunks with its other work
template <int LO, int HI>
__device__ __forceinline__ void issue_gemm2_part(uint32_t sb, uint32_t tmem) {
    constexpr uint32_t ID = umma_idesc(64, 64, 0, 1);
#pragma unroll
    for (int m = LO; m < HI; ++m) {
        const int pass = m >> 3, ks = m & 7;
        const uint32_t ao = OFF_DZ + (pass == 1 ? 2 * PANEL : 0), bo = OFF_W + (pass == 2 ? 2 * WPANEL : 0);
        umma_tf32(tmem + TM_D, umma_desc(sb + ao + (ks >> 2) * PANEL + (ks & 3) * 32, 1024, 512),
                  umma_desc(sb + bo + ks * 1024, WPANEL, 512), ID, m != 0);
    }
}
__device__ __forceinline__ void issue_gemm2(uint32_t sb, uint32_t tmem) { issue_gemm2_part<0, 24>(sb, tmem); }
// nks = ceil(valid rows / 8): rows past the minibatch contribute nothing and are skipped.
// One MMA per K step covers all five ACT panels (N = 160: h1 hi | aug | h1 lo), so the stacked
// dz2 operand is fetched from shared memory once.
__device__ __forceinline__ void issue_gemm3(uint32_t sb, uint32_t tmem, bool first, int nks) {
    constexpr uint32_t ID = umma_idesc(128, 160, 1, 1);
#pragma unroll
    for (int ks = 0; ks < R / 8; ++ks)
        if (ks < nks)
            umma_tf32(tmem + TM_G, umma_desc(sb + OFF_DZ + ks * 1024, PANEL, 512),
                      umma_desc(sb + OFF_ACT + ks * 1024, PANEL, 512), ID, !(first && ks == 0));
}
// N = 16: [1 x_hi 0.. | 0 x_lo 0..] of the aug panel
__device__ __forceinline__ void issue_gemm4(uint32_t sb, uint32_t tmem, bool first, int nks) {
    constexpr uint32_t ID = umma_idesc(128, 16, 1, 1);
#pragma unroll
    for (int ks = 0; ks < R / 8; ++ks)
        if (ks < nks)
            umma_tf32(tmem + TM_F, umma_desc(sb + OFF_DZ + ks * 1024, PANEL, 512),
                      umma_desc(sb + OFF_ACT + 2 * PANEL + ks * 1024, PANEL, 512), ID, !(first && ks == 0));
}

template <int O, int A>
struct Small {  // float offsets inside the small region
    static constexpr int NIN = A + 3;  // act[A], adv, old_logp, ret
    static constexpr int W1 = 0, B1 = W1 + H * O, B2 = B1 + H, W3 = B2 + H, B3 = W3 + A * H,
                         X = B3 + 4, MU = X + R * 4, IN = MU + 2 * A * R /* [2][R][NIN] */, SRC = IN + 2 * R * NIN,
                         NEXT = SRC + 4 * R /* int64[2][R] */, BARS = NEXT + 4 * R /* int64[2][R] */,
                         TSLOT = BARS + 8 /* 4 x u64 */, END = TSLOT + 4;
};
template <int O, int A>
constexpr size_t smem_bytes() { return OFF_SMALL + sizeof(float) * Small<O, A>::END; }

// minibatch row id of row m of the minibatch, or -1 past its end: supplied index list,
// on-the-fly Feistel permutation (b2g_perm_apply, same ids as b2g_permutation), or sequential
__device__ __forceinline__ int64_t mb_src(const MlpArgs& a, int64_t m) {
    if (m >= a.mb) return -1;
    if (a.idx) return (int64_t)a.idx[m];
    if (a.perm_n > 0) return (int64_t)b2g_perm_apply(a.perm_keys, a.perm_half_bits, (uint64_t)a.perm_n, (uint64_t)(a.start + m));
    return a.start + m;
}

template <int O, int A>
struct RowData {  // what the gather stage loads per row
    float x[4];
    float in[A + 3];
};
template <int O, int A>
__device__ __forceinline__ RowData<O, A> load_row(const MlpArgs& a, int64_t src) {
    RowData<O, A> d;
#pragma unroll
    for (int o = 0; o < 4; ++o) d.x[o] = 0.0f;
#pragma unroll
    for (int i = 0; i < A + 3; ++i) d.in[i] = 0.0f;
    if (src >= 0) {
#pragma unroll
        for (int o = 0; o < O; ++o) d.x[o] = a.obs[src * O + o];
        if (a.loss_kind == LOSS_VF) {
            d.in[A + 2] = a.ret[src];
        } else if (a.loss_kind != LOSS_NONE) {
#pragma unroll
            for (int q = 0; q < A; ++q) d.in[q] = a.act[src * A + q];
            d.in[A] = a.adv[src];
            if (a.loss_kind == LOSS_PPO) d.in[A + 1] = a.old_logp[src];
        }
    }
    return d;
}

// One CTA's state: shared-memory carve-up, TMEM base, thread roles and the register accumulators.
// Used by the one-minibatch kernel below and by the persistent training loop (train_loop.cu).
template <int O, int A>
struct Cta {
    static_assert(O <= 4 && A <= 4, "aug panel holds [1 | x] in 8 columns; obs staged as float4");
    using S = Small<O, A>;
    static constexpr int NIN = S::NIN;
    uint8_t* smem;
    float *sm, *sW1, *sB1, *sB2, *sW3, *sB3, *sX, *sMu, *sIn;
    int64_t *sSrc, *sNext;
    uint32_t sb, bar1, bar2, bar3, bar4, tmem, tiles_done;
    int tid, w, l, q4, hf, g, p, ra, rb, c0;
    bool first;  // no MMA has written the G/F accumulators since the last readout
    MlpDist dist;
    float gW3[A][8], gB3[A];
    MlpRowAcc racc;
    long long* stamps;
    int64_t src_n;          // gather pipeline: row id / data of this thread's row in the next tile
    RowData<O, A> row_n;

    __device__ __forceinline__ void stamp(int i) const {
        if (stamps && tid == 0 && i < 32) stamps[i] = clock64();
    }

    // barriers + TMEM allocation; every thread of the CTA calls it once
    __device__ __forceinline__ void init(uint8_t* smem_) {
        smem = smem_;
        sm = reinterpret_cast<float*>(smem + OFF_SMALL);
        sW1 = sm + S::W1; sB1 = sm + S::B1; sB2 = sm + S::B2; sW3 = sm + S::W3; sB3 = sm + S::B3;
        sX = sm + S::X; sMu = sm + S::MU; sIn = sm + S::IN;
        sSrc = reinterpret_cast<int64_t*>(sm + S::SRC);
        sNext = reinterpret_cast<int64_t*>(sm + S::NEXT);
        uint32_t* tslot = reinterpret_cast<uint32_t*>(sm + S::TSLOT);
        sb = smem_u32(smem);
        bar1 = smem_u32(sm + S::BARS); bar2 = bar1 + 8; bar3 = bar1 + 16; bar4 = bar1 + 24;
        tid = threadIdx.x;
        w = __shfl_sync(0xffffffffu, tid >> 5, 0);  // provably warp-uniform
        l = tid & 31;
        q4 = w & 3; hf = w >> 2; g = l >> 2; p = l & 3;
        ra = 16 * q4 + g; rb = ra + 8; c0 = 32 * hf;
        tiles_done = 0;
        first = true;
        stamps = nullptr;
        if (tid == 0) {
            if (sb & 1023u) __trap();
#pragma unroll
            for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar1 + 8 * i));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        if (w == 1) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(tslot)) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
        for (int i = tid; i < (int)(PANEL / 16); i += NT)  // aug panel: columns 16..31 are never written
            reinterpret_cast<float4*>(smem + OFF_ACT + 2 * PANEL)[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        tc_before_sync();
        __syncthreads();
        tc_after_sync();
        tmem = __shfl_sync(0xffffffffu, *tslot, 0);
    }

    __device__ __forceinline__ void fini() {
        tc_before_sync();
        __syncthreads();
        if (w == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
    }

    // W2 as tf32 hi/lo panels, the small layers as fp32; call with no MMA in flight.
    // `ld` loads one parameter (plain or L2-coherent, the training loop rewrites them)
    template <class Ld>
    __device__ __forceinline__ void load_weights(const MlpArgs& a, const float* params, Ld ld) {
        const MlpLayout L{O, H, A};
        // All CTAs read the same 18 KB right after the update: the loads are fully coalesced (a warp reads
        // 128 contiguous bytes; W2 starts at a 4-byte aligned offset, so no vector loads) and all issued
        // before the first shared-memory store (a load behind a store through a generic pointer would
        // wait for it: one L2 round trip per group instead of one in total).
        static_assert(H * O <= NT && A * H <= NT && 2 * H <= NT && H * H == 16 * NT, "thread <-> parameter mapping");
        const float* w2 = params + L.off_w2();
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = ld(w2 + tid + NT * i);
        const float w1v = tid < H * O ? ld(params + L.off_w1() + tid) : 0.0f;
        const float bv = tid < H ? ld(params + L.off_b1() + tid) : (tid < 2 * H ? ld(params + L.off_b2() + tid - H) : 0.0f);
        const float w3v = tid < A * H ? ld(params + L.off_w3() + tid) : 0.0f;
        const float b3v = tid < A ? ld(params + L.off_b3() + tid) : 0.0f;
        const float lsv = ld(params + L.off_logstd());
        {
            // element e = tid + 256 i of W2[c][k]: k = e & 63 does not depend on i; a warp writes one
            // 128-byte panel row (32 banks)
            const int k = tid & 63;
            uint8_t* hi_p = smem + OFF_W + (uint32_t)(k >> 5) * WPANEL + (k & 3) * 4;
            uint8_t* lo_p = hi_p + 2 * WPANEL;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const int c = (tid >> 6) + 4 * i;
                const uint32_t off = chunk_off(c, (k & 31) >> 2);
                float h, l2;
                split_tf32(v[i], h, l2);
                *reinterpret_cast<float*>(hi_p + off) = h;
                *reinterpret_cast<float*>(lo_p + off) = l2;
            }
        }
        if (tid < H * O) sW1[tid] = w1v;
        if (tid < H) sB1[tid] = bv;
        else if (tid < 2 * H) sB2[tid - H] = bv;
        if (tid < A * H) sW3[tid] = w3v;
        if (tid < A) sB3[tid] = b3v;
        dist = mlp_dist(a, lsv);
        mlp_dist_entropy<A>(a, dist);
        fence_async_smem();
        __syncthreads();
    }

    __device__ __forceinline__ void zero_acc() {
#pragma unroll
        for (int q = 0; q < A; ++q) {
            gB3[q] = 0.0f;
#pragma unroll
            for (int c = 0; c < 8; ++c) gW3[q][c] = 0.0f;
        }
        racc = MlpRowAcc{};
    }

    // Gather pipeline: threads 0..R-1 own one row and load its data one tile ahead; threads
    // R..2R-1 evaluate the row ids (index list or on-the-fly permutation) two tiles ahead while
    // the CTA waits for GEMM1, and hand them over in shared memory (sNext, double-buffered), so
    // no stage waits on a dependent load or on the permutation arithmetic.  prime() starts the
    // pipeline for run_tiles(a, tile0, stride); neither ids nor row data depend on the weights,
    // so the training loop primes the next minibatch before it waits for the other CTAs.
    __device__ __forceinline__ void prime(const MlpArgs& a, int64_t tile0, int64_t stride) {
        const int64_t ntiles = (a.mb + R - 1) / R;
        src_n = -1;
        if (tid < R) {
            src_n = tile0 < ntiles ? mb_src(a, tile0 * R + tid) : -1;
            row_n = load_row<O, A>(a, src_n);
        } else if (tid < 2 * R) {
            sNext[tid - R] = tile0 + stride < ntiles ? mb_src(a, (tile0 + stride) * R + (tid - R)) : -1;
        }
    }

    // h1 = tanh(W1 x + b1) of the rows published in sX, into registers: thread (w, l) owns features
    // f0..f0+3 of rows 16*rbk + rs + 4i (i < 4); store_h1() later writes them as tf32 hi/lo panels
    // (a quarter-warp writes the 8 chunks of one row: conflict-free)
    template <int I0 = 0, int I1 = 4>
    __device__ __forceinline__ void layer1(float (&hreg)[4][4]) const {
        const int pn = w & 1, rbk = w >> 1, j = l & 7, rs = l >> 3, f0 = 32 * pn + 4 * j;
        float w1[4][O], bb[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            bb[q] = sB1[f0 + q];
#pragma unroll
            for (int o = 0; o < O; ++o) w1[q][o] = sW1[(f0 + q) * O + o];
        }
#pragma unroll
        for (int i = I0; i < I1; ++i) {
            const int r = 16 * rbk + rs + 4 * i;
            const float4 xv = *reinterpret_cast<const float4*>(sX + r * 4);
            const float x[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                float z = bb[q];
#pragma unroll
                for (int o = 0; o < O; ++o) z = fmaf(w1[q][o], x[o], z);
                hreg[i][q] = TC_TANH(z);
            }
        }
    }
    __device__ __forceinline__ void store_h1(const float (&hreg)[4][4]) const {
        const int pn = w & 1, rbk = w >> 1, j = l & 7, rs = l >> 3;
        uint8_t* hp = smem + OFF_ACT + pn * PANEL;
        uint8_t* lp = smem + OFF_ACT + (3 + pn) * PANEL;
#pragma unroll
        for (int i = 0; i < 4; ++i) store_split4(hp, lp, chunk_off(16 * rbk + rs + 4 * i, j), hreg[i]);
    }
    // rows of the tile in (src_n, row_n) -> shared memory; then start loading the rows whose ids sit in sNext
    // (sSrc / sIn are double-buffered on `buf`: the loss stage of the current tile still reads the other half)
    __device__ __forceinline__ void publish_rows(const MlpArgs& a, int buf) {
        if (tid < R) {
            sSrc[buf * R + tid] = src_n;
            *reinterpret_cast<float4*>(sX + tid * 4) = make_float4(row_n.x[0], row_n.x[1], row_n.x[2], row_n.x[3]);
#pragma unroll
            for (int i = 0; i < NIN; ++i) sIn[(buf * R + tid) * NIN + i] = row_n.in[i];
            src_n = sNext[tid];
            row_n = load_row<O, A>(a, src_n);
        }
    }

    // forward + loss + backward of the 64-row tiles tile0, tile0 + stride, ... of minibatch `a`;
    // gradients accumulate in TMEM (G, F) and registers until write_partial().
    // Software pipeline over tiles: while the tensor cores run GEMM2/GEMM3 of tile t, the CUDA cores
    // publish the rows of tile t+1 and evaluate its first layer into registers; the rows of t+2 are in
    // flight and the ids of t+3.. are computed while GEMM1 runs.  Per tile only the panel stores, the
    // two epilogues and the MMAs themselves remain on the critical path.
    // STATS = false: the logging sums of the loss are not accumulated (the training loop).
    template <bool STATS = true>
    __device__ __forceinline__ void run_tiles(const MlpArgs& a, int64_t tile0, int64_t stride) {
        const int64_t ntiles = (a.mb + R - 1) / R;
        const uint32_t tfrag = tmem + ((uint32_t)(32 * q4) << 16) + TM_D + c0;
        uint8_t* const act_hi = smem + OFF_ACT + hf * PANEL;
        uint8_t* const act_lo = smem + OFF_ACT + (3 + hf) * PANEL;
        uint8_t* const dz_hi = smem + OFF_DZ + hf * PANEL;
        uint8_t* const dz_lo = smem + OFF_DZ + (2 + hf) * PANEL;
        float hreg[4][4];
        if (tile0 < ntiles) {  // prologue: rows and first layer of the first tile
            publish_rows(a, 0);
            __syncthreads();
            layer1(hreg);
        }
        int it = 0;
        for (int64_t tile = tile0; tile < ntiles; tile += stride, ++it) {
            const uint32_t par = tiles_done & 1;
            const int64_t m0 = tile * R;
            const int valid = (int)((a.mb - m0) < R ? (a.mb - m0) : R);
            const int nks = (valid + 7) >> 3;
            const bool has_next = tile + stride < ntiles;
            // ---- h1 panels from registers; aug panel row = [1 x_hi 0.. | 0 x_lo 0..]
            // The previous tile's GEMM3 (the last reader of the h1 panels) was waited for in its EPI2; only
            // its GEMM4 may still run, and that reads the DZ and aug panels: the h1 stores go first.
            store_h1(hreg);
            if (tiles_done > 0) mbar_wait(bar4, par ^ 1);  // every MMA of the previous tile has drained
            stamp(2 + 10 * it);
            if (tid < R) {
                const float4 xv = *reinterpret_cast<const float4*>(sX + tid * 4);
                const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
                float xh[8], xl[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) { xh[i] = 0.0f; xl[i] = 0.0f; }
                xh[0] = 1.0f;
#pragma unroll
                for (int o = 0; o < O; ++o) split_tf32(xs[o], xh[1 + o], xl[1 + o]);
                uint8_t* aug = smem + OFF_ACT + 2 * PANEL;
                const uint32_t o0 = chunk_off(tid, 0), o1 = chunk_off(tid, 2);
                *reinterpret_cast<float4*>(aug + o0) = make_float4(xh[0], xh[1], xh[2], xh[3]);
                *reinterpret_cast<float4*>(aug + o0 + 16) = make_float4(xh[4], xh[5], xh[6], xh[7]);
                *reinterpret_cast<float4*>(aug + o1) = make_float4(xl[0], xl[1], xl[2], xl[3]);
                *reinterpret_cast<float4*>(aug + o1 + 16) = make_float4(xl[4], xl[5], xl[6], xl[7]);
            }
            stamp(4 + 10 * it);
            fence_async_smem();
            __syncthreads();
            if (w == 0) {  // descriptors stay in uniform registers: the whole warp takes the branch
                tc_after_sync();
                if (elect_one()) {
                    issue_gemm1(sb, tmem);
                    umma_commit(bar1);
                }
                __syncwarp();
            }
            stamp(5 + 10 * it);
            // Row ids of the tile after next, off the critical path.  Warps 4 and 5 write them, warps 0
            // and 1 read them in publish_rows() at the end of EPI1: the only barrier in between is the
            // pair barrier of EPI1, which joins exactly warps (0, 4) and (1, 5) -- entry i is written by
            // thread 2R + i and read by thread i.
            if (tid >= 2 * R && tid < 3 * R) {
                const int64_t t2 = tile + 2 * stride;
                sNext[tid - 2 * R] = t2 < ntiles ? mb_src(a, t2 * R + (tid - 2 * R)) : -1;
            }
            mbar_wait(bar1, par);
            tc_after_sync();
            stamp(6 + 10 * it);
            // ---- EPI1: h2, mu, loss, dz2
            float h2[16];  // [4i + 2*row + e]
            float w3c[A][8];  // this thread's columns of W3
            {
                float z[16];
                tmem_ld_frag(tfrag, z);
                float mpa[A], mpb[A];
#pragma unroll
                for (int q = 0; q < A; ++q) { mpa[q] = 0.0f; mpb[q] = 0.0f; }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float2 b2v = *reinterpret_cast<const float2*>(sB2 + c0 + 8 * i + 2 * p);
                    const float b2c[2] = {b2v.x, b2v.y};
#pragma unroll
                    for (int q = 0; q < A; ++q) {
                        const float2 wv = *reinterpret_cast<const float2*>(sW3 + q * H + c0 + 8 * i + 2 * p);
                        w3c[q][2 * i] = wv.x;
                        w3c[q][2 * i + 1] = wv.y;
                    }
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const float ha = TC_TANH(z[4 * i + e] + b2c[e]);
                        const float hb = TC_TANH(z[4 * i + 2 + e] + b2c[e]);
                        h2[4 * i + e] = ha;
                        h2[4 * i + 2 + e] = hb;
#pragma unroll
                        for (int q = 0; q < A; ++q) {
                            mpa[q] = fmaf(w3c[q][2 * i + e], ha, mpa[q]);
                            mpb[q] = fmaf(w3c[q][2 * i + e], hb, mpb[q]);
                        }
                    }
                }
#pragma unroll
                for (int q = 0; q < A; ++q) {
                    mpa[q] += __shfl_xor_sync(0xffffffffu, mpa[q], 1);
                    mpb[q] += __shfl_xor_sync(0xffffffffu, mpb[q], 1);
                    mpa[q] += __shfl_xor_sync(0xffffffffu, mpa[q], 2);
                    mpb[q] += __shfl_xor_sync(0xffffffffu, mpb[q], 2);
                    if (p == 0) {
                        sMu[(hf * A + q) * R + ra] = mpa[q];
                        sMu[(hf * A + q) * R + rb] = mpb[q];
                    }
                }
            }
            // the two column halves of rows 16*q4 .. 16*q4+15 sit in warps q4 and q4 + 4: only that pair meets
            asm volatile("bar.sync %0, 64;" ::"r"(1 + q4) : "memory");
            {
                // the loss of a row is evaluated by the lanes p and p^2 of its quad (p&1 picks the row),
                // then handed to the other lanes: half the instructions of evaluating both rows everywhere
                float dmu[2][A];
                {
                    const int r = (p & 1) ? rb : ra;
                    float mu[A], dm[A];
#pragma unroll
                    for (int q = 0; q < A; ++q) {
                        mu[q] = sB3[q] + sMu[q * R + r] + sMu[(A + q) * R + r];
                        dm[q] = 0.0f;
                    }
                    const int cb = (it & 1) * R;
                    if (sSrc[cb + r] >= 0) {
                        const float* rin = sIn + (cb + r) * NIN;
                        MlpRowIn<A> in;
#pragma unroll
                        for (int q = 0; q < A; ++q) { in.act[q] = rin[q]; in.old_mean[q] = 0.0f; }
                        in.adv = rin[A]; in.old_logp = rin[A + 1]; in.ret = rin[A + 2];
                        mlp_row_loss<A, STATS, true>(a, in, mu, dist, dm, racc);
                    }
#pragma unroll
                    for (int q = 0; q < A; ++q) {
                        gB3[q] += dm[q];
                        dmu[0][q] = __shfl_sync(0xffffffffu, dm[q], l & ~3);
                        dmu[1][q] = __shfl_sync(0xffffffffu, dm[q], (l & ~3) | 1);
                    }
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float dz[4];  // [2*row + e]
#pragma unroll
                    for (int s = 0; s < 2; ++s)
#pragma unroll
                        for (int e = 0; e < 2; ++e) {
                            const float h = h2[4 * i + 2 * s + e];
                            float dh = 0.0f;
#pragma unroll
                            for (int q = 0; q < A; ++q) {
                                dh = fmaf(w3c[q][2 * i + e], dmu[s][q], dh);
                                gW3[q][2 * i + e] = fmaf(dmu[s][q], h, gW3[q][2 * i + e]);
                            }
                            dz[2 * s + e] = dh * (1.0f - h * h);
                        }
                    store_split2(dz_hi, dz_lo, pair_off(ra, i, p), dz[0], dz[1]);
                    store_split2(dz_hi, dz_lo, pair_off(rb, i, p), dz[2], dz[3]);
                }
            }
            if (has_next) publish_rows(a, (it + 1) & 1);  // rows of the next tile (the other sIn / sSrc half)
            stamp(7 + 10 * it);
            fence_async_smem();
            tc_before_sync();
            __syncthreads();
            // ---- the tensor cores take over: warp 0 issues GEMM2 (its lanes are held back by the
            // tensor-pipe queue meanwhile); every warp evaluates the first layer of the next tile into
            // registers; warp 4 then issues GEMM3 (weight gradients: nobody waits for it before the
            // dz1 store), so that no single warp carries both issue sequences.
            if (w == 0) {
                tc_after_sync();
                if (elect_one()) {
                    issue_gemm2(sb, tmem);
                    umma_commit(bar2);
                }
                __syncwarp();
            }
            stamp(8 + 10 * it);
            if (has_next) layer1(hreg);
            if (w == 4) {
                tc_after_sync();
                if (elect_one()) {
                    issue_gemm3(sb, tmem, first, nks);
                    umma_commit(bar3);
                }
                __syncwarp();
            }
            stamp(3 + 10 * it);
            // ---- EPI2: dz1 = dh1 (1 - h1^2)
            mbar_wait(bar2, par);
            tc_after_sync();
            stamp(9 + 10 * it);
            {
                float d[16];
                tmem_ld_frag(tfrag, d);
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int s = 0; s < 2; ++s) {
                        const uint32_t off = pair_off(s ? rb : ra, i, p);
                        const float2 hh = *reinterpret_cast<const float2*>(act_hi + off);
                        const float2 hl = *reinterpret_cast<const float2*>(act_lo + off);
                        const float h0 = hh.x + hl.x, h1 = hh.y + hl.y;
                        d[4 * i + 2 * s] *= (1.0f - h0 * h0);
                        d[4 * i + 2 * s + 1] *= (1.0f - h1 * h1);
                    }
                mbar_wait(bar3, par);  // GEMM3 still reads dz2 from the DZ panels
                stamp(10 + 10 * it);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    store_split2(dz_hi, dz_lo, pair_off(ra, i, p), d[4 * i], d[4 * i + 1]);
                    store_split2(dz_hi, dz_lo, pair_off(rb, i, p), d[4 * i + 2], d[4 * i + 3]);
                }
            }
            fence_async_smem();
            tc_before_sync();
            __syncthreads();
            if (w == 0) {
                tc_after_sync();
                if (elect_one()) {
                    issue_gemm4(sb, tmem, first, nks);
                    umma_commit(bar4);
                }
                __syncwarp();
            }
            first = false;
            ++tiles_done;
            stamp(11 + 10 * it);
        }
    }

    // drain the MMAs, fold G/F/register accumulators into this CTA's partial row
    // [P gradients | MLP_NSTATS statistics]; afterwards the accumulators start from zero again
    __device__ __forceinline__ void write_partial(float* out, bool with_stats) {
        const MlpLayout L{O, H, A};
        const int P = L.total();
        stamp(28);
        if (tiles_done > 0) mbar_wait(bar4, (tiles_done - 1) & 1);
        tc_after_sync();
        stamp(29);
        float* scr = reinterpret_cast<float*>(smem + OFF_ACT);
        constexpr int SW = 81;  // scratch row stride (odd: conflict-free)
        const uint32_t tlane = tmem + ((uint32_t)(32 * q4) << 16);
        if (first) {  // this CTA had no tile: its partial is zero
            for (int e = tid; e < P + MLP_NSTATS; e += NT) out[e] = 0.0f;
            __syncthreads();
            return;
        }
        // Stage 1: every lane of G/F folds its own hi/lo column groups into scr[lane][81]
        // (lanes 64..127 hold the lo-split rows of dz); the register accumulators (dW3, row sums)
        // are reduced with shuffles into scr2.  Stage 2 (after one barrier): all threads add the
        // two lane halves / warp partials and store coalesced.
        float* scr2 = scr + 128 * SW;  // [4][A*H] dW3 partials | [NS][8] row sums
        constexpr int NS = A + 1 + 5;
        float* scr3 = scr2 + 4 * A * H;
        {
            // warps w and w + 4 own the same TMEM lanes: each takes half of the columns
            const int m = tid & 127, gq = 2 * hf;
            float v0[16], u0[16], v1[16], u1[16];
            tmem_ld16_nowait(tlane + TM_G + 16 * gq, v0);
            tmem_ld16_nowait(tlane + TM_G + 96 + 16 * gq, u0);
            tmem_ld16_nowait(tlane + TM_G + 16 * gq + 16, v1);
            tmem_ld16_nowait(tlane + TM_G + 96 + 16 * gq + 16, u1);
            tmem_ld_wait();
            tmem_ld_use(v0); tmem_ld_use(u0); tmem_ld_use(v1); tmem_ld_use(u1);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                scr[m * SW + 16 * gq + i] = v0[i] + u0[i];
                scr[m * SW + 16 * gq + 16 + i] = v1[i] + u1[i];
            }
            if (hf == 0) {
                float v[16], f[16];
                tmem_ld16_nowait(tlane + TM_G + 64, v);
                tmem_ld16_nowait(tlane + TM_F, f);
                tmem_ld_wait();
                tmem_ld_use(v); tmem_ld_use(f);
                scr[m * SW + 64] = v[0];
#pragma unroll
                for (int i = 0; i < 8; ++i) scr[m * SW + 72 + i] = f[i] + f[8 + i];
            }
        }
#pragma unroll
        for (int q = 0; q < A; ++q)
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                float v = gW3[q][c];
                v += __shfl_xor_sync(0xffffffffu, v, 4);
                v += __shfl_xor_sync(0xffffffffu, v, 8);
                v += __shfl_xor_sync(0xffffffffu, v, 16);
                if (g == 0) scr2[(q4 * A + q) * H + c0 + 8 * (c >> 1) + 2 * p + (c & 1)] = v;
            }
        {
            // row-owned scalars: every row's loss was computed by 4 threads; count one of them
            float vals[NS];
            const float keep = (hf == 0 && p < 2) ? 1.0f : 0.0f;
#pragma unroll
            for (int q = 0; q < A; ++q) vals[q] = gB3[q] * keep;
            vals[A] = racc.gLs * keep; vals[A + 1] = racc.st_loss * keep; vals[A + 2] = racc.st_kl * keep;
            vals[A + 3] = racc.st_ent * keep; vals[A + 4] = racc.st_clip * keep; vals[A + 5] = racc.st_rows * keep;
#pragma unroll
            for (int s = 0; s < NS; ++s) {
                float v = vals[s];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
                if (l == 0) scr3[s * 8 + w] = v;
            }
        }
        tc_before_sync();
        __syncthreads();
        for (int e = tid; e < H * H; e += NT) {
            const int m = e >> 6, k = e & 63;
            out[L.off_w2() + e] = scr[m * SW + k] + scr[(m + 64) * SW + k];
        }
        if (tid < H) {
            out[L.off_b2() + tid] = scr[tid * SW + 64] + scr[(tid + 64) * SW + 64];  // G[:, 64]  (ones column)
            out[L.off_b1() + tid] = scr[tid * SW + 72] + scr[(tid + 64) * SW + 72];  // F[:, 0]
        }
        for (int e = tid; e < H * O; e += NT) {
            const int m = e / O, o = e % O;
            out[L.off_w1() + e] = scr[m * SW + 73 + o] + scr[(m + 64) * SW + 73 + o];
        }
        for (int col = tid; col < A * H; col += NT)
            out[L.off_w3() + col] = (scr2[col] + scr2[A * H + col]) + (scr2[2 * A * H + col] + scr2[3 * A * H + col]);
        if (tid < NS) {
            const float sum = (scr3[tid * 8] + scr3[tid * 8 + 1]) + (scr3[tid * 8 + 2] + scr3[tid * 8 + 3]);
            if (tid < A) out[L.off_b3() + tid] = sum;
            else if (tid == A) out[L.off_logstd()] = sum;
            else if (with_stats) out[P + (tid - A - 1)] = sum;
        }
        if (with_stats && tid >= NS && tid < NS + (MLP_NSTATS - 5)) out[P + 5 + (tid - NS)] = 0.0f;
        __syncthreads();
        first = true;
        zero_acc();
        stamp(30);
    }

    // ---- forward only (baselines, logging passes): rows -> h1 -> GEMM1 -> h2 -> mu -> loss statistics.
    // Same panels, same GEMM1 and the same per-row loss as run_tiles(); nothing of the backward half.
    // Exactly one thread evaluates a row (warps 0..3, lanes p < 2), so racc holds every row once.
    __device__ __forceinline__ void run_tiles_fwd(const MlpArgs& a, int64_t tile0, int64_t stride) {
        const int64_t ntiles = (a.mb + R - 1) / R;
        const uint32_t tfrag = tmem + ((uint32_t)(32 * q4) << 16) + TM_D + c0;
        float w3c[A][8], b2c[8];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int c = c0 + 8 * i + 2 * p + e;
                b2c[2 * i + e] = sB2[c];
#pragma unroll
                for (int q = 0; q < A; ++q) w3c[q][2 * i + e] = sW3[q * H + c];
            }
        const bool evaluator = hf == 0 && p < 2;
        const int er = (p & 1) ? rb : ra;  // the row this thread evaluates (if evaluator)
        int it = 0;
        for (int64_t tile = tile0; tile < ntiles; tile += stride, ++it) {
            const uint32_t par = tiles_done & 1;
            const int64_t m0 = tile * R;
            if (tid < R) {
                sSrc[tid] = src_n;
                *reinterpret_cast<float4*>(sX + tid * 4) = make_float4(row_n.x[0], row_n.x[1], row_n.x[2], row_n.x[3]);
#pragma unroll
                for (int i = 0; i < NIN; ++i) sIn[tid * NIN + i] = row_n.in[i];
            }
            __syncthreads();
            if (tid < R) {
                src_n = sNext[(it & 1) * R + tid];
                row_n = load_row<O, A>(a, src_n);
            }
            {
                const int pn = w & 1, rbk = w >> 1, j = l & 7, rs = l >> 3, f0 = 32 * pn + 4 * j;
                float w1[4][O], bb[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    bb[q] = sB1[f0 + q];
#pragma unroll
                    for (int o = 0; o < O; ++o) w1[q][o] = sW1[(f0 + q) * O + o];
                }
                uint8_t* hp = smem + OFF_ACT + pn * PANEL;
                uint8_t* lp = smem + OFF_ACT + (3 + pn) * PANEL;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = 16 * rbk + rs + 4 * i;
                    const float4 xv = *reinterpret_cast<const float4*>(sX + r * 4);
                    const float x[4] = {xv.x, xv.y, xv.z, xv.w};
                    float h[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        float z = bb[q];
#pragma unroll
                        for (int o = 0; o < O; ++o) z = fmaf(w1[q][o], x[o], z);
                        h[q] = TC_TANH(z);
                    }
                    store_split4(hp, lp, chunk_off(r, j), h);
                }
            }
            fence_async_smem();
            __syncthreads();
            if (w == 0) {
                tc_after_sync();
                if (elect_one()) {
                    issue_gemm1(sb, tmem);
                    umma_commit(bar1);
                }
                __syncwarp();
            }
            if (tid >= R && tid < 2 * R) {
                const int64_t t2 = tile + 2 * stride;
                sNext[((it + 1) & 1) * R + (tid - R)] = t2 < ntiles ? mb_src(a, t2 * R + (tid - R)) : -1;
            }
            // the old policy's mean of the evaluated row (KL statistic), fetched behind the MMAs
            const int64_t esrc = evaluator ? sSrc[er] : -1;
            float om[A];
#pragma unroll
            for (int q = 0; q < A; ++q) om[q] = (esrc >= 0 && a.old_mean) ? a.old_mean[esrc * A + q] : 0.0f;
            mbar_wait(bar1, par);
            tc_after_sync();
            {
                float z[16];
                tmem_ld_frag(tfrag, z);
                float mpa[A], mpb[A];
#pragma unroll
                for (int q = 0; q < A; ++q) { mpa[q] = 0.0f; mpb[q] = 0.0f; }
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const float ha = TC_TANH(z[4 * i + e] + b2c[2 * i + e]);
                        const float hb = TC_TANH(z[4 * i + 2 + e] + b2c[2 * i + e]);
#pragma unroll
                        for (int q = 0; q < A; ++q) {
                            mpa[q] = fmaf(w3c[q][2 * i + e], ha, mpa[q]);
                            mpb[q] = fmaf(w3c[q][2 * i + e], hb, mpb[q]);
                        }
                    }
#pragma unroll
                for (int q = 0; q < A; ++q) {
                    mpa[q] += __shfl_xor_sync(0xffffffffu, mpa[q], 1);
                    mpb[q] += __shfl_xor_sync(0xffffffffu, mpb[q], 1);
                    mpa[q] += __shfl_xor_sync(0xffffffffu, mpa[q], 2);
                    mpb[q] += __shfl_xor_sync(0xffffffffu, mpb[q], 2);
                    if (p == 0) {
                        sMu[(hf * A + q) * R + ra] = mpa[q];
                        sMu[(hf * A + q) * R + rb] = mpb[q];
                    }
                }
            }
            tc_before_sync();  // the next tile's GEMM1 overwrites D: order these TMEM reads before it
            __syncthreads();
            if (evaluator && esrc >= 0) {
                float mu[A], dm[A];
                MlpRowIn<A> in;
#pragma unroll
                for (int q = 0; q < A; ++q) {
                    mu[q] = sB3[q] + sMu[q * R + er] + sMu[(A + q) * R + er];
                    dm[q] = 0.0f;
                    in.act[q] = sIn[er * NIN + q];
                    in.old_mean[q] = om[q];
                }
                in.adv = sIn[er * NIN + A]; in.old_logp = sIn[er * NIN + A + 1]; in.ret = sIn[er * NIN + A + 2];
                mlp_row_loss<A, true, true>(a, in, mu, dist, dm, racc);
                if (a.mean_out) {
#pragma unroll
                    for (int q = 0; q < A; ++q) a.mean_out[(m0 + er) * A + q] = mu[q];
                }
            }
            __syncthreads();  // sIn / sSrc / sMu are rewritten by the next tile
            ++tiles_done;
        }
    }

    // this CTA's statistics row: [P gradients (not written) | MLP_NSTATS statistics]
    __device__ __forceinline__ void write_stats(float* out) {
        const int P = MlpLayout{O, H, A}.total();
        float* scr3 = reinterpret_cast<float*>(smem + OFF_ACT);
        float vals[5] = {racc.st_loss, racc.st_kl, racc.st_ent, racc.st_clip, racc.st_rows};
#pragma unroll
        for (int s = 0; s < 5; ++s) {
            float v = vals[s];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
            if (l == 0) scr3[s * 8 + w] = v;
        }
        __syncthreads();
        if (tid < 5) out[P + tid] = (scr3[tid * 8] + scr3[tid * 8 + 1]) + (scr3[tid * 8 + 2] + scr3[tid * 8 + 3]);
        else if (tid < MLP_NSTATS) out[P + tid] = 0.0f;
    }
};

template <int O, int A>
__global__ void __maxnreg__(120) mlp_train_tc_kernel(const MlpArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const MlpLayout L{O, H, A};
    Cta<O, A> cta;
    cta.init(smem);
    cta.stamps = a.timing ? a.timing + (int64_t)blockIdx.x * 32 : nullptr;
    cta.stamp(0);
    if (a.step_counter && blockIdx.x == 0 && threadIdx.x == 0) *a.step_counter += 1ull;
    cta.prime(a, blockIdx.x, gridDim.x);
    cta.load_weights(a, a.params, [](const float* q) { return __ldg(q); });
    cta.zero_acc();
    cta.stamp(1);
    cta.run_tiles(a, blockIdx.x, gridDim.x);
    cta.write_partial(a.partial + (int64_t)blockIdx.x * (L.total() + MLP_NSTATS), true);
    cta.fini();
}

// forward-only launch: statistics of the logging passes / value baselines (mean_out)
template <int O, int A>
__global__ void __maxnreg__(120) mlp_eval_tc_kernel(const MlpArgs a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const MlpLayout L{O, H, A};
    Cta<O, A> cta;
    cta.init(smem);
    cta.prime(a, blockIdx.x, gridDim.x);
    cta.load_weights(a, a.params, [](const float* q) { return __ldg(q); });
    cta.zero_acc();
    cta.run_tiles_fwd(a, blockIdx.x, gridDim.x);
    cta.write_stats(a.partial + (int64_t)blockIdx.x * (L.total() + MLP_NSTATS));
    cta.fini();
}
}  // namespace tc
