// Stand-alone probe of the tcgen05 (UMMA) semantics the minibatch kernel relies on:
// no-swizzle shared-memory descriptors (K-major and MN-major views of ONE buffer),
// kind::tf32, M=128 and M=64 accumulator lane mapping, N=8/64/72, 3xTF32 accuracy.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/umma_probe tools/umma_probe.cu
// Each case copies host-built smem images, issues nk MMAs, dumps TMEM [128 lanes x ncols].
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>

struct Mma {
    uint32_t a_off, b_off;   // byte offsets inside the smem image
    uint32_t a_lbo, a_sbo, b_lbo, b_sbo;
    uint32_t idesc, d_col, accumulate;
    uint32_t a_layout, b_layout;
};
struct Case {
    int n_mma;
    int ncols;
    uint32_t image_bytes;
};
constexpr int MAX_MMA = 192;
__constant__ Mma c_mma[MAX_MMA];

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout = 0) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;  // descriptor version (Blackwell)
    d |= (uint64_t)(layout & 7) << 61;
    return d;                // layout_type 0 = no swizzle, base_offset 0
}

__global__ void __launch_bounds__(128, 1) probe_kernel(const uint8_t* image, uint32_t image_bytes, int n_mma,
                                                       int ncols, float* out /*[128][ncols]*/) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, w = tid >> 5;
    for (uint32_t i = tid; i < image_bytes / 16; i += 128)
        reinterpret_cast<uint4*>(smem)[i] = reinterpret_cast<const uint4*>(image)[i];
    const uint32_t bar_addr = (uint32_t)__cvta_generic_to_shared(&bar);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_addr));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (w == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(
            (uint32_t)__cvta_generic_to_shared(&tmem_base_s)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;");  // generic-proxy smem writes -> async proxy (UMMA)
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = tmem_base_s;
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);
    if (tid == 0) {
        if (sbase & 1023) printf("smem base not 1024-aligned: %u\n", sbase);
        for (int i = 0; i < n_mma; ++i) {
            const Mma m = c_mma[i];
            const uint64_t da = make_desc(sbase + m.a_off, m.a_lbo, m.a_sbo, m.a_layout);
            const uint64_t db = make_desc(sbase + m.b_off, m.b_lbo, m.b_sbo, m.b_layout);
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem + m.d_col),
                "l"(da), "l"(db), "r"(m.idesc), "r"(m.accumulate));
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_addr));
    }
    // wait phase 0
    {
        uint32_t done = 0;
        while (!done) {
            asm volatile(
                "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t}\n"
                : "=r"(done)
                : "r"(bar_addr), "r"(0u));
        }
    }
    asm volatile("tcgen05.fence::after_thread_sync;");
    for (int c0 = 0; c0 < ncols; c0 += 8) {
        uint32_t v[8];
        const uint32_t taddr = tmem + ((uint32_t)(w * 32) << 16) + c0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                     : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;");
        for (int j = 0; j < 8; ++j) out[tid * ncols + c0 + j] = __uint_as_float(v[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (w == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem));
}

// ------------------------------------------------------------------ host --
static uint32_t idesc_tf32(int M, int N, int a_mn, int b_mn) {
    uint32_t d = 0;
    d |= 1u << 4;           // D format f32
    d |= 2u << 7;           // A tf32
    d |= 2u << 10;          // B tf32
    d |= (uint32_t)a_mn << 15;
    d |= (uint32_t)b_mn << 16;
    d |= (uint32_t)(N >> 3) << 17;
    d |= (uint32_t)(M >> 4) << 24;
    return d;
}
static float tf32_trunc(float x) {
    uint32_t u;
    memcpy(&u, &x, 4);
    u &= 0xFFFFE000u;
    memcpy(&x, &u, 4);
    return x;
}
// activation-style buffer: element (row, col) of an [R x C] matrix at float index ((col/4)*R + row)*4 + col%4
static void put(std::vector<float>& img, size_t base_f, int R, int row, int col, float v) {
    img[base_f + ((size_t)(col / 4) * R + row) * 4 + (col % 4)] = v;
}
static float frand() { return (float)rand() / RAND_MAX * 2.0f - 1.0f; }

static int run_case(const char* name, const std::vector<float>& img, const std::vector<Mma>& mm, int ncols,
                    std::vector<float>& out) {
    uint8_t* d_img;
    float* d_out;
    cudaMalloc(&d_img, img.size() * 4);
    cudaMalloc(&d_out, 128 * ncols * 4);
    cudaMemcpy(d_img, img.data(), img.size() * 4, cudaMemcpyHostToDevice);
    cudaMemset(d_out, 0, 128 * ncols * 4);
    cudaMemcpyToSymbol(c_mma, mm.data(), mm.size() * sizeof(Mma));
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    probe_kernel<<<1, 128, img.size() * 4>>>(d_img, (uint32_t)(img.size() * 4), (int)mm.size(), ncols, d_out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        printf("[%s] CUDA error: %s\n", name, cudaGetErrorString(e));
        return 1;
    }
    out.resize(128 * ncols);
    cudaMemcpy(out.data(), d_out, out.size() * 4, cudaMemcpyDeviceToHost);
    cudaFree(d_img);
    cudaFree(d_out);
    return 0;
}

static void report(const char* name, double maxerr, double maxref, double tol) {
    printf("[%s] max_abs_err %.3e (max |ref| %.3e) -> %s\n", name, maxerr, maxref, maxerr <= tol ? "OK" : "FAIL");
}


// decode mode: which smem float index does the hardware read as A[m][k] (MN-major A) for given LBO/SBO?
static void decode_a(uint32_t lbo, uint32_t sbo, int a_mn, int layout = 0, int full = 0) {
    const size_t NF = 16384;               // A image floats (64 KB)
    const size_t bb = NF;                  // B image behind it: K-major identity [8 x 8]
    int idx[128][8];
    for (int pass = 0; pass < 2; ++pass) {
        std::vector<float> img(NF + 64, 0.0f);
        for (size_t i = 0; i < NF; ++i) img[i] = pass == 0 ? (float)(i & 1023) : (float)(i >> 10);
        for (int n = 0; n < 8; ++n)
            for (int k = 0; k < 8; ++k) put(img, bb, 8, n, k, n == k ? 1.0f : 0.0f);
        std::vector<Mma> mm;
        Mma m{};
        m.a_off = 0; m.a_lbo = lbo; m.a_sbo = sbo; m.a_layout = layout;
        m.b_off = bb * 4; m.b_lbo = 128; m.b_sbo = 128;
        m.idesc = idesc_tf32(128, 8, a_mn, 0); m.d_col = 0; m.accumulate = 0;
        mm.push_back(m);
        std::vector<float> out;
        if (run_case("decode", img, mm, 8, out)) exit(1);
        for (int mm_ = 0; mm_ < 128; ++mm_)
            for (int k = 0; k < 8; ++k) {
                const int v = (int)out[mm_ * 8 + k];
                if (pass == 0) idx[mm_][k] = v; else idx[mm_][k] += v << 10;
            }
    }
    printf("decode A %s layout=%d lbo=%u sbo=%u : float index read as A[m][k]\n", a_mn ? "MN-major" : "K-major", layout, lbo, sbo);
    std::vector<int> ms = {0, 1, 3, 4, 5, 8, 12, 16, 127};
    if (full) { ms.clear(); for (int i = 0; i < 128; ++i) ms.push_back(i); }
    for (int mi : ms) {
        printf("  m=%3d:", mi);
        for (int k = 0; k < 8; ++k) printf(" %6d", idx[mi][k]);
        printf("\n");
    }
}


// ---- panel layout: [feat/32][row][128 B], 32-byte chunks XOR-swizzled by row%4 (descriptor layout_type 1).
// ONE buffer serves the K-major view (M/N = row, K = feat) and the MN-major view (M/N = feat, K = row).
static size_t pidx(int R, int row, int feat) {
    return (size_t)(feat / 32) * R * 32 + (size_t)row * 32 + ((((feat % 32) / 8) ^ (row % 4)) * 8) + feat % 8;
}
static int panel_tests() {
    const int R = 128, H = 64;
    srand(7);
    // buffers (float offsets): ACT = 5 panels [hi0 hi1 aug lo0 lo1] of R rows; DZ = 4 panels [hi0 hi1 lo0 lo1]; W = 2 panels of 64 rows
    const size_t ACT = 0, DZ = ACT + 5 * (size_t)R * 32, W = DZ + 4 * (size_t)R * 32, TOT = W + 2 * 64 * 32;
    std::vector<float> img(TOT, 0.0f), Hm(R * 96), Z(R * 128), Wm(H * H);
    for (auto& v : Hm) v = tf32_trunc(frand());
    for (auto& v : Z) v = tf32_trunc(frand());
    for (auto& v : Wm) v = tf32_trunc(frand());
    for (int r = 0; r < R; ++r) {
        for (int f = 0; f < 96; ++f) img[ACT + pidx(R, r, f)] = Hm[r * 96 + f];   // act "features" 0..95 (64..95 = aug panel)
        for (int f = 0; f < 128; ++f) img[DZ + pidx(R, r, f)] = Z[r * 128 + f];
    }
    for (int c = 0; c < H; ++c)
        for (int k = 0; k < H; ++k) img[W + pidx(64, c, k)] = Wm[c * H + k];
    std::vector<float> out;
    std::vector<Mma> mm;
    double me = 0, mr = 0;
    const int kmax = getenv("PROBE_KMAX") ? atoi(getenv("PROBE_KMAX")) : 8;
    const int kstart = getenv("PROBE_KSTART") ? atoi(getenv("PROBE_KSTART")) : 0;
    const int only = getenv("PROBE_ONLY") ? atoi(getenv("PROBE_ONLY")) : 0;
    const int blay = getenv("PROBE_BLAY") ? atoi(getenv("PROBE_BLAY")) : 1;
    if (only > 1) goto g2;
    // G1: D[r][c] = sum_k Hm[r][k] W[c][k]
    for (int ks = kstart; ks < kmax; ++ks) {
        Mma m{};
        m.a_off = (ACT + (ks / 4) * R * 32) * 4 + (ks % 4) * 32; m.a_lbo = 1024; m.a_sbo = 512; m.a_layout = 1;
        m.b_off = (W + (ks / 4) * 64 * 32) * 4 + (ks % 4) * 32; m.b_lbo = 1024; m.b_sbo = 512; m.b_layout = blay;
        m.idesc = idesc_tf32(128, 64, 0, 0); m.d_col = 0; m.accumulate = ks > kstart;
        mm.push_back(m);
    }
    if (run_case("G1", img, mm, 64, out)) return 1;
    me = mr = 0;
    for (int r = 0; r < R; ++r)
        for (int c = 0; c < H; ++c) {
            double ref = 0;
            for (int k = kstart * 8; k < kmax * 8; ++k) ref += (double)Hm[r * 96 + k] * Wm[c * H + k];
            me = fmax(me, fabs(ref - out[r * 64 + c])); mr = fmax(mr, fabs(ref));
        }
    report("G1 panel K/K M128 N64", me, mr, 1e-4);
    if (only == 1) return 0;
g2:
    if (only > 2) goto g3;
    // G2: D[r][k] = sum_c Z[r][c] W[c][k]  (A = DZ hi panels K-major, B = W MN-major: mn = k, kk = c)
    mm.clear();
    for (int ks = 0; ks < 8; ++ks) {
        Mma m{};
        m.a_off = (DZ + (ks / 4) * R * 32) * 4 + (ks % 4) * 32; m.a_lbo = 1024; m.a_sbo = 512; m.a_layout = 1;
        m.b_off = W * 4 + ks * 1024; m.b_lbo = 64 * 128; m.b_sbo = 512; m.b_layout = 1;
        m.idesc = idesc_tf32(128, 64, 0, 1); m.d_col = 0; m.accumulate = ks > 0;
        mm.push_back(m);
    }
    if (run_case("G2", img, mm, 64, out)) return 1;
    me = mr = 0;
    for (int r = 0; r < R; ++r)
        for (int k = 0; k < H; ++k) {
            double ref = 0;
            for (int c = 0; c < H; ++c) ref += (double)Z[r * 128 + c] * Wm[c * H + k];
            me = fmax(me, fabs(ref - out[r * 64 + k])); mr = fmax(mr, fabs(ref));
        }
    report("G2 panel K/MN M128 N64", me, mr, 1e-4);
    if (only == 2) return 0;
g3:
    // G3: D[m][n] = sum_r Z[r][m] Hm[r][n]  m < 128 (stacked hi|lo), n < 72
    mm.clear();
    for (int ks = 0; ks < 16; ++ks) {
        Mma m{};
        m.a_off = DZ * 4 + ks * 1024; m.a_lbo = R * 128; m.a_sbo = 512; m.a_layout = 1;
        m.b_off = ACT * 4 + ks * 1024; m.b_lbo = R * 128; m.b_sbo = 512; m.b_layout = 1;
        m.idesc = idesc_tf32(128, 72, 1, 1); m.d_col = 64; m.accumulate = ks > 0;
        mm.push_back(m);
    }
    // G4 in the same launch: D[m][n] = sum_r Z[r][m] Hm[r][64 + 8 + n], n < 8 (start offset 32 B inside the aug panel)
    for (int ks = 0; ks < 16; ++ks) {
        Mma m{};
        m.a_off = DZ * 4 + ks * 1024; m.a_lbo = R * 128; m.a_sbo = 512; m.a_layout = 1;
        m.b_off = (ACT + 2 * R * 32) * 4 + 32 + ks * 1024; m.b_lbo = R * 128; m.b_sbo = 512; m.b_layout = 1;
        m.idesc = idesc_tf32(128, 8, 1, 1); m.d_col = 136; m.accumulate = ks > 0;
        mm.push_back(m);
    }
    if (run_case("G3", img, mm, 144, out)) return 1;
    me = mr = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 72; ++n) {
            double ref = 0;
            for (int r = 0; r < R; ++r) ref += (double)Z[r * 128 + m] * Hm[r * 96 + n];
            me = fmax(me, fabs(ref - out[m * 144 + 64 + n])); mr = fmax(mr, fabs(ref));
        }
    report("G3 panel MN/MN M128(stacked) N72", me, mr, 1e-4);
    me = mr = 0;
    for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 8; ++n) {
            double ref = 0;
            for (int r = 0; r < R; ++r) ref += (double)Z[r * 128 + m] * Hm[r * 96 + 72 + n];
            me = fmax(me, fabs(ref - out[m * 144 + 136 + n])); mr = fmax(mr, fabs(ref));
        }
    report("G4 panel MN/MN M128 N8 (+32 B start)", me, mr, 1e-4);
    return 0;
}

int main(int argc, char** argv) {
    if (argc > 1 && !strcmp(argv[1], "panel")) return panel_tests();
    if (argc > 5 && !strcmp(argv[1], "decode1")) { decode_a(atoi(argv[2]), atoi(argv[3]), atoi(argv[4]), atoi(argv[5])); return 0; }
    if (argc > 1 && !strcmp(argv[1], "decode")) {
        decode_a(4096, 8192, 1, 1);   // MN-major, 128B swizzle with 32B atoms
        decode_a(4096, 1024, 0, 1);   // K-major with layout_type 1: legal?
        decode_a(4096, 1024, 0, 2);   // K-major SW128
        decode_a(4096, 8192, 1, 2);   // MN-major SW128
        decode_a(4096, 8192, 1, 0);   // MN-major no swizzle
        return 0;
    }
    const int R = 128, H = 64;
    srand(1);
    // ---------------- case 1: GEMM1 form. A = act [R x H] K-major, B = W [H(n) x H(k)] K-major
    // and case 2: GEMM2 form: A = act K-major, B = W viewed MN-major (N = k index of W, K = c index)
    {
        std::vector<float> img((size_t)R * H + H * H);
        std::vector<float> A(R * H), W(H * H);
        for (auto& v : A) v = tf32_trunc(frand());
        for (auto& v : W) v = tf32_trunc(frand());
        const size_t wb = (size_t)R * H;
        for (int r = 0; r < R; ++r)
            for (int k = 0; k < H; ++k) put(img, 0, R, r, k, A[r * H + k]);
        for (int c = 0; c < H; ++c)
            for (int k = 0; k < H; ++k) put(img, wb, H, c, k, W[c * H + k]);
        std::vector<Mma> mm;
        for (int ks = 0; ks < H / 8; ++ks) {
            Mma m{};
            m.a_off = ks * 2 * R * 16; m.a_lbo = R * 16; m.a_sbo = 128;
            m.b_off = wb * 4 + ks * 2 * H * 16; m.b_lbo = H * 16; m.b_sbo = 128;
            m.idesc = idesc_tf32(128, 64, 0, 0); m.d_col = 0; m.accumulate = ks > 0;
            mm.push_back(m);
        }
        std::vector<float> out;
        if (run_case("gemm1 K/K M128 N64", img, mm, 64, out)) return 1;
        double me = 0, mr = 0;
        for (int r = 0; r < R; ++r)
            for (int c = 0; c < H; ++c) {
                double ref = 0;
                for (int k = 0; k < H; ++k) ref += (double)A[r * H + k] * W[c * H + k];
                me = fmax(me, fabs(ref - out[r * 64 + c])); mr = fmax(mr, fabs(ref));
            }
        report("gemm1 K/K M128 N64", me, mr, 1e-4);

        // case 2: D[r][k] = sum_c A[r][c] W[c][k]; B (n=k, kk=c) MN-major view of the same W buffer:
        // element (mn=k, kk=c) at (k/4)*H*16 + c*16 + (k%4)*4 -> SBO (mn groups) = H*16, LBO (k groups of 8) = 128
        mm.clear();
        for (int ks = 0; ks < H / 8; ++ks) {
            Mma m{};
            m.a_off = ks * 2 * R * 16; m.a_lbo = R * 16; m.a_sbo = 128;
            m.b_off = wb * 4 + ks * 128; m.b_lbo = 128; m.b_sbo = H * 16;
            m.idesc = idesc_tf32(128, 64, 0, 1); m.d_col = 0; m.accumulate = ks > 0;
            mm.push_back(m);
        }
        if (run_case("gemm2 K/MN M128 N64", img, mm, 64, out)) return 1;
        me = mr = 0;
        for (int r = 0; r < R; ++r)
            for (int k = 0; k < H; ++k) {
                double ref = 0;
                for (int c = 0; c < H; ++c) ref += (double)A[r * H + c] * W[c * H + k];
                me = fmax(me, fabs(ref - out[r * 64 + k])); mr = fmax(mr, fabs(ref));
            }
        report("gemm2 K/MN M128 N64", me, mr, 1e-4);
    }
    // ---------------- case 3: GEMM3 form. D[c][k] = sum_r Z[r][c] Haug[r][k]; A = Z MN-major (M=c=64, K=r),
    // B = Haug MN-major (N=72, K=r). M=64: row m -> lane (m%16) + 32*(m/16)
    {
        const int NB = 72;
        std::vector<float> img((size_t)R * H + (size_t)R * NB);
        std::vector<float> Z(R * H), Hh(R * NB);
        for (auto& v : Z) v = tf32_trunc(frand());
        for (auto& v : Hh) v = tf32_trunc(frand());
        const size_t hb = (size_t)R * H;
        for (int r = 0; r < R; ++r) {
            for (int c = 0; c < H; ++c) put(img, 0, R, r, c, Z[r * H + c]);
            for (int k = 0; k < NB; ++k) put(img, hb, R, r, k, Hh[r * NB + k]);
        }
        std::vector<Mma> mm;
        for (int ks = 0; ks < R / 8; ++ks) {
            Mma m{};
            m.a_off = ks * 128; m.a_lbo = 128; m.a_sbo = R * 16;
            m.b_off = hb * 4 + ks * 128; m.b_lbo = 128; m.b_sbo = R * 16;
            m.idesc = idesc_tf32(64, NB, 1, 1); m.d_col = 64; m.accumulate = ks > 0;
            mm.push_back(m);
        }
        std::vector<float> out;
        if (run_case("gemm3 MN/MN M64 N72", img, mm, 144, out)) return 1;
        double me = 0, mr = 0;
        for (int c = 0; c < H; ++c)
            for (int k = 0; k < NB; ++k) {
                double ref = 0;
                for (int r = 0; r < R; ++r) ref += (double)Z[r * H + c] * Hh[r * NB + k];
                const int lane = (c % 16) + 32 * (c / 16);
                me = fmax(me, fabs(ref - out[lane * 144 + 64 + k])); mr = fmax(mr, fabs(ref));
            }
        report("gemm3 MN/MN M64 N72 (lane = m%16 + 32*(m/16))", me, mr, 1e-4);

        // case 3b: stacked M=128: A rows 0..63 = Z cols, rows 64..127 = second buffer Z2 directly behind it
    }
    {
        const int NB = 72;
        std::vector<float> img((size_t)2 * R * H + (size_t)R * NB);
        std::vector<float> Z(R * 2 * H), Hh(R * NB);
        for (auto& v : Z) v = tf32_trunc(frand());
        for (auto& v : Hh) v = tf32_trunc(frand());
        const size_t hb = (size_t)2 * R * H;
        for (int r = 0; r < R; ++r) {
            for (int c = 0; c < 2 * H; ++c) put(img, 0, R, r, c, Z[r * 2 * H + c]);
            for (int k = 0; k < NB; ++k) put(img, hb, R, r, k, Hh[r * NB + k]);
        }
        std::vector<Mma> mm;
        for (int ks = 0; ks < R / 8; ++ks) {
            Mma m{};
            m.a_off = ks * 128; m.a_lbo = 128; m.a_sbo = R * 16;
            m.b_off = hb * 4 + ks * 128; m.b_lbo = 128; m.b_sbo = R * 16;
            m.idesc = idesc_tf32(128, NB, 1, 1); m.d_col = 64; m.accumulate = ks > 0;
            mm.push_back(m);
        }
        std::vector<float> out;
        if (run_case("gemm3b MN/MN M128 N72", img, mm, 144, out)) return 1;
        double me = 0, mr = 0;
        for (int c = 0; c < 2 * H; ++c)
            for (int k = 0; k < NB; ++k) {
                double ref = 0;
                for (int r = 0; r < R; ++r) ref += (double)Z[r * 2 * H + c] * Hh[r * NB + k];
                me = fmax(me, fabs(ref - out[c * 144 + 64 + k])); mr = fmax(mr, fabs(ref));
            }
        report("gemm3b MN/MN M128 N72 (stacked hi|lo)", me, mr, 1e-4);
    }
    // ---------------- case 4: GEMM4 form. D[k][o] = sum_r Z[r][k] X[r][o]; N = 8, M=128 stacked
    {
        const int NB = 8;
        std::vector<float> img((size_t)2 * R * H + (size_t)R * NB);
        std::vector<float> Z(R * 2 * H), X(R * NB);
        for (auto& v : Z) v = tf32_trunc(frand());
        for (auto& v : X) v = tf32_trunc(frand());
        const size_t xb = (size_t)2 * R * H;
        for (int r = 0; r < R; ++r) {
            for (int c = 0; c < 2 * H; ++c) put(img, 0, R, r, c, Z[r * 2 * H + c]);
            for (int k = 0; k < NB; ++k) put(img, xb, R, r, k, X[r * NB + k]);
        }
        std::vector<Mma> mm;
        for (int ks = 0; ks < R / 8; ++ks) {
            Mma m{};
            m.a_off = ks * 128; m.a_lbo = 128; m.a_sbo = R * 16;
            m.b_off = xb * 4 + ks * 128; m.b_lbo = 128; m.b_sbo = R * 16;
            m.idesc = idesc_tf32(128, NB, 1, 1); m.d_col = 136; m.accumulate = ks > 0;
            mm.push_back(m);
        }
        std::vector<float> out;
        if (run_case("gemm4 MN/MN M128 N8", img, mm, 144, out)) return 1;
        double me = 0, mr = 0;
        for (int c = 0; c < 2 * H; ++c)
            for (int k = 0; k < NB; ++k) {
                double ref = 0;
                for (int r = 0; r < R; ++r) ref += (double)Z[r * 2 * H + c] * X[r * NB + k];
                me = fmax(me, fabs(ref - out[c * 144 + 136 + k])); mr = fmax(mr, fabs(ref));
            }
        report("gemm4 MN/MN M128 N8", me, mr, 1e-4);
    }
    // ---------------- case 5: 3xTF32 accuracy on GEMM1 form with full fp32 inputs
    {
        std::vector<float> img((size_t)2 * R * H + 2 * H * H);
        std::vector<float> A(R * H), W(H * H);
        for (auto& v : A) v = frand();
        for (auto& v : W) v = frand() * 0.3f;
        const size_t alo = (size_t)R * H, whi = 2 * (size_t)R * H, wlo = whi + H * H;
        for (int r = 0; r < R; ++r)
            for (int k = 0; k < H; ++k) {
                const float hi = tf32_trunc(A[r * H + k]);
                put(img, 0, R, r, k, hi);
                put(img, alo, R, r, k, A[r * H + k] - hi);
            }
        for (int c = 0; c < H; ++c)
            for (int k = 0; k < H; ++k) {
                const float hi = tf32_trunc(W[c * H + k]);
                put(img, whi, H, c, k, hi);
                put(img, wlo, H, c, k, W[c * H + k] - hi);
            }
        for (int npass = 1; npass <= 3; npass += 2) {
            std::vector<Mma> mm;
            const size_t ao[3] = {0, alo, 0}, bo[3] = {whi, whi, wlo};
            for (int p = 0; p < npass; ++p)
                for (int ks = 0; ks < H / 8; ++ks) {
                    Mma m{};
                    m.a_off = ao[p] * 4 + ks * 2 * R * 16; m.a_lbo = R * 16; m.a_sbo = 128;
                    m.b_off = bo[p] * 4 + ks * 2 * H * 16; m.b_lbo = H * 16; m.b_sbo = 128;
                    m.idesc = idesc_tf32(128, 64, 0, 0); m.d_col = 0; m.accumulate = (p | ks) > 0;
                    mm.push_back(m);
                }
            std::vector<float> out;
            if (run_case("3xtf32", img, mm, 64, out)) return 1;
            double me = 0, mr = 0, mf = 0;
            for (int r = 0; r < R; ++r)
                for (int c = 0; c < H; ++c) {
                    double ref = 0;
                    float f32 = 0.0f;
                    for (int k = 0; k < H; ++k) {
                        ref += (double)A[r * H + k] * W[c * H + k];
                        f32 = fmaf(A[r * H + k], W[c * H + k], f32);
                    }
                    me = fmax(me, fabs(ref - out[r * 64 + c])); mr = fmax(mr, fabs(ref));
                    mf = fmax(mf, fabs(ref - (double)f32));
                }
            printf("[%dxTF32 gemm1] max_abs_err %.3e (max |ref| %.3e; plain fp32 FMA chain err %.3e)\n", npass, me, mr, mf);
        }
    }
    return 0;
}
