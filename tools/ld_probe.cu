// Micro-benchmark behind the MMA-cost numbers quoted in DESIGN.md / bench.py: how many SM cycles one
// tcgen05.mma instruction of a given shape occupies the tensor pipe when a single thread issues a
// long back-to-back sequence into one TMEM accumulator (shared-memory operands, cta_group::1), and the
// latency of the tcgen05.ld that reads an accumulator back.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -o tools/ld_probe tools/ld_probe.cu && tools/ld_probe
// Output: one line per (kind, M, N): cycles per instruction, MAC per cycle.
// Operands are zeros (the pipe does not look at the values); descriptors use the 128-byte-swizzle
// K-major layout of mlp_tc.cuh.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46) | (1ull << 61);
}
// kind::tf32: a/b format 2; kind::f16 with bf16 operands: format 1; fp32 accumulate
__host__ __device__ constexpr uint32_t idesc(int M, int N, int fmt) {
    return (1u << 4) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

template <bool TF32>
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t id, uint32_t acc) {
    if (TF32)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d),
                     "l"(a), "l"(b), "r"(id), "r"(acc) : "memory");
    else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d),
                     "l"(a), "l"(b), "r"(id), "r"(acc) : "memory");
}

struct Result { long long issue, total, ld; };

template <bool TF32, int NACC>
__global__ void __launch_bounds__(128, 1) probe(int M, int N, int reps, Result* out, int smem_kb = 160, int cols = 512) {
    out += 2 * blockIdx.x;
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tslot;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x;
    for (int i = tid; i < smem_kb * 1024 / 16; i += 128) reinterpret_cast<float4*>(smem)[i] = make_float4(0, 0, 0, 0);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid < 32) {
        if (cols == 512) asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tslot)) : "memory");
        else asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&tslot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tslot, sb = smem_u32(smem);
    long long t0 = 0, t1 = 0, t2 = 0;
    if (tid == 0) {
        const uint32_t id = idesc(M, N, TF32 ? 2 : 1);
        const uint64_t da = umma_desc(sb, 1024, 512), db = umma_desc(sb + 64 * 1024, 1024, 512);
        t0 = clock64();
        // nacc > 1: consecutive instructions accumulate into different TMEM column blocks (independent chains)
        for (int i = 0; i < reps; i += NACC) {
#pragma unroll
            for (int q = 0; q < NACC; ++q) mma<TF32>(tmem + (uint32_t)(q * N), da, db, id, i != 0);
        }
        t1 = clock64();
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                         : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
        t2 = clock64();
    }
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // latency of one 32x32b.x16 accumulator read (issue -> wait::ld), warp 0
    long long l0 = 0, l1 = 0;
    if (tid < 32) {
        uint32_t u[16];
        l0 = clock64();
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                     : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]), "=r"(u[8]),
                       "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15])
                     : "r"(tmem) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        l1 = clock64();
        if (u[0] == 0x12345678u) out[1].ld = 1;  // keep the load alive
    }
    if (tid == 0) { out->issue = t1 - t0; out->total = t2 - t0; out->ld = l1 - l0; }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid < 32) {
        if (cols == 512) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem) : "memory");
    }
}

template <bool TF32, int NACC = 1>
static void run(int M, int N, int reps) {
    const int nacc = NACC;
    Result* d;
    cudaMalloc(&d, 2 * sizeof(Result));
    cudaMemset(d, 0, 2 * sizeof(Result));
    cudaFuncSetAttribute(probe<TF32, NACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    probe<TF32, NACC><<<1, 128, 160 * 1024>>>(M, N, reps, d);
    probe<TF32, NACC><<<1, 128, 160 * 1024>>>(M, N, reps, d);
    Result r{};
    const cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(&r, d, sizeof(r), cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) { printf("kind %s M%d N%d: %s\n", TF32 ? "tf32" : "bf16", M, N, cudaGetErrorString(e)); exit(1); }
    const int K = TF32 ? 8 : 16;
    const double cyc = (double)r.total / reps;
    printf("kind %-4s M%-3d N%-3d K%-2d x %4d acc %d: %7.1f cycles / instruction (issue loop %6.1f), %7.1f MAC / cycle; tcgen05.ld 32x32b.x16 + wait %lld cycles\n",
           TF32 ? "tf32" : "bf16", M, N, K, reps, nacc, cyc, (double)r.issue / reps, (double)M * N * K / cyc, r.ld);
}

// two CTAs per SM (80 KB shared memory, 256 TMEM columns each, grid = 2 x SMs), each issuing its own chain: does the
// per-instruction cost of one CTA change when the other CTA of the SM issues at the same time?
template <bool TF32>
static void run_pair(int M, int N, int reps, int per_sm) {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int grid = per_sm * sms;
    Result* d;
    cudaMalloc(&d, 2 * grid * sizeof(Result));
    cudaMemset(d, 0, 2 * grid * sizeof(Result));
    cudaFuncSetAttribute(probe<TF32, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    for (int k = 0; k < 2; ++k) probe<TF32, 1><<<grid, 128, 80 * 1024>>>(M, N, reps, d, 80, 256);
    const cudaError_t e = cudaDeviceSynchronize();
    Result* h = (Result*)malloc(2 * grid * sizeof(Result));
    cudaMemcpy(h, d, 2 * grid * sizeof(Result), cudaMemcpyDeviceToHost);
    cudaFree(d);
    if (e != cudaSuccess) { printf("pair M%d N%d: %s\n", M, N, cudaGetErrorString(e)); exit(1); }
    double sum = 0, mx = 0;
    for (int b = 0; b < grid; ++b) { const double c = (double)h[2 * b].total / reps; sum += c; if (c > mx) mx = c; }
    free(h);
    printf("kind %-4s M%-3d N%-3d x %4d, %d CTA(s) per SM issuing at once: mean %7.1f max %7.1f cycles / instruction per CTA\n",
           TF32 ? "tf32" : "bf16", M, N, reps, per_sm, sum / grid, mx);
}

int main() {
    const int shapes[][2] = {{64, 32}, {64, 64}, {64, 128}, {128, 16}, {128, 64}, {128, 128}, {128, 144}, {128, 160}, {128, 256}};
    for (auto& s : shapes) run<true>(s[0], s[1], 256);
    for (auto& s : shapes) run<false>(s[0], s[1], 256);
    // independent accumulators: is the ~47-cycle floor a dependent-accumulate latency or an issue cost?
    run<true, 2>(64, 64, 256); run<true, 2>(128, 64, 256); run<true, 2>(64, 16, 256); run<true, 2>(128, 128, 256);
    run<true, 3>(64, 64, 252); run<true, 3>(128, 64, 252); run<true, 3>(64, 16, 252); run<true, 3>(128, 128, 252);
    run<true, 4>(64, 64, 256); run<true, 4>(128, 64, 256); run<true, 4>(64, 16, 256); run<true, 4>(128, 128, 256);
    for (int per_sm : {1, 2}) {
        run_pair<true>(64, 64, 2048, per_sm);
        run_pair<true>(64, 16, 2048, per_sm);
        run_pair<true>(128, 64, 2048, per_sm);
        run_pair<true>(128, 128, 2048, per_sm);
    }
    return 0;
}
