// Shared by the persistent training loops (train_loop.cu: tcgen05 tiles, hidden 64;
// train_loop_general_*.cu: CUDA-core tiles, any built shape): per-network launch arguments and the
// device-wide barrier over the CTAs of one network.
#pragma once
#include "mlp_train.cuh"

struct NetLoop {
    MlpArgs a;  // data pointers + loss configuration (params/idx/start/mb/partial are set per step)
    float* params;
    float* adam_m;
    float* adam_v;
    unsigned long long* step;
    double lr, beta1, beta2, eps;
    int64_t B, mbs;
    int epochs, shuffle;  // shuffle: 0 sequential, 1 device permutation, 2 index lists [epochs][B]
    const int32_t* idx_all;
    uint32_t seed_lo, seed_hi;
    float* partial;       // [ctas][P + MLP_NSTATS]
    unsigned int* sync;   // barrier counter, zero at launch
    // multi-GPU (one process per GPU): peer-mapped exchange buffers, see b2g_mlp_train_loop_dist
    unsigned long long* peer_recv[B2G_MAX_RANKS];  // rank r's receive area of this network: [src rank][parity][COMM_P] slots
    int world, rank;
    unsigned int* seq_ptr;                  // device counter: minibatch steps this network has exchanged so far
    long long* timing;    // debug: [ctas][16] clock64 stamps of minibatch step `timing_step` (or NULL)
    long long* tile_timing;  // debug: [ctas][32] stage stamps of the tiles of that step
    int timing_step;
};
struct LoopArgs {
    NetLoop net[2];
    int n_nets, ctas_per_net;
    int stagger_cycles;  // the second network starts this many cycles late (see b2g_mlp_train_loop)
};

// Barrier over the CTAs of one network (all co-resident): arrive = one release-reduction (no reply
// to wait for), wait = acquire loads until the count reaches `target`.  bar.sync orders the other
// threads' writes before thread 0's release and thread 0's acquire before their later reads.
// The two halves are separate calls so that work which does not depend on the other CTAs (the rows
// of the next minibatch) runs between "my partial row is published" and "everybody's is".
__device__ __forceinline__ void net_barrier_arrive(unsigned int* ctr) {
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
}
__device__ __forceinline__ void net_barrier_wait(unsigned int* ctr, unsigned int target) {
    if (threadIdx.x == 0) {
        unsigned int v;
        unsigned int spin = 0;
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
            if (++spin > (1u << 22)) __trap();  // a lost CTA must surface as an error, not as a hung GPU
        } while (v < target);
    }
    __syncthreads();
}
__device__ __forceinline__ void net_barrier(unsigned int* ctr, unsigned int target) {
    net_barrier_arrive(ctr);
    net_barrier_wait(ctr, target);
}

