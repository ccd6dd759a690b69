// Instantiations of the general-shape kernels for ONE hidden size (GEN_H, defined by the including .cu):
// observation capacities {4, 8} x action capacities {1, 2, 4} x {FWD, BWD, JVP}.  One translation unit
// per hidden size so that the three compile in parallel.
#include "mlp_general.cuh"

namespace {
template <int OP, int AP, int MODE>
int gen_occupancy() {
    static int occ = 0;
    if (occ == 0) {
        constexpr size_t smem = gen::smem_bytes<GEN_H, OP, AP, MODE>();
        cudaFuncSetAttribute(gen::mlp_general_kernel<GEN_H, OP, AP, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, gen::mlp_general_kernel<GEN_H, OP, AP, MODE>, gen::NT, smem) !=
                cudaSuccess || n < 1)
            n = 1;
        occ = n;
    }
    return occ;
}
template <int OP, int AP, int MODE>
int gen_grid(int64_t rows) {
    constexpr int R = gen::Cfg<GEN_H>::R;
    const int64_t tiles = (rows + R - 1) / R;
    const int64_t cap = (int64_t)b2g_sm_count() * gen_occupancy<OP, AP, MODE>();
    return (int)(tiles < cap ? (tiles > 0 ? tiles : 1) : cap);
}
template <int OP, int AP, int MODE>
int gen_launch(const MlpArgs& a, cudaStream_t s) {
    constexpr size_t smem = gen::smem_bytes<GEN_H, OP, AP, MODE>();
    const int grid = gen_grid<OP, AP, MODE>(a.mb);
    gen::mlp_general_kernel<GEN_H, OP, AP, MODE><<<grid, gen::NT, smem, s>>>(a);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
}  // namespace

#define GEN_CAT2(a, b) a##b
#define GEN_CAT(a, b) GEN_CAT2(a, b)
// mode, capacities -> launch / grid; the caller (mlp.cu) has checked O <= 8, A <= 4
#define GEN_FOR_CAPS(X) X(4, 1) X(4, 2) X(4, 4) X(8, 1) X(8, 2) X(8, 4)

int GEN_CAT(b2g_gen_launch_h, GEN_H)(int mode, int op, int ap, const MlpArgs& a, cudaStream_t s) {
#define GEN_CASE(OO, AA)                                                     \
    if (op == OO && ap == AA) {                                              \
        if (mode == gen::M_FWD) return gen_launch<OO, AA, gen::M_FWD>(a, s); \
        if (mode == gen::M_BWD) return gen_launch<OO, AA, gen::M_BWD>(a, s); \
        return gen_launch<OO, AA, gen::M_JVP>(a, s);                         \
    }
    GEN_FOR_CAPS(GEN_CASE)
#undef GEN_CASE
    b2g_set_error("general MLP kernels: capacity (%d, %d) is not built", op, ap);
    return B2G_ERR_UNSUPPORTED;
}

int GEN_CAT(b2g_gen_grid_h, GEN_H)(int mode, int op, int ap, int64_t rows) {
#define GEN_CASE(OO, AA)                                                     \
    if (op == OO && ap == AA) {                                              \
        if (mode == gen::M_FWD) return gen_grid<OO, AA, gen::M_FWD>(rows);   \
        if (mode == gen::M_BWD) return gen_grid<OO, AA, gen::M_BWD>(rows);   \
        return gen_grid<OO, AA, gen::M_JVP>(rows);                           \
    }
    GEN_FOR_CAPS(GEN_CASE)
#undef GEN_CASE
    return -1;
}
