// EpisodeBatch assembly on the device: time-major [T,N] rollout buffers ->
// garage's episode-major packed layout [sum(lengths), ...] + lengths.
//   replaces sampler/vec_worker.py:139-174,206-219 (_gather_episode /
//   collect_episode), sampler/local_sampler.py:156-166 (the ">= num_samples,
//   whole episodes only" cut) and _dtypes.py:591-632 (EpisodeBatch.concatenate).
// Episode order = completion order, ties in env-index order, i.e. row-major
// order of the TERMINAL/TIMEOUT marks in step_type[T,N]; the batch ends with the
// first row t* at which the cumulative number of completed steps >= num_samples
// (all episodes finishing in row t* are included, like the reference's
// per-rollout() granularity).  Integer outputs are bit-exact by construction.
#include "common.cuh"

// workspace (int64): row_eps[T] | row_steps[T] | row_ep_off[T] | row_step_off[T]
extern "C" int64_t b2g_pack_ws_bytes(int T) { return (int64_t)4 * T * sizeof(int64_t); }

__device__ __forceinline__ int64_t block_sum_i64(int64_t v, int64_t* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        v = (l < (blockDim.x >> 5)) ? sh[l] : 0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    }
    return v;  // thread 0
}

// one block per row: number of episodes ending in the row and their total length
__global__ void __launch_bounds__(256)
pack_row_stats_kernel(const uint8_t* __restrict__ st, const int32_t* __restrict__ ep_len,
                      int64_t N, int64_t* __restrict__ ws, int T) {
    __shared__ int64_t sh[32];
    const int t = blockIdx.x;
    int64_t ne = 0, ns = 0;
    for (int64_t n = threadIdx.x; n < N; n += blockDim.x) {
        const int64_t i = (int64_t)t * N + n;
        if (st[i] >= ST_TERMINAL) { ne += 1; ns += ep_len[i]; }
    }
    ne = block_sum_i64(ne, sh);
    ns = block_sum_i64(ns, sh);
    if (threadIdx.x == 0) { ws[t] = ne; ws[T + t] = ns; }
}

// single block: exclusive scans over rows, cut row t*, totals.
// summary = [t_star, n_eps, n_steps, reached]
__global__ void __launch_bounds__(1024)
pack_plan_kernel(int64_t* __restrict__ ws, int T, int64_t num_samples,
                 int64_t* __restrict__ summary) {
    if (threadIdx.x != 0) return;  // T <= a few thousand: a serial scan is ~us
    int64_t e = 0, s = 0, t_star = -1;
    for (int t = 0; t < T; ++t) {
        ws[2 * T + t] = e;
        ws[3 * T + t] = s;
        if (t_star < 0) {
            e += ws[t];
            s += ws[T + t];
            if (s >= num_samples) t_star = t;
        }
    }
    summary[0] = t_star < 0 ? (int64_t)T - 1 : t_star;
    summary[1] = e;
    summary[2] = s;
    summary[3] = t_star >= 0 ? 1 : 0;
}

extern "C" int b2g_pack_plan(const uint8_t* step_type, const int32_t* ep_len, int T,
                             int64_t N, int64_t num_samples, void* ws, int64_t* summary4,
                             void* stream) {
    B2G_CHECK_ARG(T > 0 && N > 0, "empty rollout buffer");
    B2G_CHECK_ARG(step_type && ep_len && ws && summary4, "NULL pointer");
    cudaStream_t s = (cudaStream_t)stream;
    pack_row_stats_kernel<<<T, 256, 0, s>>>(step_type, ep_len, N, (int64_t*)ws, T);
    pack_plan_kernel<<<1, 32, 0, s>>>((int64_t*)ws, T, num_samples, summary4);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

// one block per row t <= t*: rank the finished envs of the row in env order
// (block-wide exclusive scan with a running prefix), emit per-episode metadata
// and mark the episode's packed offset at its last step.
__global__ void __launch_bounds__(256)
pack_index_kernel(const uint8_t* __restrict__ st, const int32_t* __restrict__ ep_len,
                  const float* __restrict__ last_obs, int O, int64_t N,
                  const int64_t* __restrict__ ws, int T, int64_t* __restrict__ lengths,
                  int64_t* __restrict__ ep_off, float* __restrict__ last_observations,
                  int32_t* __restrict__ ep_env, int32_t* __restrict__ ep_end) {
    const int t = blockIdx.x;
    if (ws[t] == 0) return;
    __shared__ int64_t sh_e[8], sh_s[8];
    __shared__ int64_t run_e, run_s;
    if (threadIdx.x == 0) { run_e = ws[2 * T + t]; run_s = ws[3 * T + t]; }
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int64_t base = 0; base < N; base += blockDim.x) {
        const int64_t n = base + threadIdx.x;
        const int64_t i = (int64_t)t * N + n;
        const bool done = n < N && st[i] >= ST_TERMINAL;
        const int64_t L = done ? ep_len[i] : 0;
        // warp inclusive scans of (done, L)
        int64_t pe = done ? 1 : 0, ps = L;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int64_t ye = __shfl_up_sync(0xffffffffu, pe, o);
            const int64_t ys = __shfl_up_sync(0xffffffffu, ps, o);
            if (lane >= o) { pe += ye; ps += ys; }
        }
        if (lane == 31) { sh_e[w] = pe; sh_s[w] = ps; }
        __syncthreads();
        int64_t we = 0, wsum = 0;
        for (int k = 0; k < w; ++k) { we += sh_e[k]; wsum += sh_s[k]; }
        if (done) {
            const int64_t e = run_e + we + pe - 1;
            const int64_t off = run_s + wsum + ps - L;
            lengths[e] = L;
            ep_off[i] = off;
            if (ep_env) ep_env[e] = (int32_t)n;
            if (ep_end) ep_end[e] = t;
            if (last_observations)
                for (int k = 0; k < O; ++k) last_observations[e * O + k] = last_obs[i * O + k];
        }
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) { run_e += we + pe; run_s += wsum + ps; }
        __syncthreads();
    }
}

extern "C" int b2g_pack_index(const uint8_t* step_type, const int32_t* ep_len,
                              const float* last_obs, int obs_dim, int T_used, int T_ws,
                              int64_t N, const void* ws, int64_t* lengths_out,
                              int64_t* ep_off, float* last_observations_out,
                              int32_t* ep_env_out, int32_t* ep_end_out, void* stream) {
    B2G_CHECK_ARG(T_used > 0 && T_used <= T_ws && N > 0, "bad shape");
    B2G_CHECK_ARG(step_type && ep_len && ws && lengths_out && ep_off, "NULL pointer");
    B2G_CHECK_ARG(!last_observations_out || last_obs, "last_observations_out without last_obs");
    pack_index_kernel<<<T_used, 256, 0, (cudaStream_t)stream>>>(
        step_type, ep_len, last_obs, obs_dim, N, (const int64_t*)ws, T_ws, lengths_out, ep_off,
        last_observations_out, ep_env_out, ep_end_out);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

// ---- field scatter: thread per env, walk rows t* .. 0 (coalesced reads across
// n); at a finished step pick up the episode's packed offset, then every step of
// that episode goes to off + (its index in the episode).  Steps after the last
// finished episode of a column (partial episodes) are dropped, as the reference
// drops them at the next update_agent (vec_worker.py:107-127).
constexpr int PACK_MAX_FIELDS = 12;
struct PackFields {
    int n;
    const void* src[PACK_MAX_FIELDS];
    void* dst[PACK_MAX_FIELDS];
    int bytes[PACK_MAX_FIELDS];  // bytes per step: multiple of 4 (float words) or 1 (uint8)
};

__global__ void __launch_bounds__(128)
pack_fields_kernel(const uint8_t* __restrict__ st, const int32_t* __restrict__ ep_len,
                   const int64_t* __restrict__ ep_off, int T_used, int64_t N,
                   const PackFields f) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    int64_t dst_row = -1;  // packed row of the current step, -1 = not in a finished episode
    for (int t = T_used - 1; t >= 0; --t) {
        const int64_t i = (int64_t)t * N + n;
        if (st[i] >= ST_TERMINAL) dst_row = ep_off[i] + ep_len[i] - 1;
        if (dst_row >= 0) {
#pragma unroll 1
            for (int k = 0; k < f.n; ++k) {
                const int b = f.bytes[k];
                if (b == 1) {
                    ((uint8_t*)f.dst[k])[dst_row] = ((const uint8_t*)f.src[k])[i];
                } else {
                    const int wds = b >> 2;
                    const uint32_t* s = (const uint32_t*)f.src[k] + i * wds;
                    uint32_t* d = (uint32_t*)f.dst[k] + dst_row * wds;
                    for (int q = 0; q < wds; ++q) d[q] = s[q];
                }
            }
            dst_row -= 1;
            // the step before a FIRST-of-episode boundary belongs to another
            // episode: it either carries its own mark or is partial
            if (t > 0 && st[i - N] >= ST_TERMINAL) dst_row = -1;
        }
    }
}

extern "C" int b2g_pack_fields(const uint8_t* step_type, const int32_t* ep_len,
                               const int64_t* ep_off, int T_used, int64_t N, int n_fields,
                               const void* const* src, void* const* dst,
                               const int* bytes_per_step, void* stream) {
    B2G_CHECK_ARG(T_used > 0 && N > 0, "bad shape");
    B2G_CHECK_ARG(n_fields >= 0 && n_fields <= PACK_MAX_FIELDS, "too many fields");
    B2G_CHECK_ARG(step_type && ep_len && ep_off, "NULL pointer");
    PackFields f;
    f.n = n_fields;
    for (int k = 0; k < n_fields; ++k) {
        B2G_CHECK_ARG(src[k] && dst[k], "NULL field pointer");
        B2G_CHECK_ARG(bytes_per_step[k] == 1 || (bytes_per_step[k] > 0 && bytes_per_step[k] % 4 == 0), "field width must be 1 or a multiple of 4 bytes");
        f.src[k] = src[k]; f.dst[k] = dst[k]; f.bytes[k] = bytes_per_step[k];
    }
    if (n_fields == 0) return B2G_OK;
    const unsigned blocks = (unsigned)((N + 127) / 128);
    pack_fields_kernel<<<blocks, 128, 0, (cudaStream_t)stream>>>(step_type, ep_len, ep_off, T_used, N, f);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

// ---- two-stage field packing: row map + tiled transpose ----------------------------------
// The scatter above gives every thread its own destination stream (4-byte stores, one episode
// apart) and only N threads.  The map stage keeps the serial walk (it needs the episode end that
// lies LATER in time) but only writes one int32 per step, coalesced across n; the copy stage is a
// shared-memory transpose: a 32 x 32 (t, n) tile is read along n and written along t, where
// consecutive steps of an episode are consecutive packed rows.

// row_map[t*N + n] = packed row of step (t, n), -1 for steps of unfinished (dropped) episodes
__global__ void __launch_bounds__(128)
pack_rowmap_kernel(const uint8_t* __restrict__ st, const int32_t* __restrict__ ep_len,
                   const int64_t* __restrict__ ep_off, int T_used, int64_t N, int32_t* __restrict__ row_map) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    constexpr int U = 8;
    int32_t dst_row = -1;
    for (int t1 = T_used - 1; t1 >= 0; t1 -= U) {
        uint8_t s[U + 1];  // s[u] = step_type at t1 - u; s[U] is the one below the chunk
#pragma unroll
        for (int u = 0; u <= U; ++u) s[u] = (t1 - u >= 0) ? st[(int64_t)(t1 - u) * N + n] : 0;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int t = t1 - u;
            if (t < 0) break;
            const int64_t i = (int64_t)t * N + n;
            if (s[u] >= ST_TERMINAL) dst_row = (int32_t)(ep_off[i] + ep_len[i] - 1);
            row_map[i] = dst_row;
            if (dst_row >= 0) {
                dst_row -= 1;
                if (t > 0 && s[u + 1] >= ST_TERMINAL) dst_row = -1;  // the step below ends another episode
            }
        }
    }
}

extern "C" int b2g_pack_rowmap(const uint8_t* step_type, const int32_t* ep_len, const int64_t* ep_off,
                               int T_used, int64_t N, int32_t* row_map, void* stream) {
    B2G_CHECK_ARG(T_used > 0 && N > 0 && (int64_t)T_used * N < (int64_t)1 << 31, "bad shape");
    B2G_CHECK_ARG(step_type && ep_len && ep_off && row_map, "NULL pointer");
    pack_rowmap_kernel<<<(unsigned)((N + 127) / 128), 128, 0, (cudaStream_t)stream>>>(step_type, ep_len, ep_off, T_used,
                                                                                    N, row_map);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

constexpr int PT = 32, PN = 32, PW = 4;  // tile: steps x envs; words of a field handled per pass
__global__ void __launch_bounds__(256)
pack_fields_mapped_kernel(const int32_t* __restrict__ row_map, int T_used, int64_t N, const PackFields f) {
    __shared__ uint32_t tile[PT][PN * PW + 1];
    __shared__ int32_t smap[PT][PN + 1];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int t0 = blockIdx.y * PT;
    const int64_t n0 = (int64_t)blockIdx.x * PN;
    const int nn = (int)((N - n0) < PN ? (N - n0) : PN);
    for (int tt = w; tt < PT; tt += 8) {
        const int t = t0 + tt;
        smap[tt][lane] = (t < T_used && lane < nn) ? row_map[(int64_t)t * N + n0 + lane] : -1;
    }
    for (int k = 0; k < f.n; ++k) {
        const int b = f.bytes[k];
        if (b == 1) {
            __syncthreads();
            uint8_t* tb = reinterpret_cast<uint8_t*>(&tile[0][0]);  // [PT][PN + 4] bytes
            for (int tt = w; tt < PT; tt += 8) {
                const int t = t0 + tt;
                if (t < T_used && lane < nn) tb[tt * (PN + 4) + lane] = ((const uint8_t*)f.src[k])[(int64_t)t * N + n0 + lane];
            }
            __syncthreads();
            for (int c = w; c < nn; c += 8) {
                const int32_t d = smap[lane][c];
                if (d >= 0) ((uint8_t*)f.dst[k])[d] = tb[lane * (PN + 4) + c];
            }
            continue;
        }
        const int wds = b >> 2;
        for (int q0 = 0; q0 < wds; q0 += PW) {
            const int wc = (wds - q0) < PW ? (wds - q0) : PW;
            __syncthreads();
            for (int tt = w; tt < PT; tt += 8) {
                const int t = t0 + tt;
                if (t >= T_used) continue;
                const uint32_t* s = (const uint32_t*)f.src[k] + ((int64_t)t * N + n0) * wds + q0;
                for (int e = lane; e < nn * wc; e += 32) {
                    const int nl = e / wc, q = e - nl * wc;
                    tile[tt][nl * PW + q] = s[(int64_t)nl * wds + q];
                }
            }
            __syncthreads();
            for (int c = w; c < nn; c += 8) {
                const int32_t d = smap[lane][c];
                if (d < 0) continue;
                uint32_t* o = (uint32_t*)f.dst[k] + (int64_t)d * wds + q0;
                for (int q = 0; q < wc; ++q) o[q] = tile[lane][c * PW + q];
            }
        }
    }
}

extern "C" int b2g_pack_fields_mapped(const int32_t* row_map, int T_used, int64_t N, int n_fields,
                                      const void* const* src, void* const* dst, const int* bytes_per_step,
                                      void* stream) {
    B2G_CHECK_ARG(T_used > 0 && N > 0 && row_map, "bad shape");
    B2G_CHECK_ARG(n_fields >= 0 && n_fields <= PACK_MAX_FIELDS, "too many fields");
    PackFields f;
    f.n = n_fields;
    for (int k = 0; k < n_fields; ++k) {
        B2G_CHECK_ARG(src[k] && dst[k], "NULL field pointer");
        B2G_CHECK_ARG(bytes_per_step[k] == 1 || (bytes_per_step[k] > 0 && bytes_per_step[k] % 4 == 0), "field width must be 1 or a multiple of 4 bytes");
        f.src[k] = src[k]; f.dst[k] = dst[k]; f.bytes[k] = bytes_per_step[k];
    }
    if (n_fields == 0) return B2G_OK;
    const dim3 grid((unsigned)((N + PN - 1) / PN), (unsigned)((T_used + PT - 1) / PT));
    B2G_CHECK_ARG(grid.y <= 65535, "too many time steps");
    pack_fields_mapped_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(row_map, T_used, N, f);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
