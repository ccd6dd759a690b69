/* libb2g -- C ABI of the B200-native on-policy hot path (garage v2021.03.0
 * drop-in: rollout -> EpisodeBatch -> GAE/returns -> PPO/VPG minibatch update).
 *
 * The reference is pure Python and has no FFI; the "binding" a maintainer adds
 * is a ctypes.CDLL stub.  Every entry point below names the
 * reference interface (file:line under /root/reference/src/garage) it replaces.
 *
 * Conventions
 *  - plain C, no C++/torch types.  All `*_d` / unqualified data pointers are
 *    DEVICE pointers owned by the caller (the library never allocates, frees or
 *    retains them).  Scratch memory is passed in explicitly (`ws`), its size
 *    comes from the matching `*_ws_bytes` query.
 *  - `stream` is a cudaStream_t passed as void*; every call is asynchronous on
 *    it.  Return 0 on success, >0 = cudaError_t, <0 = library code;
 *    `b2g_last_error()` gives the message (thread-local).
 *  - time-major buffers are `[T, N, ...]` contiguous (t outermost, env n next);
 *    episode-major buffers are `[N, P, ...]` / packed `[sum(lengths), ...]`
 *    exactly as garage's EpisodeBatch stores them.
 *  - not thread-safe per stream/workspace; one caller thread per rank (the
 *    reference path is single-threaded: sampler/local_sampler.py:14-18).
 */
#ifndef B2G_H_
#define B2G_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2G_VERSION 100 /* 0.1.0 */

#define B2G_OK 0
#define B2G_ERR_INVALID (-1)     /* bad argument (shape, null pointer, enum) */
#define B2G_ERR_UNSUPPORTED (-2) /* valid in garage, not built into the CUDA path */

/* garage StepType codes, _dtypes.py:15-42 */
#define B2G_STEP_FIRST 0
#define B2G_STEP_MID 1
#define B2G_STEP_TERMINAL 2
#define B2G_STEP_TIMEOUT 3

int b2g_version(void);
const char* b2g_last_error(void);
/* number of SMs of the current device (grid sizing is a multiple of it) */
int b2g_device_sm_count(void);

/* ------------------------------------------------------------------ GAE --
 * b2g_gae_time_major: fused GAE advantages + discounted returns over
 * time-major [T,N] buffers with done masks (reverse segmented scan).
 *   replaces  torch/_functions.py:25-85  compute_advantages   (conv2d, O(N*P^2))
 *        and  np/_functions.py:111-128   discount_cumsum      (scipy lfilter per row)
 *        as called from torch/algos/vpg.py:149-153, 349-379.
 *   delta_t = r_t + gamma * V_{t+1} - V_t ;  A_t = delta_t + gamma*lambda*A_{t+1}
 *   R_t = r_t + gamma * R_{t+1}
 *   At a step whose step_type is TERMINAL or TIMEOUT the carry is cut and
 *   V_{t+1} := 0 (the reference pads with zeros and never bootstraps, Q4).
 *   After the last row (t = T-1) V_T := last_val[n] (NULL -> 0) and carries 0.
 *   step_type may be NULL (no episode boundary inside the buffer).
 *   adv_out / ret_out may be NULL (GAE-only = 13 B/elem, returns-only = 8 B/elem,
 *   fused = 17 B/elem algorithmic traffic).
 */
int b2g_gae_time_major(const float* rew, const float* val, const uint8_t* step_type,
                       const float* last_val, float gamma, float lam,
                       float* adv_out, float* ret_out, int T, int64_t N, void* stream);

/* b2g_gae_episode_major: same recurrences over garage's own layout,
 * baselines/rewards [N,P] row-major, zero padded (the exact signature of
 * torch/_functions.py:25 compute_advantages; V_P := 0).  One warp per episode,
 * warp-shuffle affine scan over 32-step segments.  ret_out may be NULL;
 * rewards64 (double, may be NULL) when given is the source of the returns,
 * matching vpg.py:149-153 which runs discount_cumsum on float64 rewards.
 */
int b2g_gae_episode_major(const float* baselines, const float* rewards, float gamma,
                          float lam, float* adv_out, float* ret_out, int64_t N, int P,
                          void* stream);

/* b2g_discount_cumsum: np/_functions.py:111-128 over rows of x[N,P] (float64 in,
 * float64 out, like scipy.signal.lfilter on garage's float64 rewards). */
int b2g_discount_cumsum_f64(const double* x, double gamma, double* out, int64_t N, int P,
                            void* stream);

/* b2g_moments: count, sum, sum of squared deviations are not enough for fp32
 * parity, so this returns mean and UNBIASED variance (torch.var default) and
 * min of x[0..n) as doubles in out[3], computed with fp64 accumulation in two
 * passes (vpg.py:371-377: center_adv divides by variance + 1e-8, Q3).
 * ws: b2g_moments_ws_bytes(). */
int64_t b2g_moments_ws_bytes(void);
int b2g_moments(const float* x, int64_t n, double* out3, void* ws, void* stream);

/* y = (x - mean) / (var + 1e-8) [- min'] : vpg.py:371-377, in place allowed.
 * stats3 = device pointer written by b2g_moments.  positive_adv subtracts the
 * minimum of the centred values (== (min - mean)/(var+eps)). */
int b2g_normalize_adv(const float* x, const double* stats3, int center_adv,
                      int positive_adv, float* y, int64_t n, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2G_H_ */
