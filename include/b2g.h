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
/* *counter += delta on the stream (device-side step/epoch counters of graph replays) */
int b2g_counter_add(unsigned long long* counter, unsigned long long delta, void* stream);
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
 * warp-shuffle affine scan over 32-step segments.  ret_out may be NULL; the
 * returns scan runs in float64 and is cast at the end, as vpg.py:149-153 does
 * (discount_cumsum on float64, then torch.Tensor()).
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

/* b2g_moments_static: b2g_moments without a workspace argument and without any allocation.  It has
 * no counterpart in the reference; it was added for callers
 * that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  The scratch (594 doubles, 4,752 bytes) is a file-scope __device__ array
 * that every GPU loading the library holds until exit; its device address is resolved once per
 * device with cudaGetSymbolAddress.  Same kernels, same results as b2g_moments (bit-identical).
 * Limit: n <= b2g_moments_static_max_n() = 2^28 elements; larger n returns B2G_ERR_INVALID with a
 * message.  All calls on one device share the array: calls that overlap on the GPU (two streams)
 * race; serialise them or use b2g_moments with separate workspaces. */
int64_t b2g_moments_static_max_n(void);
int b2g_moments_static(const float* x, int64_t n, double* out3, void* stream);

/* y = (x - mean) / (var + 1e-8) [- min'] : vpg.py:371-377, in place allowed.
 * stats3 = device pointer written by b2g_moments.  positive_adv subtracts the
 * minimum of the centred values (== (min - mean)/(var+eps)). */
int b2g_normalize_adv(const float* x, const double* stats3, int center_adv,
                      int positive_adv, float* y, int64_t n, void* stream);

/* b2g_episode_stats: log_performance (_functions.py:233-275) over an episode-major packed
 * batch.  rewards[B] float32 (rewards_f64 = 0) or float64 (1, the reference's dtype);
 * step_type[B] StepType codes (may be NULL: termination column = 0); success[B] u8 or NULL.
 * Episode e spans [e*fixed_len, (e+1)*fixed_len) when fixed_len > 0, else
 * [offsets[e], offsets[e] + lengths[e]).  per_episode4 = double[4][n_eps]: undiscounted return
 * (_functions.py:260), discounted return sum_t discount^t r_t (== discount_cumsum(r)[0],
 * _functions.py:251, float64 accumulation), any step TERMINAL (0/1, :254-256), any success
 * (0/1, :257-258).  out7 (may be NULL) = the values log_performance records: mean discounted
 * return, mean / population std / max / min undiscounted return, termination rate, success
 * rate (NaN when success is NULL). */
int b2g_episode_stats(const void* rewards, int rewards_f64, const uint8_t* step_type,
                      const uint8_t* success, const int64_t* lengths, const int64_t* offsets,
                      int64_t fixed_len, double discount, int64_t n_eps, double* per_episode4,
                      double* out7, void* stream);

/* ------------------------------------------------------------- GridWorld --
 * GridWorldEnv (envs/grid_world_env.py:44-250) for N env copies of one map, and the VecWorker
 * loop (sampler/vec_worker.py:107-219) over it with a ScriptedPolicy
 * (np/policies/scripted_policy.py:6-60: action = scripted_actions[observation]).  Integer work:
 * bit-exact against the reference (tests/golden/gridworld.npz).
 * cells[n_row*n_col] u8 (device): 0 free/start, 1 wall, 2 hole, 3 goal.  state / step_cnt: int32[N]
 * device arrays owned by the caller; step_cnt < 0 means "reset() needed" (the rollout resets such
 * envs to start_state itself; b2g_gridworld_step reports them through err like the reference's
 * RuntimeError, grid_world_env.py:139-140).
 * b2g_gridworld_step == Environment.step for every env (actions int32[N] in 0..3): writes obs
 * (next state), reward (1 at the goal), step_type (StepType codes; timeout wins over done).
 * b2g_gridworld_rollout: n_steps rows of time-major [n_steps, N] buffers (obs = the state the
 * action was chosen from, act, rew, step_type; last_obs / ep_len at rows that end an episode);
 * the buffers go through b2g_pack_* unchanged (last_obs is bit-copied, O = 1). */
#define B2G_GW_FREE 0
#define B2G_GW_WALL 1
#define B2G_GW_HOLE 2
#define B2G_GW_GOAL 3
int b2g_gridworld_step(const uint8_t* cells, int n_row, int n_col, int max_episode_length,
                       int32_t* state, int32_t* step_cnt, const int32_t* action, int32_t* obs,
                       float* reward, uint8_t* step_type, int32_t* err, int64_t N, void* stream);
int b2g_gridworld_rollout(const uint8_t* cells, int n_row, int n_col, int start_state,
                          int max_episode_length, const int32_t* scripted_actions,
                          int32_t* state, int32_t* step_cnt, int n_steps, int64_t N, int32_t* obs,
                          int32_t* act, float* rew, uint8_t* step_type, int32_t* last_obs,
                          int32_t* ep_len, void* stream);

/* ------------------------------------------------------------------ envs --
 * GPU-resident batched environments.  env_id selects the dynamics; `params`
 * is a HOST array of B2G_ENV_NPARAMS doubles:
 *   B2G_ENV_POINT    (envs/point_env.py:25-162; obs 3, act 2, state 2)
 *       [goal_x, goal_y, arena_size, done_bonus, never_done]
 *   B2G_ENV_PENDULUM (gym==0.17.2 pendulum.py behind envs/gym_env.py:182-270;
 *                     obs 3, act 1, state 2 = theta, theta_dot)
 *       [max_speed 8, max_torque 2, dt .05, g 10, m 1, l 1]
 *   B2G_ENV_CARTPOLE (gym==0.17.2 cartpole.py, "CartPole-style" continuous
 *                     action; obs 4, act 1, state 4)
 *       [gravity 9.8, masscart 1, masspole .1, length .5, force_mag 10, tau .02,
 *        x_threshold 2.4, theta_threshold 12 deg in rad,
 *        act_mode (0: force = +F if a > 0 else -F; 1: force = F*clip(a,-1,1))]
 * state is float64 SoA [S, N] (gym integrates in float64; PointEnv's float32
 * values are stored exactly), step_cnt int32 [N] (-1 = reset() needed, the
 * reference's `_step_cnt = None`), episode int32 [N] counts resets per env and
 * indexes the Philox reset stream so that results do not depend on sharding:
 * global env id = env_id_offset + n.
 */
#define B2G_ENV_POINT 0
#define B2G_ENV_PENDULUM 1
#define B2G_ENV_CARTPOLE 2
#define B2G_ENV_NPARAMS 10

int b2g_env_obs_dim(int env_id);
int b2g_env_act_dim(int env_id);
int b2g_env_state_dim(int env_id);

/* Environment.reset() for all N envs: point_env.py:79-98 / gym_env.py:182-199.
 * Writes state, obs[N,O] (float32), step_cnt := 0, episode[n] += 1. */
int b2g_env_reset(int env_id, const double* params, uint64_t seed, int64_t env_id_offset,
                  double* state, float* obs, int32_t* step_cnt, int32_t* episode, int64_t N,
                  void* stream);

/* Environment.step(action) for all N envs, ONE step, no auto-reset (the
 * replay-parity entry point): point_env.py:100-162 / gym_env.py:201-270.
 * actions[N,A] float32 (stored un-clipped by the caller, point_env.py:155);
 * obs_out[N,O] float32; reward_out[N] float64 (python float in the reference);
 * step_type_out[N] uint8 via StepType.get_step_type (_dtypes.py:44-68);
 * success_out[N] uint8 or NULL (PointEnv env_info['success']).
 * An env whose step_cnt is -1 ("reset() must be called before step()",
 * point_env.py:113-114) is left untouched and *err_flag (device int32, may be
 * NULL) is set to 1 + its index; the host wrapper raises RuntimeError.  On a
 * TERMINAL/TIMEOUT step step_cnt becomes -1 (point_env.py:147-148). */
int b2g_env_step(int env_id, const double* params, int max_episode_length, double* state,
                 const float* actions, float* obs_out, double* reward_out,
                 uint8_t* step_type_out, uint8_t* success_out, int32_t* step_cnt,
                 int32_t* err_flag, int64_t N, void* stream);

/* --------------------------------------------------------------- rollout --
 * Policy parameter vector (float32, the order of GaussianMLPPolicy.state_dict(),
 * torch/policies/gaussian_mlp_policy.py:53-87 with hidden_sizes=(H,H)):
 *   [ log_std (1) | W1 [H,O] | b1 [H] | W2 [H,H] | b2 [H] | W3 [A,H] | b3 [A] ]
 * b2g_mlp_param_count gives its length.  H in {32, 64, 128}; other sizes return
 * B2G_ERR_UNSUPPORTED (never a silent fallback).
 */
int64_t b2g_mlp_param_count(int obs_dim, int hidden, int out_dim);

/* b2g_rollout: n_steps of  get_actions -> env.step -> bookkeeping  for N envs in
 * ONE persistent launch; replaces the Python loop of sampler/vec_worker.py:176-204
 * (+ _gather_episode :139-174, default_worker.py:100-122) and the policy
 * forward/sample of torch/policies/stochastic_policy.py:46-89.
 *   in/out (persist across calls): state[S,N] f64, cur_obs[N,O], step_cnt[N],
 *     episode[N] (as b2g_env_reset leaves them).
 *   min/max_log_std: log(min_std) / log(max_std) clamps of
 *     gaussian_mlp_module.py:171-176 (+-INFINITY for None).
 *   Philox counter = (env_id_offset + n, t_counter + *t_counter_dev + t): results
 *     do not depend on how envs are sharded over ranks or on chunking over time.
 *     t_counter_dev (device u64, may be NULL) lets a captured CUDA graph advance
 *     the stream between replays (b2g_counter_add).
 *   out, time-major rows t = 0..n_steps-1: obs[T,N,O] (observation the action
 *     was computed from), act[T,N,A] (un-clipped sample), rew[T,N] f32,
 *     logp[T,N] (log-prob of act under the rollout policy == PPO's old_ll, Q8),
 *     step_type[T,N] u8, mean[T,N,A] (agent_info, may be NULL),
 *     success[T,N] u8 (PointEnv env_info, may be NULL), last_obs[T,N,O]
 *     (may be NULL; written only on TERMINAL/TIMEOUT rows: the post-step
 *     observation that EpisodeBatch.last_observations holds, Q10),
 *     ep_len[T,N] i32 (may be NULL; written on TERMINAL/TIMEOUT rows only:
 *     the length of the episode that ends there, input of b2g_pack_*).
 *   An env that finishes is reset in the same step (vec_worker.py:171).
 */
int b2g_rollout(int env_id, const double* params, int max_episode_length,
                const float* policy_params, int hidden, float min_log_std, float max_log_std,
                uint64_t seed, int64_t env_id_offset, uint32_t t_counter,
                const unsigned long long* t_counter_dev, double* state,
                float* cur_obs, int32_t* step_cnt, int32_t* episode, int n_steps, int64_t N,
                float* obs, float* act, float* rew, float* logp, uint8_t* step_type,
                float* mean, uint8_t* success, float* last_obs, int32_t* ep_len, void* stream);

/* ------------------------------------------------------ normalize(env) --
 * envs/normalized_env.py:8-171 (`normalize = NormalizedEnv`), the wrapper garage's
 * benchmark launchers put around every env
 * (benchmarks/.../algos/ppo_garage_pytorch.py:36), fused into the step:
 *   B2G_NORM_ACTION  :91-101  scaled = clip(lb + (a + s) * (0.5 (ub - lb) / s), lb, ub) in
 *                    float32 like numpy on the float32 action; the host sets the flag from the
 *                    reference's bound test (:94, including its `ub != -inf` spelling).  The
 *                    stored action stays the UNSCALED one (:111).
 *   B2G_NORM_OBS     :117-146  per env copy: mean = (1-a) mean + a o; var = (1-a) var + a (o - mean)^2
 *                    (the NEW mean); o' = (o - mean) / (sqrt(var) + 1e-8); float64.  Applied to the
 *                    observation of reset() (:71-73) and of every step (:106-107).
 *   B2G_NORM_REWARD  :124-131, 148-160  same recursion on the reward; r' = r / (sqrt(var) + 1e-8).
 *   scale_reward     :112  multiplies the (normalised) reward.
 * stats: device float64 [2*O + 2, N] SoA = obs_mean[O], obs_var[O], reward_mean, reward_var of
 * every env copy (initially 0 / 1 / 0 / 1, :50-55); it persists across calls like the wrapper
 * object does.  A NULL `norm` (or flags 0, scale_reward 1) is the bare environment. */
#define B2G_NORM_ACTION 1
#define B2G_NORM_OBS 2
#define B2G_NORM_REWARD 4
typedef struct b2g_norm {
    int flags;
    float act_low[2], act_high[2]; /* action_space.low / .high (float32) */
    float expected_action_scale;
    double scale_reward, obs_alpha, reward_alpha;
    double* stats; /* device [2*O + 2, N] */
} b2g_norm;

/* b2g_env_reset / b2g_env_step / b2g_rollout with the wrapper applied (same arguments + norm). */
int b2g_env_reset_norm(int env_id, const double* params, const b2g_norm* norm, uint64_t seed,
                       int64_t env_id_offset, double* state, float* obs, int32_t* step_cnt,
                       int32_t* episode, int64_t N, void* stream);
int b2g_env_step_norm(int env_id, const double* params, const b2g_norm* norm, int max_episode_length,
                      double* state, const float* actions, float* obs_out, double* reward_out,
                      uint8_t* step_type_out, uint8_t* success_out, int32_t* step_cnt,
                      int32_t* err_flag, int64_t N, void* stream);
int b2g_rollout_norm(int env_id, const double* params, const b2g_norm* norm, int max_episode_length,
                     const float* policy_params, int hidden, float min_log_std, float max_log_std,
                     uint64_t seed, int64_t env_id_offset, uint32_t t_counter,
                     const unsigned long long* t_counter_dev, double* state,
                     float* cur_obs, int32_t* step_cnt, int32_t* episode, int n_steps, int64_t N,
                     float* obs, float* act, float* rew, float* logp, uint8_t* step_type,
                     float* mean, uint8_t* success, float* last_obs, int32_t* ep_len, void* stream);

/* ------------------------------------------------------- EpisodeBatch pack --
 * Time-major rollout buffers -> garage's episode-major packed EpisodeBatch
 * layout; replaces sampler/vec_worker.py:139-174,206-219, the ">= num_samples,
 * whole episodes" loop of sampler/local_sampler.py:156-166 and
 * EpisodeBatch.concatenate (_dtypes.py:591-632).  Episodes are ordered by
 * completion row, ties by env index.
 *  b2g_pack_plan:  per-row episode/step counts, exclusive scans, the cut row.
 *     summary4 (device int64[4]) = [t_star, n_eps, n_steps, reached]; reached=0
 *     means the buffer holds fewer than num_samples completed steps (run more
 *     rows and call again).  ws: b2g_pack_ws_bytes(T).
 *  b2g_pack_index: lengths_out[n_eps] (int64, numpy 'l' like vec_worker.py:162),
 *     ep_off[T,N] (int64, sparse: packed offset of the episode that ends at
 *     (t,n)), last_observations_out[n_eps,O], ep_env_out/ep_end_out[n_eps]
 *     (env index / last row of each episode; may be NULL).  T_used = t_star+1,
 *     T_ws = the T given to b2g_pack_plan.
 *  b2g_pack_fields: scatter up to 12 per-step fields (src [T,N,w] -> dst
 *     [n_steps,w]); bytes_per_step[k] is 1 (uint8) or a multiple of 4.  src/dst/
 *     bytes_per_step are HOST arrays of n_fields entries.
 *  b2g_pack_rowmap + b2g_pack_fields_mapped: the same result in two stages, for batches
 *     that pack several field groups from one rollout: row_map[T_used,N] (int32) = packed row
 *     of every step (-1: step of an unfinished episode, dropped), then a tiled transpose
 *     through shared memory (reads coalesced along n, writes coalesced along t).
 */
int64_t b2g_pack_ws_bytes(int T);
int b2g_pack_plan(const uint8_t* step_type, const int32_t* ep_len, int T, int64_t N,
                  int64_t num_samples, void* ws, int64_t* summary4, void* stream);
int b2g_pack_index(const uint8_t* step_type, const int32_t* ep_len, const float* last_obs,
                   int obs_dim, int T_used, int T_ws, int64_t N, const void* ws,
                   int64_t* lengths_out, int64_t* ep_off, float* last_observations_out,
                   int32_t* ep_env_out, int32_t* ep_end_out, void* stream);
int b2g_pack_fields(const uint8_t* step_type, const int32_t* ep_len, const int64_t* ep_off,
                    int T_used, int64_t N, int n_fields, const void* const* src,
                    void* const* dst, const int* bytes_per_step, void* stream);
int b2g_pack_rowmap(const uint8_t* step_type, const int32_t* ep_len, const int64_t* ep_off,
                    int T_used, int64_t N, int32_t* row_map, void* stream);
int b2g_pack_fields_mapped(const int32_t* row_map, int T_used, int64_t N, int n_fields,
                           const void* const* src, void* const* dst, const int* bytes_per_step,
                           void* stream);

/* StochasticPolicy.get_actions sampling (torch/policies/stochastic_policy.py:46-89):
 * act = mean + exp(log_std) * eps, eps = Philox(seed; env_id_offset + i, t_counter),
 * logp_out[n] (may be NULL) = Independent(Normal).log_prob(act).  log_std is the
 * already clamped scalar. */
int b2g_policy_sample(const float* mean, int act_dim, float log_std, uint64_t seed,
                      int64_t env_id_offset, uint32_t t_counter, float* act_out,
                      float* logp_out, int64_t n, void* stream);

/* ------------------------------------------------- minibatch MLP update --
 * Fused gather -> MLP forward -> loss -> backward for ONE minibatch of garage's
 * GaussianMLPPolicy / GaussianMLPValueFunction (parameter vector as above).
 *   replaces torch/algos/vpg.py:250-293 (_train_policy/_train_value_function:
 *   zero_grad, loss, backward), torch/algos/ppo.py:96-132 (clipped surrogate),
 *   vpg.py:324-347,408-455 (loss/entropy/VPG objective),
 *   value_functions/gaussian_mlp_value_function.py:81-97 (Gaussian NLL) and the
 *   autograd graph of torch/modules/multi_headed_mlp_module.py:136-151.
 *   rows: idx[mb] (int32 row ids into the flat [B,...] arrays, the reference's
 *   BatchDataset ids, np/optimizers/minibatch_dataset.py:21-35) or, with
 *   idx == NULL, rows start..start+mb.
 *   loss_kind: 0 PPO (needs act, adv, old_logp), 1 VPG (act, adv), 2 value NLL
 *   (ret; out_dim 1), [3 none: forward only, b2g_mlp_eval].
 *   flags: 1 entropy regularised, 2 softplus entropy, 4 stop entropy gradient.
 *   Result: ws_partial[n_partials][P + 8] per-CTA sums of UNSCALED gradients
 *   (d sum_rows loss_row) and statistics [sum loss_row, sum KL, sum entropy,
 *   #rows outside the clip range, #rows, 0,0,0]; n_partials =
 *   b2g_mlp_num_partials(obs_dim, hidden, out_dim, 1, mb); ws_partial holds at
 *   least b2g_mlp_ws_floats() floats.  step_counter (device u64, may be NULL)
 *   is incremented once per call: it is the Adam step of b2g_mlp_reduce /
 *   b2g_adam_step launched after it on the same stream (graph-capturable).
 */
#define B2G_LOSS_PPO 0
#define B2G_LOSS_VPG 1
#define B2G_LOSS_VF 2
#define B2G_LOSS_NONE 3
#define B2G_FLAG_ENT_REG 1
#define B2G_FLAG_SOFTPLUS_ENT 2
#define B2G_FLAG_STOP_ENT_GRAD 4
#define B2G_MLP_NSTATS 8

int b2g_mlp_num_partials(int obs_dim, int hidden, int out_dim, int backward, int64_t rows);
/* Arithmetic path of b2g_mlp_minibatch_grad: 0 = auto (hidden 64: tcgen05 kind::tf32 with a
 * 3xTF32 hi/lo split, fp32 accumulation in tensor memory; otherwise CUDA-core fp32),
 * 1 = CUDA-core fp32 for every shape, 2 = same as 0.  Also selectable with the environment
 * variable B2G_MLP_PATH=ffma.  Both paths meet the 1e-5 parity bar of the same reference code. */
int b2g_set_mlp_path(int path);
int b2g_get_mlp_path(void);
/* ---- fused training loop ------------------------------------------------------------------
 * ALL minibatch steps of VPG/PPO._train (torch/algos/vpg.py:230-293: for epoch, for minibatch:
 * optimizer.zero_grad(); loss; backward(); optimizer.step(), first for the policy, then for the
 * value function) in one cooperative launch.  One b2g_train_net per network; with two networks
 * the policy and value-function loops run concurrently on disjoint CTAs (they are independent
 * in the reference as well).  Parameters, Adam moments and the step counter are updated in
 * place.  shuffle: 0 = rows in order (minibatch = NULL/full batch), 1 = the permutation
 * b2g_permutation(shuffle_seed, counter = step counter at the start of the epoch) would write,
 * evaluated on the fly, 2 = `idx` holds [epochs][n_rows] int32 row ids (the reference's
 * BatchDataset order).  Shapes: see b2g_mlp_train_loop_kind below (tcgen05 loop for hidden 64 and the built
 * (obs, out) pairs, general CUDA-core loop for hidden 32 / 64 / 128, obs <= 8, out <= 4; anything else returns
 * B2G_ERR_UNSUPPORTED); workspace holds at least b2g_mlp_train_loop_ws_floats() floats. */
typedef struct b2g_train_net {
    int obs_dim, hidden, out_dim;
    float* params;                       /* flat parameter vector, updated in place */
    float* adam_m;
    float* adam_v;
    unsigned long long* step_counter;    /* device u64: Adam step, advanced by epochs*ceil(n_rows/minibatch) */
    double lr, beta1, beta2, eps;
    const float* obs;                    /* [n_rows, obs_dim] */
    const float* act;                    /* [n_rows, out_dim]  policy losses */
    const float* adv;                    /* [n_rows] */
    const float* old_logp;               /* [n_rows]           PPO */
    const float* ret;                    /* [n_rows]           value loss */
    int loss_kind, flags;                /* as b2g_mlp_minibatch_grad */
    float clip, ent_coeff, min_log_std, max_log_std;
    int64_t n_rows, minibatch;           /* minibatch <= 0: full batch */
    int epochs, shuffle;
    uint64_t shuffle_seed;
    const int32_t* idx;
} b2g_train_net;
int64_t b2g_mlp_train_loop_ws_floats(void);
/* Which persistent loop b2g_mlp_train_loop[_dist] runs for these networks (only obs_dim / hidden / out_dim
 * are read): 1 = tcgen05 tiles (hidden 64, (obs, out) in {(3,1), (3,2), (4,1)}, default MLP path; one or
 * several GPUs), 2 = general CUDA-core tiles (one hidden size in {32, 64, 128} for all networks, obs 1..8,
 * out 1..4; one GPU; the CTAs per network follow the minibatch, a 64-row minibatch runs on one CTA without
 * device-wide barriers), 0 = none: the caller takes the step-wise path. */
int b2g_mlp_train_loop_kind(const b2g_train_net* nets, int n_nets, int world);

/* Multi-GPU form (one process per GPU, env-sharded data parallelism; the reference has no
 * counterpart).  Every rank runs the same loop on its own shard; per minibatch step the gradient
 * slices are all-reduced INSIDE the kernel: each CTA stores its slice sums into every rank's
 * peer-mapped buffer over NVLink, raises a flag and adds the world's contributions in rank order
 * (bit-identical on all ranks), then applies Adam with 1 / (minibatch * world).  All ranks must
 * run the same number of steps (equal n_rows, minibatch, epochs).  peer_base[r]: rank r's
 * exchange buffer of b2g_train_comm_bytes() bytes, zeroed once (b2g_comm_alloc) and mapped into
 * this process (b2g_comm_open; CUDA IPC).  The step sequence numbers the flags carry live in the
 * rank's own buffer and are advanced by the kernel, so a captured launch can be replayed. */
#define B2G_MAX_RANKS 8
typedef struct b2g_train_comm {
    int world, rank;
    void* peer_base[B2G_MAX_RANKS];
} b2g_train_comm;
int64_t b2g_train_comm_bytes(void);
int b2g_mlp_train_loop_dist(const b2g_train_net* nets, int n_nets, const b2g_train_comm* comm,
                            float* workspace, int64_t workspace_floats, void* stream);
/* Small all-gather + combine of per-rank float64 scalars through the same peer-mapped buffers (one
 * block per rank, no NCCL, graph-capturable; every rank calls it in the same order and obtains
 * identical bits).  Replaces, for several ranks, the host-side arithmetic of
 *   op B2G_PEER_MOMENTS  vpg.py:371-377: in = b2g_moments' [mean, unbiased var, min] of this
 *        rank's advantages, scalar0 = their count -> out[3] = the same three over all ranks;
 *   op B2G_PEER_LOGS     _functions.py:262-273 + vpg.py:186-199: in = the 40 logging statistics of
 *        this rank (sums [0:32], return statistics [32:39], padded count [39]), scalar0 = samples,
 *        scalar1 = episodes -> out[0:40]: sums, episode-weighted means, the GLOBAL std, max / min;
 *        out[48] = samples, out[49] = episodes of all ranks, out[50] = 1 ("reduced"); out[40:48] is
 *        not written and out may alias in (>= 51 doubles);
 *   op B2G_PEER_GATHER   out[world][n] = every rank's in[n] (n <= 64). */
enum { B2G_PEER_GATHER = 0, B2G_PEER_MOMENTS = 1, B2G_PEER_LOGS = 2 };
int b2g_peer_reduce_f64(const b2g_train_comm* comm, int op, const double* in, int n, double scalar0,
                        double scalar1, double* out, void* stream);
int b2g_comm_alloc(int64_t bytes, void** ptr_out, void* handle_out_64);
int b2g_comm_open(const void* handle_64, void** ptr_out);
int b2g_comm_close(void* ptr);
int b2g_comm_free(void* ptr);
int b2g_mlp_train_loop(const b2g_train_net* nets, int n_nets, float* workspace,
                       int64_t workspace_floats, void* stream);
/* Debug only: device buffer [n_nets][CTAs per network][16] of int64 clock64() stamps of minibatch
 * step `step` of the next loops (NULL switches it off, the default). */
int b2g_mlp_train_loop_debug_timing(long long* buf, int step);

/* Debug only: device buffer [grid][32] of int64 that the tcgen05 kernel fills with clock64()
 * stamps of its stages (NULL switches it off, the default). */
int b2g_mlp_debug_timing(long long* buf);
int64_t b2g_mlp_ws_floats(int obs_dim, int hidden, int out_dim);

int b2g_mlp_minibatch_grad(int obs_dim, int hidden, int out_dim, const float* params,
                           const int32_t* idx, int64_t start, int64_t mb, const float* obs,
                           const float* act, const float* adv, const float* old_logp,
                           const float* ret, int loss_kind, int flags, float clip,
                           float ent_coeff, float min_log_std, float max_log_std,
                           float* ws_partial, unsigned long long* step_counter, void* stream);

/* Forward only over n rows: mean_out[n,out_dim] (may be NULL; policy means /
 * value baselines, vpg.py:155-156) and the same statistics as above, plus
 * sum KL(old || new) when old_mean[B,out_dim] and old_log_std (DEVICE scalar, the
 * old policy's clamped log-std) are given (vpg.py:381-406; the old policy is
 * the rollout policy, Q8).  n_partials =
 * b2g_mlp_num_partials(..., 0, n). */
int b2g_mlp_eval(int obs_dim, int hidden, int out_dim, const float* params, const int32_t* idx,
                 int64_t start, int64_t n, const float* obs, const float* act, const float* adv,
                 const float* old_logp, const float* ret, const float* old_mean,
                 const float* old_log_std, int loss_kind, int flags, float clip, float ent_coeff,
                 float min_log_std, float max_log_std, float* mean_out, float* ws_partial,
                 void* stream);

/* Reduce the per-CTA partials in fixed order: grad_out[n_params] (may be NULL) =
 * scale * sum; stats_out (device double[8], may be NULL) = sums.  With params
 * != NULL also applies torch.optim.Adam's update (make_optimizer,
 * _functions.py:25-65; amsgrad False, weight_decay 0) in place, step =
 * *step_counter. */
int b2g_mlp_reduce(const float* ws_partial, int n_partials, int n_params, float scale,
                   float* grad_out, double* stats_out, float* params, float* adam_m,
                   float* adam_v, const unsigned long long* step_counter, double lr,
                   double beta1, double beta2, double eps, void* stream);

/* Adam on a gradient vector that was all-reduced across ranks (grad * grad_scale). */
int b2g_adam_step(float* params, const float* grad, float grad_scale, float* adam_m,
                  float* adam_v, const unsigned long long* step_counter, double lr, double beta1,
                  double beta2, double eps, int n, void* stream);

/* ---- TRPO: torch/algos/trpo.py:9-144 on torch/optimizers/conjugate_gradient_optimizer.py ----
 * Fisher-vector product: what `_build_hessian_vector_product(f_constraint, params, reg_coeff)(vec)`
 * (conjugate_gradient_optimizer.py:18-66) returns for f_constraint = mean KL(old || new)
 * (torch/algos/vpg.py:381-406) at new == old, where the Hessian of the KL of two diagonal Gaussians with a
 * state-independent log-std is J^T diag(1/var) J / n on the mean parameters and 2 * out_dim on log_std
 * (0 when the clamp blocks its gradient):
 *   out[P] = that matrix times vec[P] + reg_coeff * vec.
 * One forward-mode pass writes u = (d mu / d theta . vec) / var into u_scratch [n, out_dim], one backward
 * pass with d loss / d mu := u accumulates J^T u.  Any obs_dim <= 8, out_dim <= 4, hidden in
 * {32, 64, 128}.  idx/start/n select the rows like b2g_mlp_minibatch_grad; ws_partial as there. */
int b2g_mlp_fvp(int obs_dim, int hidden, int out_dim, const float* params, const float* vec,
                const int32_t* idx, int64_t start, int64_t n, const float* obs, float min_log_std,
                float max_log_std, float reg_coeff, float* u_scratch, float* ws_partial, float* out,
                void* stream);
/* `_conjugate_gradient` (conjugate_gradient_optimizer.py:69-110) on float32 vectors of n elements, the
 * scalars in device memory `scal` (double[8]: [0] r.r, [1] converged flag, [2] step size, [3] x.Ax,
 * [4] iterations run): init sets x = 0, r = p = b; step consumes z = A p (b2g_mlp_fvp of p) and is a no-op
 * once r.r < residual_tol -- the reference's `break` without a host synchronisation. */
int b2g_cg_init(const float* b, float* x, float* r, float* p, double* scal, int n, void* stream);
int b2g_cg_step(const float* z, float* x, float* r, float* p, double* scal, int n, double residual_tol,
                void* stream);
/* `ConjugateGradientOptimizer.step` (:141-178): nan -> 0 in the step direction; then, with Ax = A x,
 * descent = sqrt(2 max_constraint_value / (x.Ax + 1e-8)) * x (step size 1 if that is nan). */
int b2g_cg_sanitize(float* x, int n, void* stream);
int b2g_cg_descent(const float* x, const float* Ax, double max_constraint_value, float* descent,
                   double* scal, int n, void* stream);
/* One point of `_backtracking_line_search` (:218-236): params = prev - ratio * descent. */
int b2g_line_search_point(const float* prev, const float* descent, float ratio, float* params, int n,
                          void* stream);

/* Random permutation of 0..n-1 as int32 ids (minibatch shuffle; stands in for
 * np.random.shuffle(self._ids), np/optimizers/minibatch_dataset.py:35, in
 * throughput mode -- parity mode passes the reference's own id lists to
 * b2g_mlp_minibatch_grad instead).  Keyed Feistel network + cycle walking, key =
 * Philox(seed; counter + *counter_dev): no sort, no state, graph-capturable
 * (counter_dev may be NULL). */
int b2g_permutation(uint64_t seed, const unsigned long long* counter_dev, uint64_t counter,
                    int64_t n, int32_t* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B2G_H_ */
