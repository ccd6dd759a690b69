// GridWorldEnv (reference envs/grid_world_env.py:44-250) batched on the device, and the
// VecWorker loop (sampler/vec_worker.py:107-219) over it with a ScriptedPolicy
// (np/policies/scripted_policy.py: action = table[observation]).  Everything is integer work:
// the parity bar is bit-exact against fixtures made by the reference's own VecWorker
// (tests/golden/gridworld.npz, the setting of tests/garage/sampler/test_vec_worker.py).
//
// cells[n_row * n_col]: B2G_GW_FREE ('F', 'S', '.'), B2G_GW_WALL ('W', 'x'), B2G_GW_HOLE ('H', 'o'),
// B2G_GW_GOAL ('G').  Actions 0 left, 1 down, 2 right, 3 up; moves are clipped at the border, a move
// into a wall (or out of a hole / goal cell) stays in place (grid_world_env.py:227-250).  Entering a
// hole ends the episode with reward 0, the goal with reward 1 (:150-163).
#include "common.cuh"

enum : uint8_t { GW_FREE = 0, GW_WALL = 1, GW_HOLE = 2, GW_GOAL = 3 };

struct GridStep {
    int next;
    float reward;
    bool done;
};

__device__ __forceinline__ GridStep grid_step(const uint8_t* __restrict__ cells, int n_row, int n_col,
                                              int state, int action) {
    const int x = state / n_col, y = state % n_col;
    const int dx = action == 1 ? 1 : (action == 3 ? -1 : 0);
    const int dy = action == 2 ? 1 : (action == 0 ? -1 : 0);
    const int nx = min(max(x + dx, 0), n_row - 1), ny = min(max(y + dy, 0), n_col - 1);
    int next = nx * n_col + ny;
    const uint8_t here = cells[state];
    if (cells[next] == GW_WALL || here == GW_HOLE || here == GW_GOAL) next = state;
    const uint8_t kind = cells[next];
    GridStep r;
    r.next = next;
    r.done = kind == GW_HOLE || kind == GW_GOAL;
    r.reward = kind == GW_GOAL ? 1.0f : 0.0f;
    return r;
}

// One Environment.step() for N envs (reset() == state = start, step_cnt = 0 is done by the host
// with a fill).  err: first env (+1) that was stepped before reset (step_cnt < 0) or got an action
// outside 0..3.
__global__ void __launch_bounds__(256)
gridworld_step_kernel(const uint8_t* __restrict__ cells, int n_row, int n_col, int max_len,
                      int32_t* __restrict__ state, int32_t* __restrict__ step_cnt,
                      const int32_t* __restrict__ action, int32_t* __restrict__ obs, float* __restrict__ rew,
                      uint8_t* __restrict__ step_type, int32_t* __restrict__ err, int64_t N) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const int cnt = step_cnt[n], a = action[n];
    if (cnt < 0 || a < 0 || a > 3) {
        atomicCAS(err, 0, (int)(n + 1));
        return;
    }
    const GridStep s = grid_step(cells, n_row, n_col, state[n], a);
    const uint8_t st = b2g_step_type(cnt + 1, max_len, s.done);
    state[n] = s.next;
    obs[n] = s.next;
    rew[n] = s.reward;
    step_type[n] = st;
    step_cnt[n] = st >= ST_TERMINAL ? -1 : cnt + 1;  // grid_world_env.py:171-172: reset() needed
}

// VecWorker.rollout for n_steps time steps: thread per env, time-major outputs (a warp writes 32
// consecutive envs of a row).  Row t holds the observation the action was chosen from, the action,
// the reward and the step type; rows that end an episode also hold the final observation and the
// episode length, and the env restarts from `start` (vec_worker.py:139-174).
__global__ void __launch_bounds__(256)
gridworld_rollout_kernel(const uint8_t* __restrict__ cells, int n_row, int n_col, int start, int max_len,
                         const int32_t* __restrict__ table, int32_t* __restrict__ state,
                         int32_t* __restrict__ step_cnt, int n_steps, int64_t N, int32_t* __restrict__ obs,
                         int32_t* __restrict__ act, float* __restrict__ rew, uint8_t* __restrict__ step_type,
                         int32_t* __restrict__ last_obs, int32_t* __restrict__ ep_len) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    int s = state[n], cnt = step_cnt[n];
    if (cnt < 0) { s = start; cnt = 0; }
    for (int t = 0; t < n_steps; ++t) {
        const int64_t i = (int64_t)t * N + n;
        const int a = table[s] & 3;
        const GridStep g = grid_step(cells, n_row, n_col, s, a);
        const uint8_t st = b2g_step_type(cnt + 1, max_len, g.done);
        obs[i] = s;
        act[i] = a;
        rew[i] = g.reward;
        step_type[i] = st;
        if (st >= ST_TERMINAL) {
            last_obs[i] = g.next;
            ep_len[i] = cnt + 1;
            s = start;
            cnt = 0;
        } else {
            ep_len[i] = 0;
            s = g.next;
            ++cnt;
        }
    }
    state[n] = s;
    step_cnt[n] = cnt;
}

static bool grid_args_ok(const uint8_t* cells, int n_row, int n_col) {
    return cells != nullptr && n_row > 0 && n_col > 0 && (int64_t)n_row * n_col <= (1 << 20);
}

extern "C" int b2g_gridworld_step(const uint8_t* cells, int n_row, int n_col, int max_episode_length,
                                  int32_t* state, int32_t* step_cnt, const int32_t* action, int32_t* obs,
                                  float* reward, uint8_t* step_type, int32_t* err, int64_t N, void* stream) {
    B2G_CHECK_ARG(grid_args_ok(cells, n_row, n_col), "bad grid");
    B2G_CHECK_ARG(N >= 0, "negative N");
    if (N == 0) return B2G_OK;
    B2G_CHECK_ARG(state && step_cnt && action && obs && reward && step_type && err, "NULL pointer");
    gridworld_step_kernel<<<(unsigned)((N + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        cells, n_row, n_col, max_episode_length, state, step_cnt, action, obs, reward, step_type, err, N);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}

extern "C" int b2g_gridworld_rollout(const uint8_t* cells, int n_row, int n_col, int start_state,
                                     int max_episode_length, const int32_t* scripted_actions,
                                     int32_t* state, int32_t* step_cnt, int n_steps, int64_t N, int32_t* obs,
                                     int32_t* act, float* rew, uint8_t* step_type, int32_t* last_obs,
                                     int32_t* ep_len, void* stream) {
    B2G_CHECK_ARG(grid_args_ok(cells, n_row, n_col), "bad grid");
    B2G_CHECK_ARG(start_state >= 0 && start_state < n_row * n_col, "start state outside the grid");
    B2G_CHECK_ARG(n_steps >= 0 && N >= 0, "negative shape");
    if (n_steps == 0 || N == 0) return B2G_OK;
    B2G_CHECK_ARG(scripted_actions && state && step_cnt && obs && act && rew && step_type && last_obs && ep_len,
                  "NULL pointer");
    gridworld_rollout_kernel<<<(unsigned)((N + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        cells, n_row, n_col, start_state, max_episode_length, scripted_actions, state, step_cnt, n_steps, N, obs,
        act, rew, step_type, last_obs, ep_len);
    B2G_CHECK_LAUNCH();
    return B2G_OK;
}
