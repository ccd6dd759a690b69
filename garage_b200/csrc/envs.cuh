// Batched environment dynamics as device functions (state SoA in HBM / registers).
//   POINT     garage envs/point_env.py:79-162 (float32 numpy arithmetic)
//   PENDULUM  gym==0.17.2 gym/envs/classic_control/pendulum.py behind
//             garage envs/gym_env.py:182-270 (float64 state, like gym)
//   CARTPOLE  gym==0.17.2 gym/envs/classic_control/cartpole.py ("CartPole-style":
//             a continuous 1-d action is mapped to the two forces, see act_mode)
// gym is NOT in /root/reference: Pendulum/CartPole are restated from the
// published 0.17.2 source ("parity unpinned" at that third-party boundary).
#pragma once
#include "common.cuh"

struct EnvParams {
    int max_len;   // max_episode_length (<=0: none)
    double p[10];  // per-env constants, see b2g.h
};

template <int ENV> struct EnvDims;
template <> struct EnvDims<B2G_ENV_POINT> { static constexpr int O = 3, A = 2, S = 2; };
template <> struct EnvDims<B2G_ENV_PENDULUM> { static constexpr int O = 3, A = 1, S = 2; };
template <> struct EnvDims<B2G_ENV_CARTPOLE> { static constexpr int O = 4, A = 1, S = 4; };

inline int env_obs_dim_(int env) { return env == B2G_ENV_CARTPOLE ? 4 : 3; }
inline int env_act_dim_(int env) { return env == B2G_ENV_POINT ? 2 : 1; }
inline int env_state_dim_(int env) { return env == B2G_ENV_CARTPOLE ? 4 : 2; }

inline bool env_params_from_host(int env_id, const double* params, int max_len, EnvParams* out) {
    if (env_id < 0 || env_id > B2G_ENV_CARTPOLE || params == nullptr) return false;
    out->max_len = max_len;
    for (int i = 0; i < B2G_ENV_NPARAMS; ++i) out->p[i] = params[i];
    return true;
}

struct StepOut {
    double reward;
    bool done;     // env's own termination signal (before the time limit)
    bool success;  // PointEnv env_info['success']
};

// ------------------------------------------------------------------ reset --
// Initial state of episode `episode` of global env `gid` (Philox reset stream).
template <int ENV>
__device__ __forceinline__ void env_reset(const EnvParams& P, uint32_t seed_lo, uint32_t seed_hi,
                                          uint32_t gid, uint32_t episode, double* s, float* obs) {
    if (ENV == B2G_ENV_POINT) {
        // point_env.py:79-98: point = 0, obs = [0, 0, ||point - goal||]
        const float gx = (float)P.p[0], gy = (float)P.p[1];
        s[0] = 0.0; s[1] = 0.0;
        const float dx = 0.0f - gx, dy = 0.0f - gy;
        obs[0] = 0.0f; obs[1] = 0.0f; obs[2] = sqrtf(dx * dx + dy * dy);
    } else if (ENV == B2G_ENV_PENDULUM) {
        // pendulum.py reset: state = U(-[pi,1], [pi,1]); obs = [cos th, sin th, thdot]
        const Philox4 r = b2g_philox(seed_lo, seed_hi, gid, episode, B2G_STREAM_RESET, 0);
        const double pi = 3.14159265358979323846;
        s[0] = (2.0 * b2g_u01d(r.x, r.y) - 1.0) * pi;
        s[1] = (2.0 * b2g_u01d(r.z, r.w) - 1.0) * 1.0;
        double sn, cs;
        sincos(s[0], &sn, &cs);
        obs[0] = (float)cs; obs[1] = (float)sn; obs[2] = (float)s[1];
    } else {
        // cartpole.py reset: state = U(-0.05, 0.05, size=4)
        const Philox4 r0 = b2g_philox(seed_lo, seed_hi, gid, episode, B2G_STREAM_RESET, 0);
        const Philox4 r1 = b2g_philox(seed_lo, seed_hi, gid, episode, B2G_STREAM_RESET, 1);
        s[0] = (2.0 * b2g_u01d(r0.x, r0.y) - 1.0) * 0.05;
        s[1] = (2.0 * b2g_u01d(r0.z, r0.w) - 1.0) * 0.05;
        s[2] = (2.0 * b2g_u01d(r1.x, r1.y) - 1.0) * 0.05;
        s[3] = (2.0 * b2g_u01d(r1.z, r1.w) - 1.0) * 0.05;
#pragma unroll
        for (int i = 0; i < 4; ++i) obs[i] = (float)s[i];
    }
}

// ------------------------------------------------------------------- step --
template <int ENV>
__device__ __forceinline__ StepOut env_step(const EnvParams& P, double* s, const float* act,
                                            float* obs) {
    StepOut o;
    o.success = false;
    if (ENV == B2G_ENV_POINT) {
        // point_env.py:100-162, all float32 like numpy on float32 arrays.
        const float gx = (float)P.p[0], gy = (float)P.p[1], arena = (float)P.p[2];
        const float lo = -0.1f, hi = 0.1f;  // action_space, point_env.py:49-52
        const float ax = fminf(fmaxf(act[0], lo), hi), ay = fminf(fmaxf(act[1], lo), hi);
        float px = (float)s[0] + ax, py = (float)s[1] + ay;
        px = fminf(fmaxf(px, -arena), arena);
        py = fminf(fmaxf(py, -arena), arena);
        const float dx = px - gx, dy = py - gy;
        const float dist = sqrtf(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
        const float thr = sqrtf(__fadd_rn(__fmul_rn(lo, lo), __fmul_rn(lo, lo)));  // ||low||
        const bool succ = dist < thr;
        double reward = -(double)dist;
        if (succ) reward += P.p[3];  // done_bonus
        o.reward = reward;
        o.success = succ;
        o.done = succ && !(P.p[4] != 0.0);  // never_done
        s[0] = px; s[1] = py;
        obs[0] = px; obs[1] = py; obs[2] = dist;
    } else if (ENV == B2G_ENV_PENDULUM) {
        const double max_speed = P.p[0], max_torque = P.p[1], dt = P.p[2], g = P.p[3],
                     m = P.p[4], l = P.p[5];
        const double pi = 3.14159265358979323846;
        const double th = s[0], thdot = s[1];
        // u = np.clip(u, -max_torque, max_torque)[0] stays float32; u**2 is float32
        const float uf = fminf(fmaxf(act[0], (float)-max_torque), (float)max_torque);
        const double u = (double)uf;
        // angle_normalize: ((x + pi) % (2 pi)) - pi, python modulo (sign of divisor)
        double a = fmod(th + pi, 2.0 * pi);
        if (a < 0.0) a += 2.0 * pi;
        a -= pi;
        const double costs = a * a + 0.1 * (thdot * thdot) + 0.001 * (double)__fmul_rn(uf, uf);
        double newthdot = thdot + (-3.0 * g / (2.0 * l) * sin(th + pi) + 3.0 / (m * (l * l)) * u) * dt;
        const double newth = th + newthdot * dt;
        newthdot = fmin(fmax(newthdot, -max_speed), max_speed);
        s[0] = newth; s[1] = newthdot;
        double sn, cs;
        sincos(newth, &sn, &cs);
        obs[0] = (float)cs; obs[1] = (float)sn; obs[2] = (float)newthdot;
        o.reward = -costs;
        o.done = false;
    } else {
        const double gravity = P.p[0], masscart = P.p[1], masspole = P.p[2], length = P.p[3],
                     force_mag = P.p[4], tau = P.p[5], x_thr = P.p[6], th_thr = P.p[7];
        const int act_mode = (int)P.p[8];
        const double total_mass = masspole + masscart;
        const double polemass_length = masspole * length;
        double force;
        if (act_mode == 0) force = (act[0] > 0.0f) ? force_mag : -force_mag;  // a>0 == action 1
        else force = force_mag * (double)fminf(fmaxf(act[0], -1.0f), 1.0f);
        double x = s[0], x_dot = s[1], theta = s[2], theta_dot = s[3];
        double sintheta, costheta;
        sincos(theta, &sintheta, &costheta);
        const double temp = (force + polemass_length * (theta_dot * theta_dot) * sintheta) / total_mass;
        const double thetaacc = (gravity * sintheta - costheta * temp) /
                                (length * (4.0 / 3.0 - masspole * (costheta * costheta) / total_mass));
        const double xacc = temp - polemass_length * thetaacc * costheta / total_mass;
        x = x + tau * x_dot;            // kinematics_integrator == 'euler'
        x_dot = x_dot + tau * xacc;
        theta = theta + tau * theta_dot;
        theta_dot = theta_dot + tau * thetaacc;
        s[0] = x; s[1] = x_dot; s[2] = theta; s[3] = theta_dot;
        o.done = (x < -x_thr) || (x > x_thr) || (theta < -th_thr) || (theta > th_thr);
        o.reward = 1.0;
        obs[0] = (float)x; obs[1] = (float)x_dot; obs[2] = (float)theta; obs[3] = (float)theta_dot;
    }
    return o;
}

// ------------------------------------------------------- normalize(env) --
// envs/normalized_env.py:8-171 as a prologue / epilogue of env_step and env_reset (see b2g.h).
struct NormParams {
    int flags;
    float lo[2], hi[2], es;
    double scale_reward, oa, ra;
    double* stats;  // [2*O + 2, N]
};

inline bool norm_from_host(const b2g_norm* h, NormParams* q) {
    *q = NormParams{};
    q->scale_reward = 1.0;
    if (!h) return true;
    if ((h->flags & (B2G_NORM_OBS | B2G_NORM_REWARD)) && !h->stats) return false;
    q->flags = h->flags;
    for (int i = 0; i < 2; ++i) { q->lo[i] = h->act_low[i]; q->hi[i] = h->act_high[i]; }
    q->es = h->expected_action_scale;
    q->scale_reward = h->scale_reward; q->oa = h->obs_alpha; q->ra = h->reward_alpha;
    q->stats = h->stats;
    return true;
}
inline bool norm_active(const NormParams& q) { return q.flags != 0 || q.scale_reward != 1.0; }

template <int O>
struct NormStats {
    double om[O], ov[O], rm, rv;
    __device__ __forceinline__ void load(const NormParams& q, int64_t N, int64_t n) {
        if (q.flags & (B2G_NORM_OBS | B2G_NORM_REWARD)) {
#pragma unroll
            for (int i = 0; i < O; ++i) { om[i] = q.stats[(int64_t)i * N + n]; ov[i] = q.stats[(int64_t)(O + i) * N + n]; }
            rm = q.stats[(int64_t)(2 * O) * N + n];
            rv = q.stats[(int64_t)(2 * O + 1) * N + n];
        }
    }
    __device__ __forceinline__ void store(const NormParams& q, int64_t N, int64_t n) const {
        if (q.flags & B2G_NORM_OBS) {
#pragma unroll
            for (int i = 0; i < O; ++i) { q.stats[(int64_t)i * N + n] = om[i]; q.stats[(int64_t)(O + i) * N + n] = ov[i]; }
        }
        if (q.flags & B2G_NORM_REWARD) {
            q.stats[(int64_t)(2 * O) * N + n] = rm;
            q.stats[(int64_t)(2 * O + 1) * N + n] = rv;
        }
    }
    // _apply_normalize_obs (:133-146): update the estimate, then normalise with the NEW estimate.
    // Separate multiplies and adds (numpy evaluates them one by one: no fused multiply-add).
    __device__ __forceinline__ void obs(const NormParams& q, float* ob) {
        if (!(q.flags & B2G_NORM_OBS)) return;
#pragma unroll
        for (int i = 0; i < O; ++i) {
            const double o = (double)ob[i];
            om[i] = __dadd_rn(__dmul_rn(1.0 - q.oa, om[i]), __dmul_rn(q.oa, o));
            const double d = o - om[i];
            ov[i] = __dadd_rn(__dmul_rn(1.0 - q.oa, ov[i]), __dmul_rn(q.oa, __dmul_rn(d, d)));
            ob[i] = (float)(d / (sqrt(ov[i]) + 1e-8));
        }
    }
    // _apply_normalize_reward (:148-160) and scale_reward (:112)
    __device__ __forceinline__ double reward(const NormParams& q, double r) {
        if (q.flags & B2G_NORM_REWARD) {
            rm = __dadd_rn(__dmul_rn(1.0 - q.ra, rm), __dmul_rn(q.ra, r));
            const double d = r - rm;
            rv = __dadd_rn(__dmul_rn(1.0 - q.ra, rv), __dmul_rn(q.ra, __dmul_rn(d, d)));
            r = r / (sqrt(rv) + 1e-8);
        }
        return r * q.scale_reward;
    }
};

// step():91-101 -- the action the wrapped env receives
template <int A>
__device__ __forceinline__ void norm_action(const NormParams& q, const float* act, float* scaled) {
#pragma unroll
    for (int i = 0; i < A; ++i) {
        float v = act[i];
        if (q.flags & B2G_NORM_ACTION) {
            const float k = __fdiv_rn(__fmul_rn(0.5f, __fsub_rn(q.hi[i], q.lo[i])), q.es);
            v = __fadd_rn(q.lo[i], __fmul_rn(__fadd_rn(v, q.es), k));
            v = fminf(fmaxf(v, q.lo[i]), q.hi[i]);
        }
        scaled[i] = v;
    }
}
