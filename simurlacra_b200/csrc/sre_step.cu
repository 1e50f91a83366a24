// sre_step.cu -- step-granular kernels behind SimPyEnv.reset / step / observe, the domain-parameter setter and
// the Philox samplers.  One thread per environment; state lives in HBM between calls (the persistent rollout
// kernel in sre_rollout.cuh keeps it in registers instead).
#include "sre_host.h"

namespace sre {

static inline unsigned nblocks(long long n, int t) { return (unsigned)((n + t - 1) / t); }

template <int KIND, typename R>
__global__ void derive_kernel(const __grid_constant__ KCfg<R> cfg, const R *params, long long N, R *st_lo, R *st_up,
                              R *act_lo, R *act_up, R *c_max) {
    using E = Env<KIND, R>;
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    R dp[E::DP];
#pragma unroll
    for (int i = 0; i < E::DP; i++) dp[i] = params ? params[n * E::DP + i] : cfg.base_params[i];
    Lane<KIND, R> L;
    L.set_params(cfg, dp);
    R slo[E::DS], sup[E::DS], alo[E::DA], aup[E::DA];
    L.bounds(cfg.flags, slo, sup, alo, aup);
#pragma unroll
    for (int i = 0; i < E::DS; i++) {
        if (st_lo) st_lo[n * E::DS + i] = slo[i];
        if (st_up) st_up[n * E::DS + i] = sup[i];
    }
#pragma unroll
    for (int i = 0; i < E::DA; i++) {
        if (act_lo) act_lo[n * E::DA + i] = alo[i];
        if (act_up) act_up[n * E::DA + i] = aup[i];
    }
    if (c_max) c_max[n] = L.c_max;
}

template <typename R>
__global__ void sample_params_kernel(const __grid_constant__ KDistr distr, unsigned long long seed,
                                     unsigned long long env_offset, unsigned episode, R *params, int dp, long long N) {
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    for (int i = 0; i < distr.n; i++)
        params[n * dp + distr.e[i].param_index] =
            (R)sample_entry(distr.e[i], seed, env_offset + (unsigned long long)n, episode, (uint32_t)i);
}

template <int KIND, typename R>
__global__ void sample_init_kernel(const __grid_constant__ KCfg<R> cfg, const R *params, unsigned long long seed,
                                   unsigned long long env_offset, unsigned episode, R *init, long long N) {
    using E = Env<KIND, R>;
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    R dp[E::DP];
#pragma unroll
    for (int i = 0; i < E::DP; i++) dp[i] = params ? params[n * E::DP + i] : cfg.base_params[i];
    R v[E::DS];
    sample_init<KIND, R>(cfg, dp, seed, env_offset + (unsigned long long)n, episode, v);
#pragma unroll
    for (int i = 0; i < E::DINIT; i++) init[n * E::DINIT + i] = v[i];
}

template <int KIND, typename R>
__global__ void reset_kernel(const __grid_constant__ KCfg<R> cfg, const R *params, const R *init, int n_cols,
                             const unsigned char *mask, R *state, R *hidden, int *curr_step, R *obs,
                             unsigned long long seed, unsigned long long env_offset, unsigned episode, long long N) {
    using E = Env<KIND, R>;
    constexpr int NHS = E::NH > 0 ? E::NH : 1;
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N || (mask && !mask[n])) return;
    R dp[E::DP];
#pragma unroll
    for (int i = 0; i < E::DP; i++) dp[i] = params ? params[n * E::DP + i] : cfg.base_params[i];
    R iv[E::DS];
#pragma unroll
    for (int i = 0; i < E::DS; i++)
        if (i < n_cols) iv[i] = init[n * n_cols + i];
    Lane<KIND, R> L;
    L.reset(cfg, dp, iv, n_cols);
#pragma unroll
    for (int i = 0; i < E::DS; i++) state[n * E::DS + i] = L.s[i];
#pragma unroll
    for (int i = 0; i < NHS; i++) hidden[n * NHS + i] = (i < E::NH) ? L.h[i] : (R)0;
    curr_step[n] = 0;
    if (obs) {
        R o[E::DO];
        L.observe(o);
        if (cfg.obs_noise) add_obs_noise<E::DO, R>(cfg, seed, env_offset + (unsigned long long)n, episode, 0u, o);
        if (cfg.obs_norm) apply_obs_norm<E::DO, R>(cfg, o);
#pragma unroll
        for (int i = 0; i < E::DO; i++) obs[n * E::DO + i] = o[i];
    }
}

template <int KIND, typename R>
__global__ void step_kernel(const __grid_constant__ KCfg<R> cfg, const R *params, const R *act, R *state, R *hidden,
                            int *curr_step, R *obs, R *rew, unsigned char *done, int *status, unsigned long long seed,
                            unsigned long long env_offset, unsigned episode, long long N) {
    using E = Env<KIND, R>;
    constexpr int NHS = E::NH > 0 ? E::NH : 1;
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    R dp[E::DP];
#pragma unroll
    for (int i = 0; i < E::DP; i++) dp[i] = params ? params[n * E::DP + i] : cfg.base_params[i];
    Lane<KIND, R> L;
    L.set_params(cfg, dp);
#pragma unroll
    for (int i = 0; i < E::DS; i++) L.s[i] = state[n * E::DS + i];
#pragma unroll
    for (int i = 0; i < E::NH; i++) L.h[i] = hidden[n * NHS + i];
    // FinalRewTask yields its final reward ONCE per episode (tasks/final_reward.py:131-135): the flag travels between
    // step launches in bit 30 of the step counter (SRE_STEP_YIELDED_FINAL), cleared by sre_reset
    const int cs = curr_step[n];
    L.curr_step = cs & SRE_STEP_COUNT_MASK;
    L.yielded_final = (cs & SRE_STEP_YIELDED_FINAL) ? 1 : 0;
    R a[E::DA];
    int st = 0;
#pragma unroll
    for (int i = 0; i < E::DA; i++) {
        const R v = act[n * E::DA + i];
        if (m_isnan(v)) st |= SRE_ST_NAN_ACT;
        a[i] = v;
    }
    const unsigned long long genv = env_offset + (unsigned long long)n;
    if (cfg.env_act_noise == SRE_ENV_ACT_NOISE_OUTER) add_env_act_noise<E::DA, R>(cfg, seed, genv, episode, (uint32_t)L.curr_step, a);
    R a_lo[E::DA], a_up[E::DA];
    L.act_bounds(cfg.flags, a_lo, a_up);
#pragma unroll
    for (int i = 0; i < E::DA; i++)
        if (cfg.act_norm) a[i] = a_lo[i] + (a[i] + (R)1) * (a_up[i] - a_lo[i]) / (R)2;
    if (cfg.env_act_noise == SRE_ENV_ACT_NOISE_INNER) add_env_act_noise<E::DA, R>(cfg, seed, genv, episode, (uint32_t)L.curr_step, a);
    R r;
    int inside;
    R o[E::DO];
    L.observe(o);  // refreshes the trig terms of the loaded state
    const bool d = L.step(cfg, a, r, inside);
    if (inside < 0) st |= SRE_ST_NAN_STATE;
#pragma unroll
    for (int i = 0; i < E::DS; i++) state[n * E::DS + i] = L.s[i];
#pragma unroll
    for (int i = 0; i < E::NH; i++) hidden[n * NHS + i] = L.h[i];
    curr_step[n] = L.curr_step | (L.yielded_final ? SRE_STEP_YIELDED_FINAL : 0);
    L.observe(o);
    if (cfg.obs_noise)
        add_obs_noise<E::DO, R>(cfg, seed, env_offset + (unsigned long long)n, episode, (uint32_t)L.curr_step, o);
    if (cfg.obs_norm) apply_obs_norm<E::DO, R>(cfg, o);
#pragma unroll
    for (int i = 0; i < E::DO; i++) obs[n * E::DO + i] = o[i];
    rew[n] = r;
    done[n] = d ? 1 : 0;
    if (status) status[n] = st;
}

template <int KIND, typename R> __global__ void observe_kernel(const R *state, R *obs, long long N) {
    using E = Env<KIND, R>;
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    R s[E::DS], o[E::DO], trig[2];
#pragma unroll
    for (int i = 0; i < E::DS; i++) s[i] = state[n * E::DS + i];
    E::observe(s, o, trig);
#pragma unroll
    for (int i = 0; i < E::DO; i++) obs[n * E::DO + i] = o[i];
}

// Task.step_rew for a batch of (state, action) pairs (tasks/desired_state.py:107-155, tasks/reward_functions.py)
template <int KIND, typename R>
__global__ void step_rew_kernel(const __grid_constant__ KCfg<R> cfg, const R *params, const R *state, const R *act,
                                R *rew, long long N) {
    using E = Env<KIND, R>;
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    R dp[E::DP];
#pragma unroll
    for (int i = 0; i < E::DP; i++) dp[i] = params ? params[n * E::DP + i] : cfg.base_params[i];
    Lane<KIND, R> L;
    L.set_params(cfg, dp);
    R s[E::DS], a[E::DA];
#pragma unroll
    for (int i = 0; i < E::DS; i++) s[i] = state[n * E::DS + i];
#pragma unroll
    for (int i = 0; i < E::DA; i++) a[i] = act[n * E::DA + i];
    rew[n] = step_rew<E::DS, E::DA, R>(cfg.task, s, a, L.c_max);
}

// Task.step_rew for a free-standing task (no env): runtime dims, same arithmetic as step_rew<DS, DA, R>
template <typename R>
__global__ void task_rew_kernel(const __grid_constant__ KTask<R> tk, int ds, int da, R c_max, const R *state,
                                const R *act, R *rew, long long N) {
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    R err[SRE_MAX_DS], a[SRE_MAX_DA];
    for (int i = 0; i < ds; i++) err[i] = tk.state_des[i] - state[n * ds + i];
    for (int i = 0; i < da; i++) a[i] = act[n * da + i];
    if (tk.radial_mask) {
        const R pi = (R)SRE_PI, two_pi = (R)(2 * SRE_PI);
        for (int i = 0; i < ds; i++)
            if (tk.radial_mask & (1u << i)) err[i] = m_fmod_pos(err[i], tk.radial_mod[i], tk.radial_inv[i]);
        for (int i = 0; i < ds; i++) {
            if (err[i] > pi) err[i] = two_pi - err[i];
            if (err[i] < -pi) err[i] = -two_pi - err[i];
        }
    }
    R cost = (R)0;
    for (int i = 0; i < ds; i++) {
        R t = (R)0;
        for (int j = 0; j < ds; j++) t += tk.Q[i * ds + j] * err[j];
        cost += err[i] * t;
    }
    for (int i = 0; i < da; i++) {
        R t = (R)0;
        for (int j = 0; j < da; j++) t += tk.Rm[i * da + j] * a[j];
        cost += a[i] * t;
    }
    rew[n] = tk.rew_kind == SRE_REW_QUADR ? -cost : (tk.rew_kind == SRE_REW_EXP ? m_exp(-cost) : m_exp(-c_max * cost));
}

template <typename R>
static int task_rew_t(const sre_task *task, int ds, int da, double c_max, const void *state, const void *act, void *rew,
                      long long n, cudaStream_t s) {
    KTask<R> k;
    memset(&k, 0, sizeof k);
    for (int i = 0; i < ds; i++) {
        k.state_des[i] = (R)task->state_des[i];
        k.radial_mod[i] = (R)std::fabs(task->radial_mod[i]);
        k.radial_inv[i] = (R)(1.0 / (double)k.radial_mod[i]);
        for (int j = 0; j < ds; j++) k.Q[i * ds + j] = (R)task->Q[i * ds + j];
    }
    for (int i = 0; i < da * da; i++) k.Rm[i] = (R)task->R[i];
    k.rew_kind = task->rew_kind;
    k.radial_mask = task->radial_mask;
    task_rew_kernel<R><<<nblocks(n, 128), 128, 0, s>>>(k, ds, da, (R)c_max, (const R *)state, (const R *)act, (R *)rew, n);
    return (int)cudaGetLastError();
}
int task_rew_dispatch(const sre_task *task, int dtype, int ds, int da, double c_max, const void *state, const void *act,
                      void *rew, long long n, cudaStream_t s) {
    return dtype == SRE_F64 ? task_rew_t<double>(task, ds, da, c_max, state, act, rew, n, s)
                            : task_rew_t<float>(task, ds, da, c_max, state, act, rew, n, s);
}

// Policy.forward for a batch (float32): weights read from global memory through the read-only path; this is
// the step-granular entry, the rollout kernel has its own register-resident fast path.
__global__ void policy_forward_kernel(const __grid_constant__ KPolicy pol, const float *obs, int n_obs, float *act,
                                      int n_act, long long N) {
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    float o[SRE_MAX_WIDTH], a[SRE_MAX_WIDTH];
    for (int i = 0; i < n_obs; i++) o[i] = obs[n * n_obs + i];
    if (pol.kind == SRE_POL_FNN) fnn_generic_forward(pol, pol.weights, o, n_obs, a, n_act);
    else if (pol.kind == SRE_POL_LINEAR) linear_forward_rt(pol, pol.weights, o, n_obs, a, n_act);
    else
        for (int i = 0; i < n_act; i++) a[i] = 0.f;
    for (int i = 0; i < n_act; i++) act[n * n_act + i] = a[i];
}

// ---- host-side dispatch over (kind, dtype) ---------------------------------------------------------------
#define SRE_DISPATCH_KIND(KINDVAR, R, BODY)                               \
    switch (KINDVAR) {                                                    \
    case SRE_QQ: { constexpr int K = SRE_QQ; BODY } break;                \
    case SRE_QCP: { constexpr int K = SRE_QCP; BODY } break;              \
    case SRE_QBB: { constexpr int K = SRE_QBB; BODY } break;              \
    case SRE_BOB: { constexpr int K = SRE_BOB; BODY } break;              \
    case SRE_OMO: { constexpr int K = SRE_OMO; BODY } break;              \
    default: { constexpr int K = SRE_PEND; BODY } break;                  \
    }


template <typename R>
static int derive_t(const sre_env_cfg *cfg, const void *params, long long n, void *st_lo, void *st_up, void *act_lo,
                    void *act_up, void *c_max, cudaStream_t s) {
    KCfg<R> k;
    make_kcfg<R>(*cfg, k);
    SRE_DISPATCH_KIND(cfg->kind, R, (derive_kernel<K, R><<<nblocks(n, 128), 128, 0, s>>>(
                                        k, (const R *)params, n, (R *)st_lo, (R *)st_up, (R *)act_lo, (R *)act_up,
                                        (R *)c_max));)
    return (int)cudaGetLastError();
}
int derive_dispatch(const sre_env_cfg *cfg, const void *params, long long n, void *st_lo, void *st_up, void *act_lo,
                    void *act_up, void *c_max, cudaStream_t s) {
    return cfg->dtype == SRE_F64 ? derive_t<double>(cfg, params, n, st_lo, st_up, act_lo, act_up, c_max, s)
                                 : derive_t<float>(cfg, params, n, st_lo, st_up, act_lo, act_up, c_max, s);
}

int sample_params_dispatch(const sre_env_cfg *cfg, const KDistr &kd, unsigned long long seed,
                           unsigned long long env_offset, unsigned episode, void *params, long long n,
                           cudaStream_t s) {
    const int dp = k_dp[cfg->kind];
    if (cfg->dtype == SRE_F64)
        sample_params_kernel<double><<<nblocks(n, 128), 128, 0, s>>>(kd, seed, env_offset, episode, (double *)params, dp, n);
    else
        sample_params_kernel<float><<<nblocks(n, 128), 128, 0, s>>>(kd, seed, env_offset, episode, (float *)params, dp, n);
    return (int)cudaGetLastError();
}

template <typename R>
static int sample_init_t(const sre_env_cfg *cfg, const void *params, unsigned long long seed,
                         unsigned long long env_offset, unsigned episode, void *init, long long n, cudaStream_t s) {
    KCfg<R> k;
    make_kcfg<R>(*cfg, k);
    SRE_DISPATCH_KIND(cfg->kind, R, (sample_init_kernel<K, R><<<nblocks(n, 128), 128, 0, s>>>(
                                        k, (const R *)params, seed, env_offset, episode, (R *)init, n));)
    return (int)cudaGetLastError();
}
int sample_init_dispatch(const sre_env_cfg *cfg, const void *params, unsigned long long seed,
                         unsigned long long env_offset, unsigned episode, void *init, long long n, cudaStream_t s) {
    return cfg->dtype == SRE_F64 ? sample_init_t<double>(cfg, params, seed, env_offset, episode, init, n, s)
                                 : sample_init_t<float>(cfg, params, seed, env_offset, episode, init, n, s);
}

template <typename R>
static int reset_t(const sre_env_cfg *cfg, const void *params, const void *init, int n_cols, const unsigned char *mask,
                   void *state, void *hidden, int *curr_step, void *obs, unsigned long long seed,
                   unsigned long long env_offset, unsigned episode, long long n, cudaStream_t s) {
    KCfg<R> k;
    make_kcfg<R>(*cfg, k);
    SRE_DISPATCH_KIND(cfg->kind, R, (reset_kernel<K, R><<<nblocks(n, 128), 128, 0, s>>>(
                                        k, (const R *)params, (const R *)init, n_cols, mask, (R *)state, (R *)hidden,
                                        curr_step, (R *)obs, seed, env_offset, episode, n));)
    return (int)cudaGetLastError();
}
int reset_dispatch(const sre_env_cfg *cfg, const void *params, const void *init, int n_cols, const unsigned char *mask,
                   void *state, void *hidden, int *curr_step, void *obs, unsigned long long seed,
                   unsigned long long env_offset, unsigned episode, long long n, cudaStream_t s) {
    return cfg->dtype == SRE_F64
               ? reset_t<double>(cfg, params, init, n_cols, mask, state, hidden, curr_step, obs, seed, env_offset, episode, n, s)
               : reset_t<float>(cfg, params, init, n_cols, mask, state, hidden, curr_step, obs, seed, env_offset, episode, n, s);
}

template <typename R>
static int step_t(const sre_env_cfg *cfg, const void *params, const void *act, void *state, void *hidden,
                  int *curr_step, void *obs, void *rew, unsigned char *done, int *status, unsigned long long seed,
                  unsigned long long env_offset, unsigned episode, long long n, cudaStream_t s) {
    KCfg<R> k;
    make_kcfg<R>(*cfg, k);
    SRE_DISPATCH_KIND(cfg->kind, R, (step_kernel<K, R><<<nblocks(n, 128), 128, 0, s>>>(
                                        k, (const R *)params, (const R *)act, (R *)state, (R *)hidden, curr_step,
                                        (R *)obs, (R *)rew, done, status, seed, env_offset, episode, n));)
    return (int)cudaGetLastError();
}
int step_dispatch(const sre_env_cfg *cfg, const void *params, const void *act, void *state, void *hidden,
                  int *curr_step, void *obs, void *rew, unsigned char *done, int *status, unsigned long long seed,
                  unsigned long long env_offset, unsigned episode, long long n, cudaStream_t s) {
    return cfg->dtype == SRE_F64
               ? step_t<double>(cfg, params, act, state, hidden, curr_step, obs, rew, done, status, seed, env_offset, episode, n, s)
               : step_t<float>(cfg, params, act, state, hidden, curr_step, obs, rew, done, status, seed, env_offset, episode, n, s);
}

template <typename R>
static int observe_t(const sre_env_cfg *cfg, const void *state, void *obs, long long n, cudaStream_t s) {
    SRE_DISPATCH_KIND(cfg->kind, R,
                      (observe_kernel<K, R><<<nblocks(n, 128), 128, 0, s>>>((const R *)state, (R *)obs, n));)
    return (int)cudaGetLastError();
}
int observe_dispatch(const sre_env_cfg *cfg, const void *state, void *obs, long long n, cudaStream_t s) {
    return cfg->dtype == SRE_F64 ? observe_t<double>(cfg, state, obs, n, s) : observe_t<float>(cfg, state, obs, n, s);
}

template <typename R>
static int step_rew_t(const sre_env_cfg *cfg, const void *params, const void *state, const void *act, void *rew,
                      long long n, cudaStream_t s) {
    KCfg<R> k;
    make_kcfg<R>(*cfg, k);
    SRE_DISPATCH_KIND(cfg->kind, R, (step_rew_kernel<K, R><<<nblocks(n, 128), 128, 0, s>>>(
                                        k, (const R *)params, (const R *)state, (const R *)act, (R *)rew, n));)
    return (int)cudaGetLastError();
}
int step_rew_dispatch(const sre_env_cfg *cfg, const void *params, const void *state, const void *act, void *rew,
                      long long n, cudaStream_t s) {
    return cfg->dtype == SRE_F64 ? step_rew_t<double>(cfg, params, state, act, rew, n, s)
                                 : step_rew_t<float>(cfg, params, state, act, rew, n, s);
}

bool policy_is_tc_shape(const KPolicy &kp, int n_obs, int n_act);
int policy_forward_tc_dispatch(const KPolicy &kp, const float *obs, int n_obs, float *act, int n_act, long long n,
                               cudaStream_t s);

int policy_forward_dispatch(const KPolicy &kp, const float *obs, int n_obs, float *act, int n_act, long long n,
                            cudaStream_t s) {
    if (policy_is_tc_shape(kp, n_obs, n_act)) {
        const int r = policy_forward_tc_dispatch(kp, obs, n_obs, act, n_act, n, s);
        if (r != -1000) return r;
    }
    policy_forward_kernel<<<nblocks(n, 128), 128, 0, s>>>(kp, obs, n_obs, act, n_act, n);
    return (int)cudaGetLastError();
}

}  // namespace sre
