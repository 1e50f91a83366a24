// sre_core.cuh -- the per-environment "lane": everything SimPyEnv holds for one env (derived constants, spaces,
// task state, state, hidden state, step counter) as a struct of registers, plus reset/step with the reference's
// semantics (environments/pysim/base.py:166-241) and the Task/RewFcn arithmetic (tasks/*.py).
#pragma once
#include "sre_envs.cuh"

namespace sre {

// Kernel-side copy of sre_task / sre_env_cfg in the working real type (host converts once per launch).
template <typename R> struct KTask {
    R state_des[SRE_MAX_DS];
    R Q[SRE_MAX_DS * SRE_MAX_DS];
    R Rm[SRE_MAX_DA * SRE_MAX_DA];
    R radial_mod[SRE_MAX_DS];
    R radial_inv[SRE_MAX_DS];  // 1 / radial_mod (host-computed, for m_fmod_pos)
    R final_factor;
    double min_rew;
    int rew_kind;
    unsigned radial_mask;
    int final_flags;
    int q_diag;  // Q and R are diagonal: 3*ds + 3*da flops instead of the dense form
};

#define SRE_MAX_POL_IN (SRE_MAX_DO + SRE_MAX_AUG)  // widest policy input with StateAugmentationWrapper
template <typename R> struct KWrap {
    int active, delay, delay_inner, delay_first, down, keep_mask, vf_n, n_aug, aug_fixed;
    int vf_pos[SRE_MAX_VF], vf_vel[SRE_MAX_VF], aug_param[SRE_MAX_AUG];
    R vf_b0, vf_b1, vf_a1, vf_zi;
    R aug_nominal[SRE_MAX_AUG];
};

template <typename R> struct KCfg {
    int flags, wild_init, act_norm, obs_noise, has_init_box, obs_norm, env_act_noise;
    long long max_steps;  // <= 0: inf
    R dt;
    KTask<R> task;
    R base_params[SRE_MAX_DP];
    R noise_std[SRE_MAX_DO], noise_mean[SRE_MAX_DO];
    R obs_lb[SRE_MAX_DO], obs_span[SRE_MAX_DO];  // ObsNormWrapper: lower bound and (ub - lb)
    R an_std[SRE_MAX_DA], an_mean[SRE_MAX_DA];    // GaussianActNoiseWrapper
    double init_box_lo[SRE_MAX_DS], init_box_up[SRE_MAX_DS];
    KWrap<R> wrap;
};

struct KDistr {
    int n;
    sre_distr e[SRE_MAX_DISTR];
    signed char entry_of_param[SRE_MAX_DP];  // last table entry that writes parameter j, or -1 (host-filled)
};

// ------------------------------------------------------------------------------------------------------
// Task.step_rew: tasks/desired_state.py:107-110 (DesStateTask), :146-155 (RadiallySymmDesStateTask: fmod on the
// selected indices, then the +-pi wrap on EVERY element), reward functions tasks/reward_functions.py:202-222
// (Quadr), :237-244 (ExpQuadr), :276-282 (ScaledExpQuadr).  act is the UNCLIPPED action (base.py:220).
// ------------------------------------------------------------------------------------------------------
// PLAIN (compile time): diagonal Q / R known to the caller -- the dense branch is not compiled in.
template <int DS, int DA, typename R, bool PLAIN = false>
SRE_DEV R step_rew(const KTask<R> &tk, const R *state, const R *act, R c_max) {
    R err[DS];
#pragma unroll
    for (int i = 0; i < DS; i++) err[i] = tk.state_des[i] - state[i];
    if (tk.radial_mask) {
        const R pi = (R)SRE_PI, two_pi = (R)(2 * SRE_PI);
#pragma unroll
        for (int i = 0; i < DS; i++)
            if (tk.radial_mask & (1u << i)) err[i] = m_fmod_pos(err[i], tk.radial_mod[i], tk.radial_inv[i]);
#pragma unroll
        for (int i = 0; i < DS; i++) {
            if (err[i] > pi) err[i] = two_pi - err[i];
            if (err[i] < -pi) err[i] = -two_pi - err[i];
        }
    }
    R cost = (R)0;
    if (PLAIN || tk.q_diag) {
#pragma unroll
        for (int i = 0; i < DS; i++) cost += tk.Q[i * DS + i] * (err[i] * err[i]);
#pragma unroll
        for (int i = 0; i < DA; i++) cost += tk.Rm[i * DA + i] * (act[i] * act[i]);
    } else {
#pragma unroll
        for (int i = 0; i < DS; i++) {
            R t = (R)0;
#pragma unroll
            for (int j = 0; j < DS; j++) t += tk.Q[i * DS + j] * err[j];
            cost += err[i] * t;
        }
#pragma unroll
        for (int i = 0; i < DA; i++) {
            R t = (R)0;
#pragma unroll
            for (int j = 0; j < DA; j++) t += tk.Rm[i * DA + j] * act[j];  // (-a)^T R (-a) == a^T R a
            cost += act[i] * t;
        }
    }
    if (tk.rew_kind == SRE_REW_QUADR) return -cost;
    if (tk.rew_kind == SRE_REW_EXP) return m_exp(-cost);
    return m_exp(-c_max * cost);
}

// ScaledExpQuadrErrRewFcn.reset (reward_functions.py:284-297): c_max from the CURRENT (param dependent) spaces,
// computed in fp64 once per episode
template <int DS, int DA, typename R>
SRE_DEV R compute_cmax(const KTask<R> &tk, const R *st_lo, const R *st_up, const R *act_lo, const R *act_up) {
    if (tk.rew_kind != SRE_REW_SCALED_EXP) return (R)1;
    double smax[DS], amax[DA];
#pragma unroll
    for (int i = 0; i < DS; i++) smax[i] = fmax(fabs((double)st_lo[i]), fabs((double)st_up[i]));
#pragma unroll
    for (int i = 0; i < DA; i++) amax[i] = fmax(fabs((double)act_lo[i]), fabs((double)act_up[i]));
    double cost = 0.0;
#pragma unroll
    for (int i = 0; i < DS; i++) {
        double t = 0.0;
#pragma unroll
        for (int j = 0; j < DS; j++) t += (double)tk.Q[i * DS + j] * smax[j];
        cost += smax[i] * t;
    }
#pragma unroll
    for (int i = 0; i < DA; i++) {
        double t = 0.0;
#pragma unroll
        for (int j = 0; j < DA; j++) t += (double)tk.Rm[i * DA + j] * amax[j];
        cost += amax[i] * t;
    }
    return (R)(-1.0 * log(tk.min_rew) / cost);
}

// ------------------------------------------------------------------------------------------------------
// One environment.
// ------------------------------------------------------------------------------------------------------
template <int KIND, typename R> struct Lane {
    using E = Env<KIND, R>;
    typename E::Consts c;  // the spaces are rebuilt from these on use (E::bounds): most bounds are compile-time constants,
                           // the few that depend on a domain parameter are members of Consts -- no bound lives in a register
    R c_max;
    R s[E::DS];
    R h[E::NH > 0 ? E::NH : 1];
    R trig[2];  // sin / cos the dynamics share with the observation of the current state (QQ: al, QCP: th)
    int curr_step;
    int yielded_final;

    // Env.observe (environments/base.py:203 and the overrides); also refreshes `trig`
    SRE_DEV void observe(R *obs) { E::observe(s, obs, trig); }

    // A prepared episode start parked in shared memory (rollout kernel): everything set_params + reset produce, one
    // column per lane (`p` points at the lane's column, consecutive fields are `stride` reals apart: conflict-free).
    static constexpr int NC = sizeof(typename E::Consts) / sizeof(R);
    static constexpr int NHS = E::NH > 0 ? E::NH : 1;
    static constexpr int NSLOT = NC + E::DS + 1 + E::NH;
    SRE_DEV void bounds(int flags, R *st_lo, R *st_up, R *act_lo, R *act_up) const { E::bounds(c, flags, st_lo, st_up, act_lo, act_up); }
    SRE_DEV void act_bounds(int flags, R *act_lo, R *act_up) const {
        R st_lo[E::DS], st_up[E::DS];
        E::bounds(c, flags, st_lo, st_up, act_lo, act_up);
    }
    SRE_DEV void store_slot(R *p, int stride) const {
        const R *cc = reinterpret_cast<const R *>(&c);
        int k = 0;
#pragma unroll
        for (int i = 0; i < NC; i++) p[(k++) * stride] = cc[i];
#pragma unroll
        for (int i = 0; i < E::DS; i++) p[(k++) * stride] = s[i];
        p[(k++) * stride] = c_max;
#pragma unroll
        for (int i = 0; i < E::NH; i++) p[(k++) * stride] = h[i];
    }
    SRE_DEV void load_slot(const R *p, int stride) {
        R *cc = reinterpret_cast<R *>(&c);
        int k = 0;
#pragma unroll
        for (int i = 0; i < NC; i++) cc[i] = p[(k++) * stride];
#pragma unroll
        for (int i = 0; i < E::DS; i++) s[i] = p[(k++) * stride];
        c_max = p[(k++) * stride];
#pragma unroll
        for (int i = 0; i < E::NH; i++) h[i] = p[(k++) * stride];
        if (E::NH == 0) h[0] = (R)0;
        curr_step = 0;
        yielded_final = 0;
    }

    // domain_param setter: _calc_constants + _create_spaces + task.reset (base.py:112-124)
    SRE_DEV void set_params(const KCfg<R> &cfg, const R *dp) {
        E::derive(dp, cfg.flags, c);
        R st_lo[E::DS], st_up[E::DS], act_lo[E::DA], act_up[E::DA];
        E::bounds(c, cfg.flags, st_lo, st_up, act_lo, act_up);
        c_max = compute_cmax<E::DS, E::DA, R>(cfg.task, st_lo, st_up, act_lo, act_up);
    }

    // reset with a GIVEN init (base.py:166-203); QCP zeroes th_ddot (quanser_cartpole.py:101-106), QBB runs the
    // inverse kinematics (quanser_ball_balancer.py:231-245)
    SRE_DEV void reset(const KCfg<R> &cfg, const R *dp, const R *init, int n_cols) {
        reset_state(init, n_cols);
        E::reset_hidden(dp, cfg.flags, s, h);
    }
    // the part of reset() that does not depend on the hidden state (the rollout kernel computes QBB's plate angles
    // warp-cooperatively afterwards)
    SRE_DEV void reset_state(const R *init, int n_cols) {
        curr_step = 0;
        yielded_final = 0;
        E::state_from_init(init, n_cols, s);
        h[0] = (R)0;
    }

    // BoxSpace.contains (spaces/box.py:138-167): inclusive bounds; 1 inside, 0 outside, -1 NaN
    SRE_DEV int contains(int flags) const {
        R st_lo[E::DS], st_up[E::DS], act_lo[E::DA], act_up[E::DA];
        E::bounds(c, flags, st_lo, st_up, act_lo, act_up);
        bool nan = false, in = true;
#pragma unroll
        for (int i = 0; i < E::DS; i++) {
            nan |= m_isnan(s[i]);
            in &= (s[i] >= st_lo[i]) && (s[i] <= st_up[i]);
        }
        return nan ? -1 : (in ? 1 : 0);
    }

    // FinalRewTask.compute_final_rew (tasks/final_reward.py:111-275) with success_fcn = never_succeeded
    SRE_DEV R final_rew(const KCfg<R> &cfg, bool failed, long long remaining) {
        const int ff = cfg.task.final_flags;
        if (!(ff & SRE_FR_ENABLED) || yielded_final) return (R)0;
        yielded_final = 1;
        if (!failed || (ff & SRE_FR_ALWAYS_POS)) return (R)0;
        const R scale = (ff & SRE_FR_TIME_DEP) ? (R)remaining : cfg.task.final_factor;
        if (ff & SRE_FR_STATE_DEP) {
            R zero_act[E::DA];
#pragma unroll
            for (int i = 0; i < E::DA; i++) zero_act[i] = (R)0;
            return -scale * m_abs(step_rew<E::DS, E::DA, R>(cfg.task, s, zero_act, c_max));
        }
        return -scale;
    }

    // Env.limit_act (environments/base.py:215-222): BoxSpace.project_to clips (box.py:180); the DiscreteSpace of
    // BallOnBeamDiscSim keeps an action that np.isclose()s one of its three elements and else takes the closest
    // element, the first one on ties (discrete.py:119-127).
    SRE_DEV void limit_act(const KCfg<R> &cfg, const R *act, R *a) const {
        R act_lo[E::DA], act_up[E::DA];
        act_bounds(cfg.flags, act_lo, act_up);
        if (KIND == SRE_BOB && (cfg.flags & SRE_F_DISC_ACT)) {  // the flag exists for the ball-on-beam only
            const R lo = act_lo[0], up = act_up[0], x = act[0], mid = lo + (up - lo) / (R)2;
            const R d0 = m_abs(x - lo), d1 = m_abs(x - mid), d2 = m_abs(x - up);
            const R tol = (R)1e-8 + (R)1e-5 * m_abs(x);
            R e = lo, d = d0;
            if (d1 < d) { e = mid; d = d1; }
            if (d2 < d) e = up;
            a[0] = (d0 <= tol || d1 <= tol || d2 <= tol) ? x : e;
            return;
        }
#pragma unroll
        for (int i = 0; i < E::DA; i++) a[i] = m_min(m_max(act[i], act_lo[i]), act_up[i]);
    }

    // SimPyEnv.step (base.py:217-241).  act: inner-env units, unclipped.  Returns done; *inside gets the
    // contains() code (-1 = NaN state, the reference raises ValueErr at spaces/box.py:142).
    // PLAIN (compile time): no FinalRewTask and diagonal Q / R (the caller checked), nothing of either is compiled in
    template <bool PLAIN = false> SRE_DEV bool step(const KCfg<R> &cfg, const R *act, R &rew, int &inside) {
        const long long remaining = (!PLAIN && cfg.max_steps > 0) ? cfg.max_steps - (curr_step + 1) : 0;
        R r = step_rew<E::DS, E::DA, R, PLAIN>(cfg.task, s, act, c_max);
        R a[E::DA];
        limit_act(cfg, act, a);
        E::step(c, cfg.dt, cfg.flags, s, h, a, trig);  // trig belongs to the pre-step state
        curr_step += 1;
        inside = contains(cfg.flags);
        bool done = (inside != 1);  // never_succeeded: done == has_failed (tasks/base.py:159-172)
        const bool failed = done;
        if (cfg.max_steps > 0 && curr_step >= cfg.max_steps) done = true;
        if (!PLAIN && done) r += final_rew(cfg, failed, remaining);
        rew = r;
        return done;
    }
};

// init_space.sample_uniform() for (seed, env, episode): spaces/box.py:169-178, polar.py:108-113,
// compound.py:83-87.  fp64 like numpy; result in the working type.
template <int KIND, typename R>
SRE_DEV void sample_init(const KCfg<R> &cfg, const R *dp, uint64_t seed, uint64_t env, uint32_t episode, R *init) {
    using E = Env<KIND, R>;
    const u32x4 x = rng_block(seed, env, episode, SRE_STREAM_INIT, 0);
    const u32x4 y = rng_block(seed, env, episode, SRE_STREAM_INIT, 1);
    const double u[4] = {u53(x.x, x.y), u53(x.z, x.w), u53(y.x, y.y), u53(y.z, y.w)};
    double lo[4], up[4];
    if (cfg.has_init_box) {
#pragma unroll
        for (int i = 0; i < E::DINIT; i++) { lo[i] = cfg.init_box_lo[i]; up[i] = cfg.init_box_up[i]; }
    } else {
        bool right = false;
        if (KIND == SRE_BOB) right = (rng_block(seed, env, episode, SRE_STREAM_INIT, 2).x >> 31) != 0;
        E::init_box(dp, cfg.wild_init, right, lo, up);
    }
    double v[4];
#pragma unroll
    for (int i = 0; i < E::DINIT; i++) v[i] = lo[i] + (up[i] - lo[i]) * u[i];
    if (KIND == SRE_QBB) {  // Polar2DPosVelSpace.sample_uniform, spaces/polar.py:108-113
        double sn, cs;
        sincos(v[1], &sn, &cs);
        const double r = v[0];
        v[0] = r * cs;
        v[1] = r * sn;
    }
#pragma unroll
    for (int i = 0; i < E::DINIT; i++) init[i] = (R)v[i];
}

// GaussianObsNoiseWrapper._process_obs (environment_wrappers/observation_noise.py:68-73): obs + randn*std + mean,
// at reset and at every step; Philox block index = 2*t + i/4
template <int DO, typename R>
SRE_DEV void add_obs_noise(const KCfg<R> &cfg, uint64_t seed, uint64_t env, uint32_t episode, uint32_t t, R *obs) {
#pragma unroll
    for (int i = 0; i < DO; i += 4) {
        const u32x4 x = rng_block(seed, env, episode, SRE_STREAM_OBS_NOISE, t * 2u + (uint32_t)(i / 4));
        R z[4];
        normal4<R>(x, z);
#pragma unroll
        for (int j = 0; j < 4; j++)
            if (i + j < DO) obs[i + j] += z[j] * cfg.noise_std[i + j] + cfg.noise_mean[i + j];
    }
}

// ObsNormWrapper._process_obs (observation_normalization.py:119-122), same operation order
template <int DO, typename R> SRE_DEV void apply_obs_norm(const KCfg<R> &cfg, R *obs) {
#pragma unroll
    for (int i = 0; i < DO; i++) obs[i] = (obs[i] - cfg.obs_lb[i]) / cfg.obs_span[i] * (R)2 - (R)1;
}

// GaussianActNoiseWrapper._process_act (action_noise.py:73-78): act + (randn * std + mean) in the working type
template <int DA, typename R>
SRE_DEV void add_env_act_noise(const KCfg<R> &cfg, uint64_t seed, uint64_t env, uint32_t episode, uint32_t t, R *act) {
    static_assert(DA <= 2, "one Philox block carries the noise of at most two action dimensions");
    const u32x4 x = rng_block(seed, env, episode, SRE_STREAM_ENV_ACT_NOISE, t);
    R z[2];
    normal2_r(x.x, x.y, z[0], z[1]);
#pragma unroll
    for (int i = 0; i < DA; i++) act[i] = act[i] + (z[i] * cfg.an_std[i] + cfg.an_mean[i]);
}

// StochasticActionExplStrat.forward (exploration/stochastic_action.py:68-94): act = dist(policy(obs) + mean).rsample() in
// float32 with torch's roundings -- Normal.rsample: loc + eps * scale; Uniform.rsample: low + rand * (high - low) with
// low/high = loc -/+ halfspan.  One Philox block per (env, episode, step): words (x, y) -> the normal pair, or x, y ->
// the two uniforms.
template <int DA>
SRE_DEV void add_act_noise(int kind, const float *scale, const float *mean, uint64_t seed, uint64_t env, uint32_t episode,
                           uint32_t t, float *act) {
    static_assert(DA <= 2, "one Philox block carries the noise of at most two action dimensions");
    const u32x4 x = rng_block(seed, env, episode, SRE_STREAM_ACT_NOISE, t);
    if (kind == SRE_ACT_NOISE_NORMAL) {
        float z[2];
        normal2_f(x.x, x.y, z[0], z[1]);
#pragma unroll
        for (int i = 0; i < DA; i++) act[i] = __fadd_rn(__fadd_rn(act[i], mean[i]), __fmul_rn(z[i], scale[i]));
    } else {
        const float u[2] = {u24(x.x), u24(x.y)};
#pragma unroll
        for (int i = 0; i < DA; i++) {
            const float loc = __fadd_rn(act[i], mean[i]);
            const float lo = __fsub_rn(loc, scale[i]), hi = __fadd_rn(loc, scale[i]);
            act[i] = __fadd_rn(lo, __fmul_rn(u[i], __fsub_rn(hi, lo)));
        }
    }
}

}  // namespace sre
