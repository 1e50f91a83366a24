// sre_host.h -- host-side glue shared by the translation units: conversion of the C-ABI structs into the
// kernel-side structs, error reporting, per-(env, dtype) dispatch declarations.
#pragma once
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstring>

#include "sre_rollout.cuh"

namespace sre {

void set_error(const char *fmt, ...);
extern thread_local int g_launch_count;

#define SRE_CHECK_ARG(cond, code, ...)   \
    do {                                 \
        if (!(cond)) {                   \
            sre::set_error(__VA_ARGS__); \
            return (code);               \
        }                                \
    } while (0)

#define SRE_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            sre::set_error("%s failed: %s", #call, cudaGetErrorString(e_));              \
            return (int)e_;                                                              \
        }                                                                                \
    } while (0)

static const int k_ds[SRE_N_KINDS] = {4, 4, 8, 4, 2, 2};
static const int k_do[SRE_N_KINDS] = {6, 5, 8, 4, 2, 3};
static const int k_da[SRE_N_KINDS] = {1, 1, 2, 1, 1, 1};
static const int k_dp[SRE_N_KINDS] = {11, 17, 20, 8, 3, 5};
static const int k_nh[SRE_N_KINDS] = {0, 1, 2, 0, 0, 0};
static const int k_dinit[SRE_N_KINDS] = {4, 4, 4, 4, 2, 2};

template <typename R> inline void make_kcfg(const sre_env_cfg &c, KCfg<R> &k) {
    memset(&k, 0, sizeof(k));
    const int ds = k_ds[c.kind], da = k_da[c.kind];
    k.flags = c.flags;
    k.wild_init = c.wild_init;
    k.act_norm = c.act_norm;
    k.obs_noise = c.obs_noise;
    k.has_init_box = c.has_init_box;
    k.obs_norm = c.obs_norm;
    k.env_act_noise = c.env_act_noise;
    for (int i = 0; i < SRE_MAX_DO; i++) {
        k.obs_lb[i] = (R)c.obs_lb[i];
        k.obs_span[i] = (R)(c.obs_ub[i] - c.obs_lb[i]);
    }
    for (int i = 0; i < SRE_MAX_DA; i++) {
        k.an_std[i] = (R)c.env_act_noise_std[i];
        k.an_mean[i] = (R)c.env_act_noise_mean[i];
    }
    k.max_steps = c.max_steps;
    k.dt = (R)c.dt;
    bool diag = true;
    for (int i = 0; i < ds; i++) {
        k.task.state_des[i] = (R)c.task.state_des[i];
        k.task.radial_mod[i] = (R)std::fabs(c.task.radial_mod[i]);  // fmod(x, y) == fmod(x, |y|)
        k.task.radial_inv[i] = (R)(1.0 / (double)k.task.radial_mod[i]);
        for (int j = 0; j < ds; j++) {
            k.task.Q[i * ds + j] = (R)c.task.Q[i * ds + j];
            if (i != j && c.task.Q[i * ds + j] != 0.0) diag = false;
        }
    }
    for (int i = 0; i < da; i++)
        for (int j = 0; j < da; j++) {
            k.task.Rm[i * da + j] = (R)c.task.R[i * da + j];
            if (i != j && c.task.R[i * da + j] != 0.0) diag = false;
        }
    k.task.q_diag = diag ? 1 : 0;
    k.task.final_factor = (R)std::fabs(c.task.final_factor);  // final_reward.py:99
    k.task.min_rew = c.task.min_rew;
    k.task.rew_kind = c.task.rew_kind;
    k.task.radial_mask = c.task.radial_mask;
    k.task.final_flags = c.task.final_flags;
    for (int i = 0; i < SRE_MAX_DP; i++) k.base_params[i] = (R)c.base_params[i];
    for (int i = 0; i < SRE_MAX_DO; i++) {
        k.noise_std[i] = (R)c.noise_std[i];
        k.noise_mean[i] = (R)c.noise_mean[i];
    }
    for (int i = 0; i < SRE_MAX_DS; i++) {
        k.init_box_lo[i] = c.init_box_lo[i];
        k.init_box_up[i] = c.init_box_up[i];
    }
    const sre_step_wrappers &w = c.wrap;
    k.wrap.delay = w.act_delay;
    k.wrap.delay_inner = w.act_delay_inner;
    k.wrap.delay_first = w.delay_first;
    k.wrap.down = w.downsample > 1 ? w.downsample : 0;
    k.wrap.keep_mask = w.obs_keep_mask;
    k.wrap.vf_n = w.vf_n;
    k.wrap.n_aug = w.n_aug;
    k.wrap.aug_fixed = w.aug_fixed;
    for (int i = 0; i < SRE_MAX_VF; i++) {
        k.wrap.vf_pos[i] = w.vf_pos[i];
        k.wrap.vf_vel[i] = w.vf_vel[i];
    }
    k.wrap.vf_b0 = (R)w.vf_b0; k.wrap.vf_b1 = (R)w.vf_b1; k.wrap.vf_a1 = (R)w.vf_a1; k.wrap.vf_zi = (R)w.vf_zi;
    for (int i = 0; i < SRE_MAX_AUG; i++) {
        k.wrap.aug_param[i] = w.aug_param[i];
        k.wrap.aug_nominal[i] = (R)w.aug_nominal[i];
    }
    k.wrap.active = (w.act_delay > 0 || w.downsample > 1 || w.obs_keep_mask != 0 || w.vf_n > 0 || w.n_aug > 0) ? 1 : 0;
}

int check_cfg(const sre_env_cfg *cfg);
int reject_step_wrappers(const sre_env_cfg *cfg, const char *what);
int make_kdistr(const sre_env_cfg *cfg, const sre_distr *distr, int n_distr, KDistr &kd);
int make_kpolicy(const sre_policy *pol, int n_obs, int n_act, KPolicy &kp);

// per-env translation units (sre_rollout_<env>.cu)
#define SRE_DECL_ROLLOUT(KIND)                                                                                       \
    int rollout_f32_##KIND(const KCfg<float> &, const KDistr &, KPolicy, const KIO &, cudaStream_t);                  \
    int rollout_f64_##KIND(const KCfg<double> &, const KDistr &, KPolicy, const KIO &, cudaStream_t);
SRE_DECL_ROLLOUT(0)
SRE_DECL_ROLLOUT(1)
SRE_DECL_ROLLOUT(2)
SRE_DECL_ROLLOUT(3)
SRE_DECL_ROLLOUT(4)
SRE_DECL_ROLLOUT(5)

}  // namespace sre
