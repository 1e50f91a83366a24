// sre_api.cu -- the extern "C" surface declared in include/sre.h: argument checks, struct conversion, dispatch.
#include <cstdarg>

#include "sre_host.h"

namespace sre {

static thread_local char g_err[512] = "";
thread_local int g_launch_count = 0;

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

int check_cfg(const sre_env_cfg *cfg) {
    SRE_CHECK_ARG(cfg != nullptr, SRE_E_ARG, "cfg is NULL");
    SRE_CHECK_ARG(cfg->kind >= 0 && cfg->kind < SRE_N_KINDS, SRE_E_KIND, "unknown env kind %d", cfg->kind);
    SRE_CHECK_ARG(cfg->dtype == SRE_F32 || cfg->dtype == SRE_F64, SRE_E_DTYPE, "unknown dtype %d", cfg->dtype);
    // Env.__init__ (environments/base.py:55-60): dt must be a positive number
    SRE_CHECK_ARG(cfg->dt > 0.0 && std::isfinite(cfg->dt), SRE_E_ARG, "dt must be > 0, got %g", cfg->dt);
    SRE_CHECK_ARG(cfg->task.rew_kind >= 0 && cfg->task.rew_kind <= 2, SRE_E_ARG, "unknown reward kind %d",
                  cfg->task.rew_kind);
    SRE_CHECK_ARG(cfg->env_act_noise >= SRE_ENV_ACT_NOISE_NONE && cfg->env_act_noise <= SRE_ENV_ACT_NOISE_INNER, SRE_E_ARG,
                  "unknown env_act_noise %d", cfg->env_act_noise);
    if (cfg->obs_norm)  // observation_normalization.py:69-80: every (overridden) bound must be finite
        for (int i = 0; i < k_do[cfg->kind]; i++)
            SRE_CHECK_ARG(std::isfinite(cfg->obs_lb[i]) && std::isfinite(cfg->obs_ub[i]) && cfg->obs_ub[i] > cfg->obs_lb[i],
                          SRE_E_ARG, "obs_norm bounds of observation %d are [%g, %g]", i, cfg->obs_lb[i], cfg->obs_ub[i]);
    const sre_step_wrappers &w = cfg->wrap;
    SRE_CHECK_ARG(w.act_delay >= 0 && w.act_delay <= SRE_MAX_DELAY, SRE_E_ARG, "act_delay %d outside [0, %d]", w.act_delay,
                  SRE_MAX_DELAY);
    SRE_CHECK_ARG(w.downsample >= 0 && w.downsample <= SRE_MAX_DOWN, SRE_E_ARG, "downsample %d outside [0, %d]",
                  w.downsample, SRE_MAX_DOWN);
    SRE_CHECK_ARG(w.vf_n >= 0 && w.vf_n <= SRE_MAX_VF && w.n_aug >= 0 && w.n_aug <= SRE_MAX_AUG, SRE_E_ARG,
                  "vf_n %d / n_aug %d out of range", w.vf_n, w.n_aug);
    SRE_CHECK_ARG((w.obs_keep_mask >> k_do[cfg->kind]) == 0 && w.obs_keep_mask >= 0, SRE_E_ARG,
                  "obs_keep_mask 0x%x has bits beyond the %d observations", w.obs_keep_mask, k_do[cfg->kind]);
    for (int i = 0; i < w.vf_n; i++)
        SRE_CHECK_ARG(w.vf_pos[i] >= 0 && w.vf_pos[i] < k_ds[cfg->kind] && w.vf_vel[i] >= 0 && w.vf_vel[i] < k_do[cfg->kind],
                      SRE_E_ARG, "velocity filter pair %d: state %d / observation %d", i, w.vf_pos[i], w.vf_vel[i]);
    for (int i = 0; i < w.n_aug; i++)
        SRE_CHECK_ARG(w.aug_param[i] >= 0 && w.aug_param[i] < k_dp[cfg->kind], SRE_E_ARG, "augmented parameter %d: index %d",
                      i, w.aug_param[i]);
    SRE_CHECK_ARG(!(cfg->obs_noise || cfg->obs_norm) || !(w.downsample > 1 || w.obs_keep_mask || w.vf_n || w.n_aug),
                  SRE_E_UNSUPPORTED, "observation noise / normalisation cannot be combined with the fused observation wrappers");
    return 0;
}

// the step-granular entry points do not carry wrapper state between launches
int reject_step_wrappers(const sre_env_cfg *cfg, const char *what) {
    const sre_step_wrappers &w = cfg->wrap;
    SRE_CHECK_ARG(!(w.act_delay > 0 || w.downsample > 1 || w.obs_keep_mask != 0 || w.vf_n > 0 || w.n_aug > 0),
                  SRE_E_UNSUPPORTED, "%s: cfg.wrap is only implemented by sre_rollout", what);
    return 0;
}

int make_kdistr(const sre_env_cfg *cfg, const sre_distr *distr, int n_distr, KDistr &kd) {
    memset(&kd, 0, sizeof kd);
    SRE_CHECK_ARG(n_distr >= 0 && n_distr <= SRE_MAX_DISTR, SRE_E_ARG, "n_distr %d outside [0, %d]", n_distr,
                  SRE_MAX_DISTR);
    SRE_CHECK_ARG(n_distr == 0 || distr != nullptr, SRE_E_ARG, "distr is NULL");
    kd.n = n_distr;
    for (int j = 0; j < SRE_MAX_DP; j++) kd.entry_of_param[j] = -1;
    for (int i = 0; i < n_distr; i++) {
        SRE_CHECK_ARG(distr[i].param_index >= 0 && distr[i].param_index < k_dp[cfg->kind], SRE_E_ARG,
                      "distr[%d].param_index %d outside [0, %d)", i, distr[i].param_index, k_dp[cfg->kind]);
        SRE_CHECK_ARG(distr[i].kind >= 0 && distr[i].kind <= 2, SRE_E_ARG, "distr[%d].kind %d unknown", i, distr[i].kind);
        kd.e[i] = distr[i];
        kd.entry_of_param[distr[i].param_index] = (signed char)i;  // later entries win, like the sequential loop
    }
    return 0;
}

int make_kpolicy(const sre_policy *pol, int n_obs, int n_act, KPolicy &kp) {
    memset(&kp, 0, sizeof kp);
    SRE_CHECK_ARG(pol != nullptr, SRE_E_POLICY, "policy is NULL");
    kp.kind = pol->kind;
    SRE_CHECK_ARG(pol->act_noise >= SRE_ACT_NOISE_NONE && pol->act_noise <= SRE_ACT_NOISE_UNIFORM, SRE_E_POLICY,
                  "unknown act_noise %d", pol->act_noise);
    SRE_CHECK_ARG(pol->act_noise == SRE_ACT_NOISE_NONE || pol->kind == SRE_POL_FNN || pol->kind == SRE_POL_LINEAR,
                  SRE_E_UNSUPPORTED, "exploration noise needs an FNN or Linear policy");
    kp.act_noise = pol->act_noise;
    for (int i = 0; i < SRE_MAX_DA; i++) {
        if (pol->act_noise != SRE_ACT_NOISE_NONE && i < n_act)
            SRE_CHECK_ARG(pol->act_noise_scale[i] > 0.f, SRE_E_POLICY, "act_noise_scale[%d] = %g must be > 0", i,
                          (double)pol->act_noise_scale[i]);
        kp.noise_scale[i] = pol->act_noise_scale[i];
        kp.noise_mean[i] = pol->act_noise_mean[i];
    }
    if (pol->kind == SRE_POL_ACTIONS || pol->kind == SRE_POL_IDLE) return 0;
    SRE_CHECK_ARG(pol->weights != nullptr, SRE_E_POLICY, "policy weights are NULL");
    long long need = 0;
    if (pol->kind == SRE_POL_FNN) {
        SRE_CHECK_ARG(pol->n_layers >= 1 && pol->n_layers <= SRE_MAX_LAYERS, SRE_E_POLICY, "n_layers %d outside [1, %d]",
                      pol->n_layers, SRE_MAX_LAYERS);
        for (int i = 0; i <= pol->n_layers; i++) {
            SRE_CHECK_ARG(pol->sizes[i] >= 1 && pol->sizes[i] <= SRE_MAX_WIDTH, SRE_E_POLICY,
                          "layer size %d outside [1, %d]", pol->sizes[i], SRE_MAX_WIDTH);
            kp.sizes[i] = pol->sizes[i];
        }
        const int n_in = pol->cartpole_feats ? 5 : n_obs;
        SRE_CHECK_ARG(pol->sizes[0] == n_in, SRE_E_POLICY, "FNN input size %d != %d", pol->sizes[0], n_in);
        SRE_CHECK_ARG(pol->sizes[pol->n_layers] == n_act, SRE_E_POLICY, "FNN output size %d != act dim %d",
                      pol->sizes[pol->n_layers], n_act);
        for (int i = 0; i < pol->n_layers; i++) need += (long long)pol->sizes[i] * pol->sizes[i + 1] + pol->sizes[i + 1];
    } else if (pol->kind == SRE_POL_LINEAR) {
        SRE_CHECK_ARG(pol->n_feats >= 1 && pol->n_feats <= SRE_MAX_FEATS, SRE_E_POLICY, "n_feats %d outside [1, %d]",
                      pol->n_feats, SRE_MAX_FEATS);
        int nf = 0;
        SRE_CHECK_ARG(pol->n_feat_params >= 0 && pol->n_feat_params <= SRE_MAX_FEAT_PARAMS &&
                          (pol->n_feat_params == 0 || pol->feat_params != nullptr),
                      SRE_E_POLICY, "n_feat_params %d outside [0, %d] or feat_params NULL", pol->n_feat_params,
                      SRE_MAX_FEAT_PARAMS);
        for (int i = 0; i < pol->n_feat_params; i++) kp.fparams[i] = pol->feat_params[i];
        for (int i = 0; i < pol->n_feats; i++) {
            const int fk = pol->feats[i];
            SRE_CHECK_ARG(fk >= 0 && fk <= SRE_FEAT_RBF, SRE_E_POLICY, "unknown feature %d", fk);
            kp.feats[i] = fk;
            int dim = fk == SRE_FEAT_CONST ? 1 : n_obs;
            if (fk >= SRE_FEAT_SIG) {  // parametrised: the block has to lie inside feat_params and index the observation
                const int a = pol->feat_arg[i], left = pol->n_feat_params - a;
                SRE_CHECK_ARG(a >= 0 && left >= 1, SRE_E_POLICY, "feature %d: parameter block at %d outside the %d floats", i,
                              a, pol->n_feat_params);
                const float *p = pol->feat_params + a;
                if (fk == SRE_FEAT_MULT) {
                    const int n = (int)p[0];
                    SRE_CHECK_ARG(n >= 2 && n <= 8 && left >= 1 + n, SRE_E_POLICY, "MultFeat with %d indices", n);
                    for (int j = 0; j < n; j++)
                        SRE_CHECK_ARG((int)p[1 + j] >= 0 && (int)p[1 + j] < n_obs, SRE_E_POLICY, "MultFeat index %d", (int)p[1 + j]);
                    dim = 1;
                } else if (fk == SRE_FEAT_ATAN2) {
                    SRE_CHECK_ARG(left >= 2 && (int)p[0] >= 0 && (int)p[0] < n_obs && (int)p[1] >= 0 && (int)p[1] < n_obs,
                                  SRE_E_POLICY, "ATan2Feat indices out of range");
                    dim = 1;
                } else if (fk == SRE_FEAT_RF) {
                    const int D = (int)p[0];
                    SRE_CHECK_ARG(D >= 1 && left >= 2 + D * n_obs + D, SRE_E_POLICY, "RFFeat with %d features does not fit", D);
                    dim = D;
                } else if (fk == SRE_FEAT_RBF) {
                    const int F = (int)p[0];
                    SRE_CHECK_ARG(F >= 2 && left >= 2 + n_obs + F * n_obs, SRE_E_POLICY, "RBFFeat with %d centers does not fit", F);
                    dim = F * n_obs;
                }
                kp.feat_arg[i] = a;
            }
            nf += dim;
        }
        kp.n_lin_feats = nf;
        need = (long long)nf * n_act;
    } else {
        set_error("unknown policy kind %d", pol->kind);
        return SRE_E_POLICY;
    }
    SRE_CHECK_ARG(pol->n_weights == need, SRE_E_POLICY, "policy has %lld weights, layout needs %lld",
                  (long long)pol->n_weights, need);
    SRE_CHECK_ARG(pol->group_size >= 0 && (pol->group_size == 0 || pol->n_groups >= 1), SRE_E_POLICY,
                  "group_size %lld / n_groups %lld invalid", (long long)pol->group_size, (long long)pol->n_groups);
    kp.group_size = pol->group_size;
    kp.cta_groups = 0;  // decided per launch (launch_rollout)
    kp.n_layers = pol->n_layers;
    kp.n_feats = pol->n_feats;
    kp.hidden_nonlin = pol->hidden_nonlin;
    kp.output_nonlin = pol->output_nonlin;
    kp.cartpole_feats = pol->cartpole_feats;
    SRE_CHECK_ARG(pol->precision >= 0 && pol->precision <= SRE_PREC_TF32, SRE_E_POLICY, "unknown precision %d", pol->precision);
    kp.precision = pol->precision;
    kp.weights = pol->weights;
    kp.n_weights = pol->n_weights;
    return 0;
}

// defined in sre_step.cu
int derive_dispatch(const sre_env_cfg *, const void *, long long, void *, void *, void *, void *, void *, cudaStream_t);
int sample_params_dispatch(const sre_env_cfg *, const KDistr &, unsigned long long, unsigned long long, unsigned, void *,
                           long long, cudaStream_t);
int sample_init_dispatch(const sre_env_cfg *, const void *, unsigned long long, unsigned long long, unsigned, void *,
                         long long, cudaStream_t);
int reset_dispatch(const sre_env_cfg *, const void *, const void *, int, const unsigned char *, void *, void *, int *,
                   void *, unsigned long long, unsigned long long, unsigned, long long, cudaStream_t);
int step_dispatch(const sre_env_cfg *, const void *, const void *, void *, void *, int *, void *, void *,
                  unsigned char *, int *, unsigned long long, unsigned long long, unsigned, long long, cudaStream_t);
int observe_dispatch(const sre_env_cfg *, const void *, void *, long long, cudaStream_t);
int step_rew_dispatch(const sre_env_cfg *, const void *, const void *, const void *, void *, long long, cudaStream_t);
int policy_forward_dispatch(const KPolicy &, const float *, int, float *, int, long long, cudaStream_t);
int task_rew_dispatch(const sre_task *, int, int, int, double, const void *, const void *, void *, long long, cudaStream_t);

static int cuda_ret(int e, const char *what) {
    if (e != 0) set_error("%s: %s", what, cudaGetErrorString((cudaError_t)e));
    return e;
}

}  // namespace sre

using namespace sre;

extern "C" {

int sre_abi_version(void) { return SRE_ABI_VERSION; }
const char *sre_last_error(void) { return g_err; }
int sre_last_launch_count(void) { return g_launch_count; }

int sre_device_count(int *count) {
    SRE_CHECK_ARG(count != nullptr, SRE_E_ARG, "count is NULL");
    SRE_CUDA(cudaGetDeviceCount(count));
    return 0;
}

int sre_env_dims(int kind, int *ds, int *dobs, int *da, int *dp, int *nh, int *dinit) {
    SRE_CHECK_ARG(kind >= 0 && kind < SRE_N_KINDS, SRE_E_KIND, "unknown env kind %d", kind);
    if (ds) *ds = k_ds[kind];
    if (dobs) *dobs = k_do[kind];
    if (da) *da = k_da[kind];
    if (dp) *dp = k_dp[kind];
    if (nh) *nh = k_nh[kind];
    if (dinit) *dinit = k_dinit[kind];
    return 0;
}

int sre_nominal_params(int kind, int long_pole, double *dp) {
    SRE_CHECK_ARG(kind >= 0 && kind < SRE_N_KINDS, SRE_E_KIND, "unknown env kind %d", kind);
    SRE_CHECK_ARG(dp != nullptr, SRE_E_ARG, "dp is NULL");
    for (int i = 0; i < SRE_MAX_DP; i++) dp[i] = 0.0;
    switch (kind) {
    case SRE_QQ: {  // quanser_qube.py:54-68
        const double v[11] = {9.81, 8.4, 0.042, 0.095, 0.085, 5e-6, 0.024, 0.129, 1e-6, 0.0, 0.0};
        memcpy(dp, v, sizeof v);
    } break;
    case SRE_QCP: {  // quanser_cartpole.py:111-143 (long / short pole)
        const double m_pole = long_pole ? 0.23 : 0.127, l_pole = long_pole ? 0.641 / 2 : 0.3365 / 2;
        const double v[17] = {9.81, 0.58, 0.814, 0.9, 0.9, 3.71, 3.9e-7, 6.35e-3, 2.6, 7.67e-3, 0.0024, 5.4,
                              m_pole, l_pole, 0.02, 0.0, 0.0};
        memcpy(dp, v, sizeof v);
    } break;
    case SRE_QBB: {  // quanser_ball_balancer.py:169-202, default thresholds :135-138
        const double v[20] = {9.81, 0.003, 0.019625, 0.275, 0.0254, 70.0, 0.9, 5.2822e-5, 4.6063e-7, 0.0077,
                              2.6,  0.69,  0.015,    0.05,  0.28,   -0.10, 0.28, -0.074,   0.0,       0.0};
        memcpy(dp, v, sizeof v);
    } break;
    case SRE_BOB: {  // ball_on_beam.py:75-87
        const double v[8] = {9.81, 0.5, 0.1, 3.0, 2.0, 0.1, 0.05, 0.0};
        memcpy(dp, v, sizeof v);
    } break;
    case SRE_OMO: {  // one_mass_oscillator.py:81-87
        const double v[3] = {1.0, 30.0, 0.5};
        memcpy(dp, v, sizeof v);
    } break;
    default: {  // pendulum.py:96-104
        const double v[5] = {9.81, 1.0, 1.0, 0.05, 3.5};
        memcpy(dp, v, sizeof v);
    } break;
    }
    return 0;
}

// _create_task defaults: quanser_qube.py:183-188, quanser_cartpole.py:573-588, quanser_ball_balancer.py:121-131,
// ball_on_beam.py:100-108, one_mass_oscillator.py:68-79
int sre_default_task(int kind, sre_task *t) {
    SRE_CHECK_ARG(kind >= 0 && kind < SRE_N_KINDS, SRE_E_KIND, "unknown env kind %d", kind);
    SRE_CHECK_ARG(t != nullptr, SRE_E_ARG, "task is NULL");
    memset(t, 0, sizeof *t);
    const int ds = k_ds[kind], da = k_da[kind];
    for (int i = 0; i < SRE_MAX_DS; i++) t->radial_mod[i] = 2 * SRE_PI;
    t->min_rew = 1e-4;
    double q[SRE_MAX_DS] = {0}, r[SRE_MAX_DA] = {0};
    switch (kind) {
    case SRE_QQ:
        t->state_des[1] = SRE_PI; q[0] = 1.0; q[1] = 1.0; q[2] = 2e-2; q[3] = 5e-3; r[0] = 4e-3;
        t->rew_kind = SRE_REW_EXP; t->radial_mask = 1u << 1;
        break;
    case SRE_QCP:
        t->state_des[1] = SRE_PI; q[0] = 3e-1; q[1] = 5e-1; q[2] = 5e-3; q[3] = 1e-3; r[0] = 1e-3;
        t->rew_kind = SRE_REW_EXP; t->radial_mask = 1u << 1;
        break;
    case SRE_QBB:
        q[0] = 1e0; q[1] = 1e0; q[2] = 5e3; q[3] = 5e3; q[4] = 1e-2; q[5] = 1e-2; q[6] = 5e-1; q[7] = 5e-1;
        r[0] = 1e-2; r[1] = 1e-2; t->rew_kind = SRE_REW_SCALED_EXP;
        break;
    case SRE_BOB:
        q[0] = 1e5; q[1] = 1e3; q[2] = 1e3; q[3] = 1e2; r[0] = 1.0; t->rew_kind = SRE_REW_SCALED_EXP;
        break;
    case SRE_OMO:
        q[0] = 1e1; q[1] = 1e-2; r[0] = 1e-6; t->rew_kind = SRE_REW_QUADR;
        t->final_flags = SRE_FR_ENABLED | SRE_FR_ALWAYS_NEG; t->final_factor = 1e3;
        break;
    default:  // pendulum.py:82-88 -- idcs=[1] modulates theta_dot, as the reference does
        t->state_des[0] = SRE_PI; q[0] = 1e-0; q[1] = 1e-3; r[0] = 1e-2;
        t->rew_kind = SRE_REW_EXP; t->radial_mask = 1u << 1;
        break;
    }
    for (int i = 0; i < ds; i++) t->Q[i * ds + i] = q[i];
    for (int i = 0; i < da; i++) t->R[i * da + i] = r[i];
    return 0;
}

int sre_record_width(int kind) {
    if (kind < 0 || kind >= SRE_N_KINDS) return SRE_E_KIND;
    return k_do[kind] + 2 * k_da[kind] + 1 + k_ds[kind] + k_nh[kind];
}

int sre_record_layout(int kind, int *off_obs, int *off_act, int *off_act_app, int *off_rew, int *off_state,
                      int *off_hidden) {
    SRE_CHECK_ARG(kind >= 0 && kind < SRE_N_KINDS, SRE_E_KIND, "unknown env kind %d", kind);
    int o = 0;
    if (off_obs) *off_obs = o;
    o += k_do[kind];
    if (off_act) *off_act = o;
    o += k_da[kind];
    if (off_act_app) *off_act_app = o;
    o += k_da[kind];
    if (off_rew) *off_rew = o;
    o += 1;
    if (off_state) *off_state = o;
    o += k_ds[kind];
    if (off_hidden) *off_hidden = o;
    return 0;
}

int sre_stats_init(double *stats, void *stream) {
    SRE_CHECK_ARG(stats != nullptr, SRE_E_ARG, "stats is NULL");
    double h[SRE_N_STATS] = {0, 0, 0, 0, 0, 0, 1e300, -1e300};
    SRE_CUDA(cudaMemcpyAsync(stats, h, sizeof h, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    SRE_CUDA(cudaStreamSynchronize((cudaStream_t)stream));  // h is a stack buffer
    return 0;
}

int sre_derive(const sre_env_cfg *cfg, const void *params, int64_t n, void *st_lo, void *st_up, void *act_lo,
               void *act_up, void *c_max, void *stream) {
    int e = check_cfg(cfg);
    if (e) return e;
    SRE_CHECK_ARG(n >= 0, SRE_E_ARG, "n < 0");
    if (n == 0) return 0;
    return cuda_ret(derive_dispatch(cfg, params, n, st_lo, st_up, act_lo, act_up, c_max, (cudaStream_t)stream),
                    "sre_derive");
}

int sre_sample_params(const sre_env_cfg *cfg, const sre_distr *distr, int n_distr, uint64_t seed,
                      uint64_t env_offset, uint32_t episode, void *params, int64_t n, void *stream) {
    int e = check_cfg(cfg);
    if (e) return e;
    SRE_CHECK_ARG(params != nullptr && n >= 0, SRE_E_ARG, "params is NULL or n < 0");
    KDistr kd;
    e = make_kdistr(cfg, distr, n_distr, kd);
    if (e) return e;
    if (n == 0 || n_distr == 0) return 0;
    return cuda_ret(sample_params_dispatch(cfg, kd, seed, env_offset, episode, params, n, (cudaStream_t)stream),
                    "sre_sample_params");
}

int sre_sample_init(const sre_env_cfg *cfg, const void *params, uint64_t seed, uint64_t env_offset,
                    uint32_t episode, void *init, int64_t n, void *stream) {
    int e = check_cfg(cfg);
    if (e) return e;
    SRE_CHECK_ARG(init != nullptr && n >= 0, SRE_E_ARG, "init is NULL or n < 0");
    if (n == 0) return 0;
    return cuda_ret(sample_init_dispatch(cfg, params, seed, env_offset, episode, init, n, (cudaStream_t)stream),
                    "sre_sample_init");
}

int sre_reset(const sre_env_cfg *cfg, const void *params, const void *init, int n_init_cols, const uint8_t *mask,
              void *state, void *hidden, int32_t *curr_step, void *obs, uint64_t seed, uint64_t env_offset,
              uint32_t episode, int64_t n, void *stream) {
    int e = check_cfg(cfg);
    if (e) return e;
    if ((e = reject_step_wrappers(cfg, "sre_reset"))) return e;
    SRE_CHECK_ARG(init && state && hidden && curr_step && n >= 0, SRE_E_ARG, "NULL pointer or n < 0");
    // base.py:184-193: full state or an element of the init space
    SRE_CHECK_ARG(n_init_cols == k_ds[cfg->kind] || n_init_cols == k_dinit[cfg->kind], SRE_E_ARG,
                  "init has %d columns, expected %d (state) or %d (init space)", n_init_cols, k_ds[cfg->kind],
                  k_dinit[cfg->kind]);
    if (n == 0) return 0;
    return cuda_ret(reset_dispatch(cfg, params, init, n_init_cols, mask, state, hidden, curr_step, obs, seed,
                                   env_offset, episode, n, (cudaStream_t)stream),
                    "sre_reset");
}

int sre_step(const sre_env_cfg *cfg, const void *params, const void *act, void *state, void *hidden,
             int32_t *curr_step, void *obs, void *rew, uint8_t *done, int32_t *status, uint64_t seed,
             uint64_t env_offset, uint32_t episode, int64_t n, void *stream) {
    int e = check_cfg(cfg);
    if (e) return e;
    if ((e = reject_step_wrappers(cfg, "sre_step"))) return e;
    SRE_CHECK_ARG(act && state && hidden && curr_step && obs && rew && done && n >= 0, SRE_E_ARG,
                  "NULL pointer or n < 0");
    if (n == 0) return 0;
    return cuda_ret(step_dispatch(cfg, params, act, state, hidden, curr_step, obs, rew, done, status, seed, env_offset,
                                  episode, n, (cudaStream_t)stream),
                    "sre_step");
}

int sre_observe(const sre_env_cfg *cfg, const void *state, void *obs, int64_t n, void *stream) {
    int e = check_cfg(cfg);
    if (e) return e;
    SRE_CHECK_ARG(state && obs && n >= 0, SRE_E_ARG, "NULL pointer or n < 0");
    if (n == 0) return 0;
    return cuda_ret(observe_dispatch(cfg, state, obs, n, (cudaStream_t)stream), "sre_observe");
}

int sre_step_rew(const sre_env_cfg *cfg, const void *params, const void *state, const void *act, void *rew, int64_t n,
                 void *stream) {
    int e = check_cfg(cfg);
    if (e) return e;
    SRE_CHECK_ARG(state && act && rew && n >= 0, SRE_E_ARG, "NULL pointer or n < 0");
    if (n == 0) return 0;
    return cuda_ret(step_rew_dispatch(cfg, params, state, act, rew, n, (cudaStream_t)stream), "sre_step_rew");
}

int sre_task_step_rew(const sre_task *task, int dtype, int ds, int da, double c_max, const void *state,
                      const void *act, void *rew, int64_t n, void *stream) {
    SRE_CHECK_ARG(task && state && act && rew && n >= 0, SRE_E_ARG, "NULL pointer or n < 0");
    SRE_CHECK_ARG(dtype == SRE_F32 || dtype == SRE_F64, SRE_E_DTYPE, "unknown dtype %d", dtype);
    SRE_CHECK_ARG(ds >= 1 && ds <= SRE_MAX_DS && da >= 1 && da <= SRE_MAX_DA, SRE_E_ARG, "ds %d / da %d out of range", ds, da);
    SRE_CHECK_ARG(task->rew_kind >= 0 && task->rew_kind <= 2, SRE_E_ARG, "unknown reward kind %d", task->rew_kind);
    if (n == 0) return 0;
    return cuda_ret(task_rew_dispatch(task, dtype, ds, da, c_max, state, act, rew, n, (cudaStream_t)stream),
                    "sre_task_step_rew");
}

int sre_policy_forward(const sre_policy *pol, const float *obs, int n_obs, float *act, int n_act, int64_t n,
                       void *stream) {
    SRE_CHECK_ARG(obs && act && n >= 0, SRE_E_ARG, "NULL pointer or n < 0");
    SRE_CHECK_ARG(n_obs >= 1 && n_obs <= SRE_MAX_WIDTH && n_act >= 1 && n_act <= SRE_MAX_WIDTH, SRE_E_ARG,
                  "n_obs / n_act outside [1, %d]", SRE_MAX_WIDTH);
    KPolicy kp;
    int e = make_kpolicy(pol, n_obs, n_act, kp);
    if (e) return e;
    if (n == 0) return 0;
    return cuda_ret(policy_forward_dispatch(kp, obs, n_obs, act, n_act, n, (cudaStream_t)stream), "sre_policy_forward");
}

int sre_rollout(const sre_env_cfg *cfg, const sre_policy *pol, const sre_rollout_io *io, void *stream) {
    g_launch_count = 0;
    int e = check_cfg(cfg);
    if (e) return e;
    SRE_CHECK_ARG(io != nullptr, SRE_E_ARG, "io is NULL");
    SRE_CHECK_ARG(io->n_envs >= 0 && io->n_rows >= 1, SRE_E_ARG, "n_envs %lld / n_rows %lld invalid",
                  (long long)io->n_envs, (long long)io->n_rows);
    const int kind = cfg->kind;
    KPolicy kp;
    // policy input width behind ObsPartialWrapper / StateAugmentationWrapper
    int n_pol_in = k_do[kind];
    if (cfg->wrap.obs_keep_mask) n_pol_in = __builtin_popcount((unsigned)cfg->wrap.obs_keep_mask);
    n_pol_in += cfg->wrap.n_aug;
    SRE_CHECK_ARG(!pol || pol->kind != SRE_POL_LINEAR || n_pol_in == k_do[kind], SRE_E_UNSUPPORTED,
                  "the fused linear policy needs the full observation (got %d of %d entries)", n_pol_in, k_do[kind]);
    e = make_kpolicy(pol, n_pol_in, k_da[kind], kp);
    if (e) return e;
    // small layers of the tensor-core policy as launch arguments (FnnTC2's layout: W1^T, b1, b2 prescaled, W3, b3)
    if (pol->weights_host && pol->kind == SRE_POL_FNN && pol->n_layers == 3 && pol->sizes[1] == 64 && pol->sizes[2] == 64 &&
        pol->group_size == 0 && pol->sizes[0] <= 8 && pol->sizes[3] <= 2 && pol->hidden_nonlin == 0 && pol->output_nonlin == 0) {
        const int IN = pol->sizes[0], OUT = pol->sizes[3], H = 64;
        const float PRE = SRE_TANH_PRESCALE;
        const float *W1 = pol->weights_host, *b1 = W1 + IN * H, *W2 = b1 + H, *b2 = W2 + H * H, *W3 = b2 + H, *b3 = W3 + OUT * H;
        float *c = kp.cw;
        for (int i = 0; i < IN * H; i++) c[(i % IN) * H + i / IN] = PRE * W1[i];
        for (int i = 0; i < H; i++) { c[IN * H + i] = PRE * b1[i]; c[IN * H + H + i] = PRE * b2[i]; }
        for (int i = 0; i < OUT * H; i++) c[IN * H + 2 * H + i] = W3[i];
        for (int i = 0; i < 4; i++) c[IN * H + 2 * H + OUT * H + i] = i < OUT ? b3[i] : 0.f;
        kp.has_cw = 1;
    }
    if (pol->weights_host && pol->kind == SRE_POL_FNN && pol->n_layers == 3 && pol->sizes[1] == 16 && pol->sizes[2] == 16 &&
        pol->group_size == 0 && pol->sizes[0] <= 8 && pol->sizes[3] <= 2 && pol->hidden_nonlin == 0 && pol->output_nonlin == 0) {
        // FnnLane16's layout (sre_policy_tc.cuh): W1^T [IN][16], b1, W2^T [16][16], b2 (prescaled), W3^T [16][2 OP], b3
        const int IN = pol->sizes[0], OUT = pol->sizes[3], H = 16, OP = (OUT + 1) / 2;
        const int oB1 = IN * H, oW2 = oB1 + H, oB2 = oW2 + H * H, oW3 = oB2 + H, oB3 = oW3 + H * 2 * OP;
        const float PRE = SRE_TANH_PRESCALE;
        const float *W1 = pol->weights_host, *b1 = W1 + IN * H, *W2 = b1 + H, *b2 = W2 + H * H, *W3 = b2 + H, *b3 = W3 + OUT * H;
        float *c = kp.cw;
        for (int i = 0; i < IN * H; i++) c[(i % IN) * H + i / IN] = PRE * W1[i];
        for (int i = 0; i < H * H; i++) c[oW2 + (i % H) * H + i / H] = PRE * W2[i];
        for (int i = 0; i < H; i++) { c[oB1 + i] = PRE * b1[i]; c[oB2 + i] = PRE * b2[i]; }
        for (int i = 0; i < H * 2 * OP; i++) c[oW3 + i] = (i % (2 * OP)) < OUT ? W3[(i % (2 * OP)) * H + i / (2 * OP)] : 0.f;
        for (int i = 0; i < 2 * OP; i++) c[oB3 + i] = i < OUT ? b3[i] : 0.f;
        kp.has_cw = 2;
    }
    SRE_CHECK_ARG(!kp.cartpole_feats, SRE_E_UNSUPPORTED,
                  "cartpole_feats (the fork's 4-dim FNNPolicy input) is only available in sre_policy_forward");
    if (kp.group_size > 0)
        SRE_CHECK_ARG((io->n_envs + kp.group_size - 1) / kp.group_size <= pol->n_groups, SRE_E_POLICY,
                      "%lld lanes in groups of %lld need more than the %lld parameter sets given", (long long)io->n_envs,
                      (long long)kp.group_size, (long long)pol->n_groups);
    KDistr kd;
    e = make_kdistr(cfg, io->distr, io->n_distr, kd);
    if (e) return e;
    if (io->init)
        SRE_CHECK_ARG(io->n_init_cols == k_ds[kind] || io->n_init_cols == k_dinit[kind], SRE_E_ARG,
                      "init has %d columns, expected %d or %d", io->n_init_cols, k_ds[kind], k_dinit[kind]);
    SRE_CHECK_ARG(io->traj_env_stride == 0 ||
                      (io->traj_env_offset >= 0 && io->traj_env_offset + io->n_envs <= io->traj_env_stride),
                  SRE_E_ARG, "lanes [%lld, %lld) do not fit a record buffer of %lld columns", (long long)io->traj_env_offset,
                  (long long)(io->traj_env_offset + io->n_envs), (long long)io->traj_env_stride);
    if (io->n_envs == 0) return 0;
    KIO k;
    memset(&k, 0, sizeof k);
    k.n_envs = io->n_envs; k.n_rows = io->n_rows; k.seed = io->seed; k.env_offset = io->env_offset;
    k.episode = io->episode; k.stop_on_done = io->stop_on_done; k.auto_reset = io->auto_reset;
    k.params = io->params; k.init = io->init; k.actions = io->actions; k.n_init_cols = io->n_init_cols;
    k.traj = io->traj; k.row_flag = io->row_flag; k.ret = io->ret; k.length = io->length;
    k.traj_stride = io->traj_env_stride > 0 ? io->traj_env_stride : io->n_envs;
    k.traj_off = io->traj_env_stride > 0 ? io->traj_env_offset : 0;
    k.params_out = io->params_out; k.stats = io->stats; k.status = io->status;
    k.idle_policy = (pol->kind == SRE_POL_IDLE);
    cudaStream_t s = (cudaStream_t)stream;
    int r;
    if (cfg->dtype == SRE_F64) {
        KCfg<double> kc;
        make_kcfg<double>(*cfg, kc);
        switch (kind) {
        case SRE_QQ: r = rollout_f64_0(kc, kd, kp, k, s); break;
        case SRE_QCP: r = rollout_f64_1(kc, kd, kp, k, s); break;
        case SRE_QBB: r = rollout_f64_2(kc, kd, kp, k, s); break;
        case SRE_BOB: r = rollout_f64_3(kc, kd, kp, k, s); break;
        case SRE_OMO: r = rollout_f64_4(kc, kd, kp, k, s); break;
        default: r = rollout_f64_5(kc, kd, kp, k, s); break;
        }
    } else {
        KCfg<float> kc;
        make_kcfg<float>(*cfg, kc);
        switch (kind) {
        case SRE_QQ: r = rollout_f32_0(kc, kd, kp, k, s); break;
        case SRE_QCP: r = rollout_f32_1(kc, kd, kp, k, s); break;
        case SRE_QBB: r = rollout_f32_2(kc, kd, kp, k, s); break;
        case SRE_BOB: r = rollout_f32_3(kc, kd, kp, k, s); break;
        case SRE_OMO: r = rollout_f32_4(kc, kd, kp, k, s); break;
        default: r = rollout_f32_5(kc, kd, kp, k, s); break;
        }
    }
    if (r == 0) g_launch_count = 1;
    return cuda_ret(r, "sre_rollout");
}

}  // extern "C"
