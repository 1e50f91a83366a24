/*
 * sre.h -- C-ABI of the B200-native batched rollout engine for Pyrado's pysim environments.
 *
 * The reference (swami1995/SimuRLacra) has NO FFI on this path: pysim envs are plain Python classes
 * and the boundary is the duck-typed Env API.  This header is therefore the boundary a maintainer would
 * bind (ctypes) UNDER Python classes that keep the reference signatures.  Each entry
 * point cites the reference interface it replaces; paths are relative to /root/reference/Pyrado/pyrado/ .
 *
 * Conventions
 *   - every function returns int: 0 = ok, < 0 = argument error (SRE_E_*), > 0 = CUDA error code;
 *     sre_last_error() gives a thread-local message for the last non-zero return
 *   - "real" arrays are float (cfg.dtype == SRE_F32) or double (SRE_F64); policy weights are always float
 *   - all data pointers are DEVICE pointers owned by the caller unless the parameter says "host"
 *   - `stream` is a cudaStream_t passed as void* (NULL = the default stream); calls are asynchronous
 *   - nothing crosses the ABI but plain pointers, sizes and the POD structs below; no torch types
 *   - there is no CPU fallback: without a CUDA device every compute call returns a CUDA error
 */
#ifndef SRE_H
#define SRE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library is built with -fvisibility=hidden; only this header is exported */
#endif

#define SRE_ABI_VERSION 1

/* environment kinds (environments/pysim/*.py) */
#define SRE_QQ 0  /* QQubeSwingUpSim       quanser_qube.py:152 */
#define SRE_QCP 1 /* QCartPoleSwingUpSim   quanser_cartpole.py:507 */
#define SRE_QBB 2 /* QBallBalancerSim      quanser_ball_balancer.py:50 */
#define SRE_BOB 3 /* BallOnBeamSim         ball_on_beam.py:41 */
#define SRE_OMO 4 /* OneMassOscillatorSim  one_mass_oscillator.py:43 */
#define SRE_PEND 5 /* PendulumSim           pendulum.py:43 */
#define SRE_N_KINDS 6

#define SRE_F32 0
#define SRE_F64 1

#define SRE_MAX_DS 8
#define SRE_MAX_DO 8
#define SRE_MAX_DA 2
#define SRE_MAX_DP 20
#define SRE_MAX_NH 2
#define SRE_MAX_DISTR 32
#define SRE_MAX_LAYERS 8
#define SRE_MAX_WIDTH 128
#define SRE_MAX_FEATS 12

/* cfg.flags */
#define SRE_F_SIMPLE_DYN 1 /* simple_dynamics=True (quanser_cartpole.py:515, quanser_ball_balancer.py:64) */
#define SRE_F_QCP_STAB 2   /* QCartPoleStabSim state space: theta in pi +- 15 deg, |theta_dot| <= 2 pi (quanser_cartpole.py:473-482) */
#define SRE_F_DISC_ACT 4   /* BallOnBeamDiscSim: actions are projected onto the 3-element DiscreteSpace linspace(lo, up, 3) (ball_on_beam.py:155-162, spaces/discrete.py:119-127) */

/* reward kinds: tasks/reward_functions.py:176 (Quadr), :224 (ExpQuadr), :247 (ScaledExpQuadr) */
#define SRE_REW_QUADR 0
#define SRE_REW_EXP 1
#define SRE_REW_SCALED_EXP 2

/* FinalRewMode bits: tasks/final_reward.py:41-48 */
#define SRE_FR_ENABLED 1
#define SRE_FR_STATE_DEP 2
#define SRE_FR_TIME_DEP 4
#define SRE_FR_ALWAYS_POS 8
#define SRE_FR_ALWAYS_NEG 16

/* policy kinds */
#define SRE_POL_ACTIONS 0 /* injected actions [H, N, da] */
#define SRE_POL_FNN 1     /* policies/feed_back/fnn.py:43 (FNN), :163 (FNNPolicy) */
#define SRE_POL_LINEAR 2  /* policies/feed_back/linear.py:41 + policies/features.py:42 */
#define SRE_POL_IDLE 3    /* policies/feed_forward/dummy.py (IdlePolicy): zero action */

/* sre_policy.precision */
#define SRE_ENV_ACT_NOISE_NONE 0
#define SRE_ENV_ACT_NOISE_OUTER 1
#define SRE_ENV_ACT_NOISE_INNER 2
#define SRE_ACT_NOISE_NONE 0
#define SRE_ACT_NOISE_NORMAL 1
#define SRE_ACT_NOISE_UNIFORM 2
#define SRE_PREC_FP32 0   /* FP32 FMA on the CUDA cores (bias first, inputs ascending) */
#define SRE_PREC_TF32X3 1 /* tcgen05 tensor cores, 3 x TF32 split (hi*hi + hi*lo + lo*hi): fp32-level accuracy */
#define SRE_PREC_TF32 2   /* tcgen05 tensor cores, single TF32 pass (about 1e-3 relative) */

/* feature functions of the Linear policy, policies/features.py:127-175 */
#define SRE_FEAT_CONST 0
#define SRE_FEAT_IDENTITY 1
#define SRE_FEAT_SIGN 2
#define SRE_FEAT_ABS 3
#define SRE_FEAT_SQUARED 4
#define SRE_FEAT_CUBIC 5
#define SRE_FEAT_SIN 6
#define SRE_FEAT_COS 7
#define SRE_FEAT_SINSIN 8
#define SRE_FEAT_SINCOS 9
/* parametrised features (policies/features.py:157-161, :181-445); their parameters sit in sre_policy.feat_params at
 * feat_arg[k] (floats):
 *   SIG   sigmoid(scale x)            [scale]                         n_obs features
 *   BELL  exp(-scale x^2)             [scale]                         n_obs features
 *   MULT  prod_i x[idx_i]             [n, idx_0 .. idx_{n-1}]         1 feature          (MultFeat)
 *   ATAN2 atan2(x[i_sin], x[i_cos])   [i_sin, i_cos]                  1 feature          (ATan2Feat)
 *   RF    scale cos(freq x + shift)   [D, scale, freq[D][n_obs], shift[D]]   D features  (RFFeat)
 *   RBF   exp(-scale_i (x_i - c_fi)^2), L1-normalised per input dimension (state_wise = 1) or over all of them (0),
 *         feature index i F + f       [F, state_wise, scale[n_obs], centers[F][n_obs]]   F n_obs features (RBFFeat) */
#define SRE_FEAT_SIG 10
#define SRE_FEAT_BELL 11
#define SRE_FEAT_MULT 12
#define SRE_FEAT_ATAN2 13
#define SRE_FEAT_RF 14
#define SRE_FEAT_RBF 15
#define SRE_MAX_FEAT_PARAMS 1024

/* distribution kinds, domain_randomization/domain_parameter.py:135 (Uniform), :171 (Normal), :246 (Bernoulli) */
#define SRE_DISTR_NORMAL 0    /* a = mean, b = std */
#define SRE_DISTR_UNIFORM 1   /* a = lo,   b = hi  */
#define SRE_DISTR_BERNOULLI 2 /* a = val_0, b = val_1, c = prob_1 */

/* row flags written next to the trajectory records */
#define SRE_ROW_STEP 0      /* a step was taken from this row, episode continues */
#define SRE_ROW_DONE 1      /* a step was taken from this row and it ended the episode (done == True) */
#define SRE_ROW_TERMINAL 2  /* closing row of a finished episode: obs/state/hidden only */
#define SRE_ROW_TRUNCATED 3 /* closing row of an episode cut by the row budget (not done) */
#define SRE_ROW_IDLE 4      /* lane idle (single-episode mode, after the closing row): record is zero */

/* per-env status bits (instead of the exceptions the reference raises mid-flight) */
#define SRE_ST_NAN_STATE 1 /* spaces/box.py:142 would raise ValueErr */
#define SRE_ST_NAN_ACT 2   /* sampling/rollout.py:223 would raise ValueErr */
#define SRE_ST_NAN_OBS 4   /* sampling/rollout.py:193 would raise ValueErr */

/* curr_step words of the step-granular API: the step count sits in the low 30 bits; bit 30 says that the episode's
 * FinalRewTask has already yielded its final reward (tasks/final_reward.py:131-135: once per episode).  sre_reset
 * clears the word, sre_step maintains both parts. */
#define SRE_STEP_YIELDED_FINAL (1 << 30)
#define SRE_STEP_COUNT_MASK 0x3FFFFFFF

/* error codes */
#define SRE_E_ARG (-1)
#define SRE_E_KIND (-2)
#define SRE_E_DTYPE (-3)
#define SRE_E_POLICY (-4)
#define SRE_E_UNSUPPORTED (-5)

/* rollout statistics vector (doubles), the numbers the NCCL all-reduce carries:
 * sums first (all-reduce SUM), then min/max */
#define SRE_STAT_EPISODES 0
#define SRE_STAT_SUM_RET 1
#define SRE_STAT_SUM_RET2 2
#define SRE_STAT_SUM_LEN 3
#define SRE_STAT_STEPS 4
#define SRE_STAT_FAILED 5 /* episodes ended by leaving the state space */
#define SRE_STAT_MIN_RET 6
#define SRE_STAT_MAX_RET 7
#define SRE_N_STATS 8

/* Task description: tasks/desired_state.py:40 (DesStateTask), :116 (RadiallySymmDesStateTask),
 * tasks/final_reward.py:69 (FinalRewTask), reward functions above.  success_fcn is never_succeeded
 * (tasks/utils.py:32) as in all five pysim envs. */
typedef struct sre_task {
    double state_des[SRE_MAX_DS];
    double Q[SRE_MAX_DS * SRE_MAX_DS]; /* row-major [ds, ds] */
    double R[SRE_MAX_DA * SRE_MAX_DA]; /* row-major [da, da] */
    double radial_mod[SRE_MAX_DS];     /* modulation factor per state index (default 2*pi) */
    double min_rew;                    /* ScaledExpQuadrErrRewFcn.min_rew, reward_functions.py:257 */
    double final_factor;               /* FinalRewTask factor, final_reward.py:99 */
    int32_t rew_kind;
    uint32_t radial_mask;              /* bit i: state index i is modulated (desired_state.py:146) */
    int32_t final_flags;
    int32_t reserved;
} sre_task;

/* One batch of identical environments: ctor arguments of SimPyEnv (environments/pysim/base.py:47) plus the
 * wrappers that are fused into the kernels. */
/* Stateful / shape-changing step-path wrappers run INSIDE sre_rollout (one wrapper state per environment lane, kept
 * for the whole launch).  Order of the observation path, innermost first:
 *   observe -> ObsVelFiltWrapper -> DownsamplingWrapper (mean) -> ObsPartialWrapper -> StateAugmentationWrapper -> policy
 * and of the action path:  policy -> {Downsampling hold, ActDelay (outer)} -> ActNormWrapper -> ActDelay (inner) -> env.
 * The recorded observation is the policy's (first min(n_in, DO) entries); the recorded action stays the policy's.
 * sre_step / sre_reset reject a cfg with any of these set (SRE_E_UNSUPPORTED): between step launches the host applies them. */
#define SRE_MAX_DELAY 8
#define SRE_MAX_DOWN 8
#define SRE_MAX_VF 4
#define SRE_MAX_AUG 8
typedef struct sre_step_wrappers {
    int32_t act_delay;       /* ActDelayWrapper._process_act (environment_wrappers/action_delay.py:97-110): the action of
                              * `act_delay` steps ago reaches the env, the queue holds zeros after a reset; 0: off */
    int32_t act_delay_inner; /* 1: the wrapper sits inside ActNormWrapper (env units), 0: outside (policy units) */
    int32_t delay_first;     /* 1: ActDelay (outer) is applied before the Downsampling hold, 0: after it */
    int32_t downsample;      /* DownsamplingWrapper (downsampling.py:134-162): every factor-th action is held, the
                              * observation is the mean over the last `factor` observations; <= 1: off */
    int32_t obs_keep_mask;   /* ObsPartialWrapper (observation_partial.py:63-75): bit i = entry i is kept; 0: all */
    int32_t vf_n;            /* ObsVelFiltWrapper (observation_velfilter.py:107-145): number of (pos, vel) pairs */
    int32_t vf_pos[SRE_MAX_VF]; /* state index of the position */
    int32_t vf_vel[SRE_MAX_VF]; /* observation index that is replaced by the filtered velocity */
    double vf_b0, vf_b1, vf_a1, vf_zi; /* first-order filter y = b0 x + z, z' = b1 x - a1 y (a0 = 1), z(reset) = zi x */
    int32_t n_aug;           /* StateAugmentationWrapper (state_augmentation.py:75-88): appended domain parameters */
    int32_t aug_fixed;       /* append aug_nominal instead of the episode's values */
    int32_t aug_param[SRE_MAX_AUG];
    double aug_nominal[SRE_MAX_AUG];
} sre_step_wrappers;

typedef struct sre_env_cfg {
    int32_t kind;
    int32_t dtype;
    int32_t flags;
    int32_t wild_init; /* QCP init box: 0 -> 'True', 1 -> 'False', 2 -> else (quanser_cartpole.py:552-560) */
    double dt;
    int64_t max_steps; /* <= 0: inf */
    sre_task task;
    double base_params[SRE_MAX_DP]; /* domain parameters used where no per-env value is given */
    int32_t act_norm;               /* ActNormWrapper, environment_wrappers/action_normalization.py:66 */
    int32_t obs_noise;              /* GaussianObsNoiseWrapper, environment_wrappers/observation_noise.py:39 */
    double noise_std[SRE_MAX_DO];
    double noise_mean[SRE_MAX_DO];
    int32_t has_init_box;           /* init_space setter (base.py:99): a box replacing the default init space */
    int32_t reserved;
    double init_box_lo[SRE_MAX_DS];
    double init_box_up[SRE_MAX_DS];
    /* ObsNormWrapper._process_obs (environment_wrappers/observation_normalization.py:119-122), applied AFTER the
     * observation noise: obs = (obs - obs_lb) / (obs_ub - obs_lb) * 2 - 1 with the (overridden) bounds */
    int32_t obs_norm;
    /* GaussianActNoiseWrapper._process_act (environment_wrappers/action_noise.py:73-78): act + randn * std + mean.
     * SRE_ENV_ACT_NOISE_OUTER: the wrapper sits outside ActNormWrapper (noise in normalised units, before the
     * de-normalisation); SRE_ENV_ACT_NOISE_INNER: inside it (env units).  Philox stream 4, block = step index. */
    int32_t env_act_noise;
    double obs_lb[SRE_MAX_DO];
    double obs_ub[SRE_MAX_DO];
    double env_act_noise_std[SRE_MAX_DA];
    double env_act_noise_mean[SRE_MAX_DA];
    sre_step_wrappers wrap; /* stateful step-path wrappers of the persistent rollout (all zero: none) */
} sre_env_cfg;

/* DomainParam (domain_randomization/domain_parameter.py:46): float32 sample -> clamp -> optional round */
typedef struct sre_distr {
    int32_t param_index;
    int32_t kind;
    float a, b, c;
    float clip_lo, clip_up;
    int32_t roundint;
} sre_distr;

typedef struct sre_policy {
    int32_t kind;
    int32_t n_layers;                 /* FNN: number of Linear layers (hidden + output) */
    int32_t sizes[SRE_MAX_LAYERS + 1]; /* FNN: in, h1, ..., out */
    int32_t n_feats;                  /* Linear: number of feature functions */
    int32_t feats[SRE_MAX_FEATS];
    int32_t hidden_nonlin;            /* 0 tanh, 1 relu, 2 identity */
    int32_t output_nonlin;            /* 0 none, 1 tanh */
    int32_t cartpole_feats;           /* the fork's FNNPolicy input transform, fnn.py:214-218 */
    int32_t precision;                /* SRE_PREC_*: arithmetic of the 64x64 hidden layer (float builds, [in,64,64,out] tanh) */
    const float *weights;             /* device; torch layout W[out][in], bias[out], layers concatenated */
    int64_t n_weights;
    /* Step-based exploration fused into sre_rollout (StochasticActionExplStrat.forward,
     * exploration/stochastic_action.py:68-94): act = dist(policy(obs) + mean).rsample(), float32 like torch.
     * SRE_ACT_NOISE_NORMAL: NormalActNoiseExplStrat / DiagNormalNoise (stochastic_action.py:121, normal_noise.py:155),
     * act_noise_scale = std.  SRE_ACT_NOISE_UNIFORM: UniformActNoiseExplStrat (stochastic_action.py:183,
     * uniform_noise.py:160), act_noise_scale = halfspan.  Philox stream SRE act-noise, block = step index. */
    int32_t act_noise;
    int32_t reserved;
    float act_noise_scale[SRE_MAX_DA];
    float act_noise_mean[SRE_MAX_DA];
    /* Parameter-exploration sampling (ParameterExplorationSampler, sampling/parameter_exploration_sampler.py:151-345:
     * many policy-parameter sets x the same domains / init states): if group_size > 0, `weights` holds n_groups
     * consecutive sets of n_weights floats and lane n of a sre_rollout launch evaluates set n / group_size.  A
     * group_size that is a multiple of the CTA size (128 with `precision` != FP32 on a [in,64,64,out] net, else 256;
     * 128 for other shapes and the linear policy) runs the same kernels as a single set, one staged set per CTA;
     * any other group_size reads the weights per lane from global memory. */
    int64_t group_size;
    int64_t n_groups;
    /* parameters of the parametrised features of a Linear policy: HOST pointer to n_feat_params floats
     * (<= SRE_MAX_FEAT_PARAMS, copied into the launch), feat_arg[k] = offset of feature k's block (see SRE_FEAT_SIG ..) */
    const float *feat_params;
    int32_t n_feat_params;
    int32_t feat_arg[SRE_MAX_FEATS];
    /* optional HOST copy of `weights` (same layout, n_weights floats, read during the call): lets sre_rollout pass the small
     * layers of a [in,64,64,out] tensor-core policy as launch arguments instead of staging them in shared memory */
    const float *weights_host;
} sre_policy;

/* Inputs / outputs of one rollout launch.  Replaces rollout() (sampling/rollout.py:63) run for N envs by
 * ParallelRolloutSampler.sample (sampling/parallel_rollout_sampler.py:248). */
typedef struct sre_rollout_io {
    int64_t n_envs;           /* N (this rank's shard) */
    int64_t n_rows;           /* T: rows per env lane in `traj`/`row_flag`; a lane takes at most T - 1 steps */
    uint64_t seed;            /* Philox key */
    uint64_t env_offset;      /* global index of lane 0 (multi-GPU shard offset) */
    uint32_t episode;         /* first episode index of every lane */
    int32_t stop_on_done;     /* rollout(stop_on_done=...) */
    int32_t auto_reset;       /* 1: a lane starts episode+1 after its closing row (min_steps sampling) */
    int32_t n_distr;
    const sre_distr *distr;   /* HOST pointer to n_distr entries, or NULL */
    const void *params;       /* real [N, dp] or NULL -> cfg.base_params */
    const void *init;         /* real [N, n_init_cols] or NULL -> sample the init space */
    int32_t n_init_cols;
    int32_t reserved;
    const void *actions;      /* real [T-1, N, da] for SRE_POL_ACTIONS (NULL -> zeros) */
    /* outputs, any may be NULL */
    void *traj;               /* real [T, N, W] records, W = sre_record_width(kind) */
    uint8_t *row_flag;        /* [T, N] SRE_ROW_* */
    void *ret;                /* real [N]   undiscounted return of the lane's first episode */
    int32_t *length;          /* [N]        length of the lane's first episode */
    void *params_out;         /* real [N, dp] parameters of the lane's first episode */
    double *stats;            /* [SRE_N_STATS] accumulated with atomics; caller initialises (sre_stats_init) */
    int32_t *status;          /* [N] SRE_ST_* bits */
    /* Placement of this launch's lanes inside a WIDER record buffer: `traj` is real [T, traj_env_stride, W] and
     * `row_flag` [T, traj_env_stride], lane n writes column traj_env_offset + n.  0 / 0: the buffers are [T, N, ...].
     * This is how a rank writes its shard of the rollouts straight into the learner's buffer over NVLink (the pointers
     * may be peer-mapped device memory): the gather of ParallelRolloutSampler's results overlaps the simulation tile by
     * tile instead of following it as a collective. */
    int64_t traj_env_stride;
    int64_t traj_env_offset;
} sre_rollout_io;

/* ---- introspection (host only) ---------------------------------------------------------------------- */
int sre_abi_version(void);
const char *sre_last_error(void);
int sre_device_count(int *count);
/* dims of an env kind: state, obs, act, domain-param, hidden-state, init-space dims (SimPyEnv spaces) */
int sre_env_dims(int kind, int *ds, int *dobs, int *da, int *dp, int *nh, int *dinit);
/* get_nominal_domain_param(): quanser_qube.py:54, quanser_cartpole.py:111, quanser_ball_balancer.py:169,
 * ball_on_beam.py:75, one_mass_oscillator.py:81.  dp: host double[SRE_MAX_DP] */
int sre_nominal_params(int kind, int long_pole, double *dp);
/* default task of the env (_create_task): fills cfg->task.  host */
int sre_default_task(int kind, sre_task *task);
/* record layout of `traj`: W and the offsets of obs, act, act_app, rew, state, hidden inside a record */
int sre_record_width(int kind);
int sre_record_layout(int kind, int *off_obs, int *off_act, int *off_act_app, int *off_rew, int *off_state,
                      int *off_hidden);
int sre_stats_init(double *stats, void *stream);

/* ---- domain parameters ------------------------------------------------------------------------------ */
/* domain_param setter side effects (base.py:112-124): _calc_constants + _create_spaces + ScaledExp c_max.
 * params real [N, dp]; outputs real: st_lo/st_up [N, ds], act_lo/act_up [N, da], c_max [N]; any may be NULL */
int sre_derive(const sre_env_cfg *cfg, const void *params, int64_t n, void *st_lo, void *st_up, void *act_lo,
               void *act_up, void *c_max, void *stream);
/* DomainRandomizer.randomize + get_params (domain_randomizer.py:123,159) for episode `episode` of envs
 * env_offset .. env_offset + n - 1: overwrites the listed columns of params real [N, dp]. distr: HOST */
int sre_sample_params(const sre_env_cfg *cfg, const sre_distr *distr, int n_distr, uint64_t seed,
                      uint64_t env_offset, uint32_t episode, void *params, int64_t n, void *stream);
/* init_space.sample_uniform() (spaces/box.py:169, polar.py:108, compound.py:83): init real [N, dinit] */
int sre_sample_init(const sre_env_cfg *cfg, const void *params, uint64_t seed, uint64_t env_offset,
                    uint32_t episode, void *init, int64_t n, void *stream);

/* ---- step-granular API: SimPyEnv.reset / step (base.py:166, :217) ------------------------------------- */
/* mask: uint8 [N] or NULL (all).  init real [N, n_init_cols] (n_init_cols == ds: full state, else init space).
 * Writes state [N, ds], hidden [N, max(nh,1)], curr_step int32 [N], obs [N, do] (obs may be NULL).
 * seed/env_offset/episode key the observation noise when cfg.obs_noise (Philox block t = 0). */
int sre_reset(const sre_env_cfg *cfg, const void *params, const void *init, int n_init_cols, const uint8_t *mask,
              void *state, void *hidden, int32_t *curr_step, void *obs, uint64_t seed, uint64_t env_offset,
              uint32_t episode, int64_t n, void *stream);
/* act real [N, da] in the units of the outermost fused wrapper (normalised if cfg.act_norm).
 * Updates state/hidden/curr_step in place (curr_step: count | SRE_STEP_YIELDED_FINAL, see above); writes obs [N, do],
 * rew [N], done uint8 [N], status int32 [N].
 * noise_seed/env_offset/episode key the observation noise when cfg.obs_noise. */
int sre_step(const sre_env_cfg *cfg, const void *params, const void *act, void *state, void *hidden,
             int32_t *curr_step, void *obs, void *rew, uint8_t *done, int32_t *status, uint64_t seed,
             uint64_t env_offset, uint32_t episode, int64_t n, void *stream);
/* Env.observe (environments/base.py:203 and the overrides): obs [N, do] from state [N, ds] */
int sre_observe(const sre_env_cfg *cfg, const void *state, void *obs, int64_t n, void *stream);
/* Task.step_rew (tasks/base.py:108, tasks/desired_state.py:107,146) for a batch: state [N, ds], act [N, da]
 * (UNCLIPPED, inner-env units) -> rew [N]; params as in sre_derive (c_max of ScaledExp depends on them) */
int sre_step_rew(const sre_env_cfg *cfg, const void *params, const void *state, const void *act, void *rew, int64_t n,
                 void *stream);
/* The same for a free-standing task without an env (Task / RewFcn objects, tasks/reward_functions.py:202-297):
 * task->Q is row-major [ds, ds], task->R [da, da]; c_max only matters for SRE_REW_SCALED_EXP */
int sre_task_step_rew(const sre_task *task, int dtype, int ds, int da, double c_max, const void *state,
                      const void *act, void *rew, int64_t n, void *stream);
/* Policy.forward for a batch: obs float [N, n_obs] -> act float [N, n_act] (always float32) */
int sre_policy_forward(const sre_policy *pol, const float *obs, int n_obs, float *act, int n_act, int64_t n,
                       void *stream);

/* ---- post-processing of the record buffer on the device ------------------------------------------------ */
/* discounted_reverse_cumsum / discounted_value / gae_returns (sampling/step_sequence.py:1049-1116) for every episode
 * of every lane of a time-major buffer.  rew: real, element (t, n) at rew[t * rew_row_stride + n * rew_lane_stride]
 * (N*W, W with an offset pointer for the `rew` column of sre_rollout_io.traj; N, 1 for a contiguous [T, N] array);
 * row_flag [T, N] as written by sre_rollout; values real [T, N] (value estimate of the observation of every row,
 * closing rows included) or NULL; out_ret real [T, N] discounted return-to-go; out_adv real [T, N] generalised
 * advantage estimate (NULL iff values is NULL).  Rows without a step (closing / idle) get 0.  A SRE_ROW_TRUNCATED
 * closing row bootstraps the episode before it from its value estimate. */
int sre_returns(int dtype, const void *rew, int64_t rew_row_stride, int64_t rew_lane_stride, const uint8_t *row_flag,
                const void *values, double gamma, double lamb, void *out_ret, void *out_adv, int64_t n_rows,
                int64_t n_envs, void *stream);

/* ---- persistent-horizon rollout ----------------------------------------------------------------------- */
int sre_rollout(const sre_env_cfg *cfg, const sre_policy *pol, const sre_rollout_io *io, void *stream);
/* number of kernels the last sre_rollout call on this thread launched (for bench.py's gpu_launches) */
int sre_last_launch_count(void);

/* ---- differentiable cart-pole step (quanser_cartpole.py:257-358, state_based=True) ------------------- */
/* Forward over a horizon with injected actions.  act real [H, N] in inner-env units BEFORE the clamp
 * (act + act_diff of step_diff_state); s0 real [N, 4]; thdd0 real [N] or NULL (zeros).
 * Outputs: states real [H+1, N, 4] (row 0 = s0), thdd real [H+1, N] (carry before each step), rew real [H, N],
 * done uint8 [H, N].  states/thdd are the tape the backward pass reads. */
int sre_qcp_diff_forward(const sre_env_cfg *cfg, const void *params, const void *s0, const void *thdd0,
                         const void *act, int64_t n, int64_t horizon, void *states, void *thdd, void *rew,
                         uint8_t *done, void *stream);
/* Adjoint through the horizon.  Inputs: the tape, g_states real [H+1, N, 4] (dL/d states, row 0 = dL/d s0
 * direct term), g_rew real [H, N], g_thdd real [H+1, N] or NULL.  Outputs: g_s0 real [N, 4], g_act real [H, N],
 * g_thdd0 real [N] (may be NULL). */
int sre_qcp_diff_backward(const sre_env_cfg *cfg, const void *params, const void *states, const void *thdd,
                          const void *act, const void *g_states, const void *g_rew, const void *g_thdd,
                          int64_t n, int64_t horizon, void *g_s0, void *g_act, void *g_thdd0, void *stream);
/* Per-step Jacobians d(next_state, rew)/d(state, act): jac real [N, 5, 5] (rows s'0..3, rew; cols s0..3, act),
 * the quantities rollout_tensor_full collects (sampling/rollout.py:832-837). */
int sre_qcp_diff_jacobian(const sre_env_cfg *cfg, const void *params, const void *state, const void *thdd,
                          const void *act, int64_t n, void *jac, void *stream);

/* ---- no-allocation entry point ------------------------------------------------------------------------ */
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * sre_qcp_diff_forward followed by sre_qcp_diff_backward in one call (same kernels, same arguments, same results),
 * with the tape between them -- states real [H+1, N, 4] and thdd real [H+1, N], which the other two entry points take
 * from the caller -- in a file-scope static device array of SRE_STATIC_TAPE_BYTES that every GPU loading the library
 * holds until exit.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine); its only
 * set-up is one cudaGetSymbolAddress per device, on the first call with that device current.
 * Limit: align256((H+1) * N * 4 * sizeof(real)) + (H+1) * N * sizeof(real) <= SRE_STATIC_TAPE_BYTES (float32: about
 * 1.67 M (row, env) pairs, e.g. 4096 envs x 400 steps; float64: half of that); larger sizes return SRE_E_ARG with a
 * message, device ordinals >= SRE_STATIC_MAX_DEVICES return SRE_E_UNSUPPORTED.
 * Every call on a device shares the one tape: calls that overlap on the GPU (two streams, two host threads) RACE --
 * serialise them on one stream.  Nothing in the tape has to start at zero (the forward pass writes every element the
 * adjoint reads), so it is not cleared; it keeps the last call's contents.
 * Inputs as sre_qcp_diff_forward (params, s0, thdd0, act) and sre_qcp_diff_backward (g_states, g_rew, g_thdd, any of
 * them NULL = zeros); outputs rew real [H, N] / done uint8 [H, N] (may be NULL), g_s0 real [N, 4], g_act real [H, N],
 * g_thdd0 real [N] (may be NULL). */
#define SRE_STATIC_TAPE_BYTES (32u << 20)
#define SRE_STATIC_MAX_DEVICES 64
int sre_qcp_diff_grad_static(const sre_env_cfg *cfg, const void *params, const void *s0, const void *thdd0,
                             const void *act, const void *g_states, const void *g_rew, const void *g_thdd, int64_t n,
                             int64_t horizon, void *rew, uint8_t *done, void *g_s0, void *g_act, void *g_thdd0,
                             void *stream);

/* ---- measurement helpers (roofline denominators measured on the box) ---------------------- */
/* FMA-only kernel: every thread runs `iters` x 32 dependent-free FFMA (dtype F32) or DFMA (F64); dtype 2: the same on
 * packed fma.rn.f32x2; dtype 3 / 4: FFMA / FFMA2 whose three operands are distinct, changing registers (the form
 * simulation code issues: no immediate, constant-bank or reuse-cache operand); dtype 5 / 6 / 7: per iteration 16 such FFMA2,
 * 8 MUFU.EX2, or both (do the FMA and the XU pipe overlap?); dtype 8 .. 11: 16 packed instructions per iteration by operand
 * form (fma2 with a loop-invariant register / a launch-uniform operand, mul2 of two registers, add2 with an immediate).
 * Returns FLOPs issued in *flops; time it with CUDA events on `stream`. */
int sre_bench_fma(int dtype, int64_t iters, int blocks, int threads, float *sink, double *flops, void *stream);
/* streaming write of `bytes` with 128-bit stores (the HBM-write denominator) */
int sre_bench_hbm_write(void *dst, int64_t bytes, void *stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* SRE_H */
