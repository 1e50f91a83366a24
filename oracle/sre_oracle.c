/*
 * sre_oracle.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement (plain C, fp64) of the reference's
 * pysim hot path.  It is the parity oracle for the CUDA engine and the timed "port" CPU baseline.
 * Nothing in the product package (simurlacra_b200/) links, imports or executes this file; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
 *
 * Parity status: PINNED.  The reference is Python and is importable in the build container
 * (oracle/ref_shim.py); oracle/gen_golden.py ran the unmodified reference and committed its
 * trajectories under tests/golden/.  tests/test_oracle_golden.py checks this file against those
 * vectors (fp64, rel 1e-12) and against the reference's own known-answer tests
 * (Pyrado/tests/test_tasks.py:82-99).
 *
 * Every function cites the reference lines it follows.  Paths are relative to
 * /root/reference/Pyrado/pyrado/ .  The operation ORDER of the reference is kept on purpose (e.g. the
 * QQube pseudo-RK4 is executed stage by stage), so the CUDA kernels -- which use algebraically
 * reduced forms -- are checked against an independent formulation.
 *
 * The counter-based RNG (Philox4x32-10, Salmon et al., SC'11, "Parallel random numbers: as easy as
 * 1, 2, 3"; Random123 v1.14 constants) and the sample mapping are the ENGINE's own specification
 * (the reference uses torch/numpy Mersenne-Twister streams that cannot be reproduced per-env); they
 * are restated here independently from the published algorithm and pinned by the Random123
 * known-answer vectors in tests/test_oracle_golden.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_QQ 0
#define ORC_QCP 1
#define ORC_QBB 2
#define ORC_BOB 3
#define ORC_OMO 4
#define ORC_PEND 5

#define ORC_MAX_DS 8
#define ORC_MAX_DO 8
#define ORC_MAX_DA 2
#define ORC_MAX_DP 20
#define ORC_MAX_NH 2

/* flags */
#define ORC_F_SIMPLE_DYN 1 /* QCartPole / QBB simple_dynamics=True */
#define ORC_F_QCP_STAB 2   /* QCartPoleStabSim state space, quanser_cartpole.py:473-489 */
#define ORC_F_DISC_ACT 4   /* BallOnBeamDiscSim: 3-element DiscreteSpace over the action bounds, ball_on_beam.py:155-162 */

/* reward function kinds: tasks/reward_functions.py:176 (Quadr), :224 (ExpQuadr), :247 (ScaledExpQuadr) */
#define ORC_REW_QUADR 0
#define ORC_REW_EXP 1
#define ORC_REW_SCALED_EXP 2

/* FinalRewMode bits: tasks/final_reward.py:41-48 */
#define ORC_FR_ENABLED 1
#define ORC_FR_STATE_DEP 2
#define ORC_FR_TIME_DEP 4
#define ORC_FR_ALWAYS_POS 8
#define ORC_FR_ALWAYS_NEG 16

static const int k_ds[6] = {4, 4, 8, 4, 2, 2};
static const int k_do[6] = {6, 5, 8, 4, 2, 3};
static const int k_da[6] = {1, 1, 2, 1, 1, 1};
static const int k_dp[6] = {11, 17, 20, 8, 3, 5};
static const int k_nh[6] = {0, 1, 2, 0, 0, 0};
static const int k_dinit[6] = {4, 4, 4, 4, 2, 2};

typedef struct {
    int kind, flags;
    double dt;
    long max_steps; /* <=0 means inf */
    double dp[ORC_MAX_DP];
    /* derived constants (per-env meaning) */
    double c[8];
    /* spaces (param dependent) */
    double st_lo[ORC_MAX_DS], st_up[ORC_MAX_DS];
    double act_lo[ORC_MAX_DA], act_up[ORC_MAX_DA];
    /* task */
    double state_des[ORC_MAX_DS];
    double Q[ORC_MAX_DS * ORC_MAX_DS];
    double R[ORC_MAX_DA * ORC_MAX_DA];
    int rew_kind;
    double min_rew, c_max;
    unsigned radial_mask; /* bit i set: state index i is modulated (tasks/desired_state.py:146) */
    double radial_mod[ORC_MAX_DS]; /* modulation factor per state index, default 2*pi (desired_state.py:127,144) */
    int final_flags;
    double final_factor;
    int yielded_final;
    /* runtime */
    double state[ORC_MAX_DS];
    double hidden[ORC_MAX_NH];
    long curr_step;
} orc_env;

int orc_ds(int kind) { return k_ds[kind]; }
int orc_do(int kind) { return k_do[kind]; }
int orc_da(int kind) { return k_da[kind]; }
int orc_dp(int kind) { return k_dp[kind]; }
int orc_nh(int kind) { return k_nh[kind]; }
int orc_sizeof_env(void) { return (int)sizeof(orc_env); }

/* ---------------------------------------------------------------------------------------------
 * Nominal domain parameters, in the key order of the reference's get_nominal_domain_param():
 *   QQ  environments/pysim/quanser_qube.py:54-68
 *   QCP environments/pysim/quanser_cartpole.py:111-143 (long / short pole)
 *   QBB environments/pysim/quanser_ball_balancer.py:169-202 (+ default thresholds :135-138)
 *   BOB environments/pysim/ball_on_beam.py:75-87
 *   OMO environments/pysim/one_mass_oscillator.py:81-87
 * ------------------------------------------------------------------------------------------- */
void orc_nominal_params(int kind, int long_pole, double *dp) {
    switch (kind) {
    case ORC_QQ: {
        const double v[11] = {9.81, 8.4, 0.042, 0.095, 0.085, 5e-6, 0.024, 0.129, 1e-6, 0.0, 0.0};
        memcpy(dp, v, sizeof v);
    } break;
    case ORC_QCP: {
        double m_pole = long_pole ? 0.23 : 0.127;
        double l_pole = long_pole ? 0.641 / 2 : 0.3365 / 2;
        const double v[17] = {9.81, 0.58, 0.814, 0.9, 0.9, 3.71, 3.9e-7, 6.35e-3, 2.6, 7.67e-3, 0.0024, 5.4,
                              m_pole, l_pole, 0.02, 0.0, 0.0};
        memcpy(dp, v, sizeof v);
    } break;
    case ORC_QBB: {
        const double v[20] = {9.81,  0.003, 0.019625, 0.275, 0.0254, 70.0, 0.9,  5.2822e-5, 4.6063e-7, 0.0077,
                              2.6,   0.69,  0.015,    0.05,  0.28,   -0.10, 0.28, -0.074,    0.0,       0.0};
        memcpy(dp, v, sizeof v);
    } break;
    case ORC_BOB: {
        const double v[8] = {9.81, 0.5, 0.1, 3.0, 2.0, 0.1, 0.05, 0.0};
        memcpy(dp, v, sizeof v);
    } break;
    case ORC_OMO: {
        const double v[3] = {1.0, 30.0, 0.5};
        memcpy(dp, v, sizeof v);
    } break;
    case ORC_PEND: { /* pendulum.py:96-104 */
        const double v[5] = {9.81, 1.0, 1.0, 0.05, 3.5};
        memcpy(dp, v, sizeof v);
    } break;
    }
}

/* ---------------------------------------------------------------------------------------------
 * _calc_constants + _create_spaces + (ScaledExp) rew_fcn.reset : what the domain_param setter does,
 * environments/pysim/base.py:112-124.
 * ------------------------------------------------------------------------------------------- */
static double quad_form(const double *M, const double *v, int n) {
    /* err.dot(M.dot(err)), tasks/reward_functions.py:210 */
    double tmp[ORC_MAX_DS];
    for (int i = 0; i < n; i++) {
        double s = 0.0;
        for (int j = 0; j < n; j++) s += M[i * n + j] * v[j];
        tmp[i] = s;
    }
    double r = 0.0;
    for (int i = 0; i < n; i++) r += v[i] * tmp[i];
    return r;
}

static void orc_update_cmax(orc_env *e) {
    /* ScaledExpQuadrErrRewFcn.reset, tasks/reward_functions.py:284-297 */
    if (e->rew_kind != ORC_REW_SCALED_EXP) return;
    int ds = k_ds[e->kind], da = k_da[e->kind];
    double smax[ORC_MAX_DS], amax[ORC_MAX_DA];
    for (int i = 0; i < ds; i++) smax[i] = fmax(fabs(e->st_lo[i]), fabs(e->st_up[i])); /* bound_abs_up */
    for (int i = 0; i < da; i++) amax[i] = fmax(fabs(e->act_lo[i]), fabs(e->act_up[i]));
    double max_cost = quad_form(e->Q, smax, ds) + quad_form(e->R, amax, da);
    e->c_max = -1.0 * log(e->min_rew) / max_cost;
}

static void orc_derive(orc_env *e) {
    const double *p = e->dp;
    const double pi = M_PI;
    switch (e->kind) {
    case ORC_QQ: {
        /* quanser_qube.py:70-87 */
        double g = p[0], mr = p[3], lr = p[4], mp = p[6], lp = p[7];
        double Jr = mr * (lr * lr) / 12;
        double Jp = mp * (lp * lp) / 12;
        e->c[0] = Jr + mp * (lr * lr);
        e->c[1] = 0.25 * mp * (lp * lp);
        e->c[2] = 0.5 * mp * lp * lr;
        e->c[3] = Jp + e->c[1];
        e->c[4] = 0.5 * mp * lp * g;
        /* quanser_qube.py:167-181 */
        const double ms[4] = {115.0 / 180 * pi, 4 * pi, 20 * pi, 20 * pi};
        for (int i = 0; i < 4; i++) { e->st_lo[i] = -ms[i]; e->st_up[i] = ms[i]; }
        e->act_lo[0] = -4.5; e->act_up[0] = 4.5; /* MAX_ACT_QQ, environments/quanser/__init__.py:34 */
    } break;
    case ORC_QCP: {
        /* quanser_cartpole.py:145-155 */
        double m_c = p[1], eta_g = p[4], K_g = p[5], J_m = p[6], r_mp = p[7], m_p = p[12], l_p = p[13];
        e->c[0] = (l_p * l_p) * m_p / 3.0;                            /* J_pole */
        e->c[1] = m_c + (eta_g * (K_g * K_g) * J_m) / (r_mp * r_mp); /* J_eq */
        /* quanser_cartpole.py:541-549, x_buffer = 0.15 (:79) */
        double l_rail = p[2];
        const double up[4] = {+l_rail / 2.0 - 0.15, +4 * pi, 1 * l_rail, 20 * pi};
        const double lo[4] = {-l_rail / 2.0 + 0.15, -4 * pi, -1 * l_rail, -20 * pi};
        for (int i = 0; i < 4; i++) { e->st_lo[i] = lo[i]; e->st_up[i] = up[i]; }
        if (e->flags & ORC_F_QCP_STAB) {
            /* QCartPoleStabSim._create_spaces, quanser_cartpole.py:473-482: stab_thold = 15 deg (:467) */
            double thold = 15 / 180.0 * pi;
            e->st_lo[1] = pi - thold; e->st_up[1] = pi + thold;
            e->st_lo[3] = -2 * pi; e->st_up[3] = +2 * pi;
        }
        e->act_lo[0] = -6.0; e->act_up[0] = 6.0; /* MAX_ACT_QCP */
    } break;
    case ORC_QBB: {
        /* quanser_ball_balancer.py:204-223 */
        double m_ball = p[1], r_ball = p[2], l_plate = p[3], r_arm = p[4], K_g = p[5], eta_g = p[6], J_l = p[7],
               J_m = p[8], k_m = p[9], R_m = p[10], eta_m = p[11], B_eq = p[12];
        e->c[0] = 2.0 / 5 * m_ball * (r_ball * r_ball);                        /* J_ball */
        e->c[1] = eta_g * (K_g * K_g) * J_m + J_l;                             /* J_eq */
        e->c[2] = 2.0 * r_arm / l_plate;                                       /* c_kin */
        e->c[3] = eta_g * K_g * eta_m * k_m / R_m;                             /* A_m */
        e->c[4] = eta_g * (K_g * K_g) * eta_m * (k_m * k_m) / R_m + B_eq;      /* B_eq_v */
        e->c[5] = m_ball * (r_ball * r_ball) + e->c[0];                        /* zeta */
        /* quanser_ball_balancer.py:93-117 */
        const double ms[8] = {pi / 4.0, pi / 4.0, l_plate / 2.0, l_plate / 2.0, 5 * pi, 5 * pi, 0.5, 0.5};
        for (int i = 0; i < 8; i++) { e->st_lo[i] = -ms[i]; e->st_up[i] = ms[i]; }
        e->act_lo[0] = e->act_lo[1] = -3.0; e->act_up[0] = e->act_up[1] = 3.0; /* MAX_ACT_QBB */
    } break;
    case ORC_BOB: {
        /* ball_on_beam.py:89-98 */
        double g = p[0], m_ball = p[1], r_ball = p[2], m_beam = p[3], l_beam = p[4], d_beam = p[5];
        e->c[0] = 2.0 / 5 * m_ball * (r_ball * r_ball);                 /* J_ball */
        e->c[1] = 1.0 / 12 * m_beam * (l_beam * l_beam + d_beam * d_beam); /* J_beam */
        e->c[2] = m_ball + e->c[0] / (r_ball * r_ball);                 /* zeta_ball */
        /* ball_on_beam.py:48-58 */
        const double ms[4] = {l_beam / 2.0, pi / 4.0, 10.0, pi};
        for (int i = 0; i < 4; i++) { e->st_lo[i] = -ms[i]; e->st_up[i] = ms[i]; }
        double max_act = l_beam / 2.0 * g * 3.0;
        e->act_lo[0] = -max_act; e->act_up[0] = max_act;
    } break;
    case ORC_OMO: {
        /* one_mass_oscillator.py:89-103 */
        double m = p[0], k = p[1], d = p[2];
        e->c[0] = sqrt(k / m);                /* omega */
        e->c[1] = d / (2.0 * sqrt(m * k));    /* zeta */
        /* one_mass_oscillator.py:54-66 */
        e->st_lo[0] = -1.0; e->st_up[0] = 1.0; e->st_lo[1] = -10.0; e->st_up[1] = 10.0;
        e->act_lo[0] = -1.0 * k; e->act_up[0] = 1.0 * k;
    } break;
    case ORC_PEND: {
        /* pendulum.py:66-74: max_state = [4 pi, 4 pi], act bound = torque_thold */
        e->st_lo[0] = -4 * pi; e->st_up[0] = 4 * pi; e->st_lo[1] = -4 * pi; e->st_up[1] = 4 * pi;
        e->act_lo[0] = -p[4]; e->act_up[0] = p[4];
    } break;
    }
    orc_update_cmax(e);
}

/* Default tasks:
 *   QQ  quanser_qube.py:183-188   QCP quanser_cartpole.py:573-588   QBB quanser_ball_balancer.py:121-131
 *   BOB ball_on_beam.py:100-108   OMO one_mass_oscillator.py:68-79 */
static void orc_default_task(orc_env *e) {
    int ds = k_ds[e->kind], da = k_da[e->kind];
    memset(e->state_des, 0, sizeof e->state_des);
    memset(e->Q, 0, sizeof e->Q);
    memset(e->R, 0, sizeof e->R);
    e->radial_mask = 0; e->final_flags = 0; e->final_factor = 0.0; e->min_rew = 1e-4; e->c_max = 1.0;
    for (int i = 0; i < ORC_MAX_DS; i++) e->radial_mod[i] = 2 * M_PI;
    double q[ORC_MAX_DS] = {0}, r[ORC_MAX_DA] = {0};
    switch (e->kind) {
    case ORC_QQ:
        e->state_des[1] = M_PI; q[0] = 1.0; q[1] = 1.0; q[2] = 2e-2; q[3] = 5e-3; r[0] = 4e-3;
        e->rew_kind = ORC_REW_EXP; e->radial_mask = 1u << 1; break;
    case ORC_QCP:
        e->state_des[1] = M_PI; q[0] = 3e-1; q[1] = 5e-1; q[2] = 5e-3; q[3] = 1e-3; r[0] = 1e-3;
        e->rew_kind = ORC_REW_EXP; e->radial_mask = 1u << 1; break;
    case ORC_QBB:
        q[0] = 1e0; q[1] = 1e0; q[2] = 5e3; q[3] = 5e3; q[4] = 1e-2; q[5] = 1e-2; q[6] = 5e-1; q[7] = 5e-1;
        r[0] = 1e-2; r[1] = 1e-2; e->rew_kind = ORC_REW_SCALED_EXP; break;
    case ORC_BOB:
        q[0] = 1e5; q[1] = 1e3; q[2] = 1e3; q[3] = 1e2; r[0] = 1.0; e->rew_kind = ORC_REW_SCALED_EXP; break;
    case ORC_OMO:
        q[0] = 1e1; q[1] = 1e-2; r[0] = 1e-6; e->rew_kind = ORC_REW_QUADR;
        e->final_flags = ORC_FR_ENABLED | ORC_FR_ALWAYS_NEG; e->final_factor = 1e3; break;
    case ORC_PEND: /* pendulum.py:82-88: RadiallySymmDesStateTask(..., idcs=[1]) as the reference has it */
        e->state_des[0] = M_PI; q[0] = 1e-0; q[1] = 1e-3; r[0] = 1e-2;
        e->rew_kind = ORC_REW_EXP; e->radial_mask = 1u << 1; break;
    }
    for (int i = 0; i < ds; i++) e->Q[i * ds + i] = q[i];
    for (int i = 0; i < da; i++) e->R[i * da + i] = r[i];
}

void orc_init(orc_env *e, int kind, double dt, long max_steps, int flags, int long_pole) {
    memset(e, 0, sizeof *e);
    e->kind = kind; e->dt = dt; e->max_steps = max_steps; e->flags = flags;
    orc_nominal_params(kind, long_pole, e->dp);
    orc_default_task(e);
    orc_derive(e);
}

orc_env *orc_create(int kind, double dt, long max_steps, int flags, int long_pole) {
    orc_env *e = (orc_env *)malloc(sizeof(orc_env));
    orc_init(e, kind, dt, max_steps, flags, long_pole);
    return e;
}
void orc_destroy(orc_env *e) { free(e); }

void orc_set_params(orc_env *e, const double *dp) {
    memcpy(e->dp, dp, sizeof(double) * k_dp[e->kind]);
    orc_derive(e);
}
void orc_get_params(const orc_env *e, double *dp) { memcpy(dp, e->dp, sizeof(double) * k_dp[e->kind]); }

void orc_set_task(orc_env *e, const double *state_des, const double *Q, const double *R, int rew_kind,
                  double min_rew, unsigned radial_mask, int final_flags, double final_factor) {
    int ds = k_ds[e->kind], da = k_da[e->kind];
    memcpy(e->state_des, state_des, sizeof(double) * ds);
    memcpy(e->Q, Q, sizeof(double) * ds * ds);
    memcpy(e->R, R, sizeof(double) * da * da);
    e->rew_kind = rew_kind; e->min_rew = min_rew; e->radial_mask = radial_mask;
    e->final_flags = final_flags; e->final_factor = fabs(final_factor); /* final_reward.py:99 */
    orc_update_cmax(e);
}

/* per-index modulation factors of RadiallySymmDesStateTask (desired_state.py:127,144); entries for indices
 * outside radial_mask are ignored */
void orc_set_radial_mod(orc_env *e, const double *mod) { memcpy(e->radial_mod, mod, sizeof(double) * k_ds[e->kind]); }

void orc_get_spaces(const orc_env *e, double *st_lo, double *st_up, double *act_lo, double *act_up) {
    memcpy(st_lo, e->st_lo, sizeof(double) * k_ds[e->kind]);
    memcpy(st_up, e->st_up, sizeof(double) * k_ds[e->kind]);
    memcpy(act_lo, e->act_lo, sizeof(double) * k_da[e->kind]);
    memcpy(act_up, e->act_up, sizeof(double) * k_da[e->kind]);
}
double orc_get_cmax(const orc_env *e) { return e->c_max; }
void orc_get_state(const orc_env *e, double *state, double *hidden) {
    memcpy(state, e->state, sizeof(double) * k_ds[e->kind]);
    if (hidden) memcpy(hidden, e->hidden, sizeof(double) * ORC_MAX_NH);
}
void orc_set_state(orc_env *e, const double *state, const double *hidden) {
    memcpy(e->state, state, sizeof(double) * k_ds[e->kind]);
    if (hidden) memcpy(e->hidden, hidden, sizeof(double) * ORC_MAX_NH);
}
long orc_curr_step(const orc_env *e) { return e->curr_step; }

/* ---------------------------------------------------------------------------------------------
 * observe: environments/base.py:203-213 (copy); quanser_qube.py:148-149; quanser_cartpole.py:108-109
 * ------------------------------------------------------------------------------------------- */
void orc_observe(const orc_env *e, const double *s, double *obs) {
    switch (e->kind) {
    case ORC_QQ:
        obs[0] = sin(s[0]); obs[1] = cos(s[0]); obs[2] = sin(s[1]); obs[3] = cos(s[1]); obs[4] = s[2]; obs[5] = s[3];
        break;
    case ORC_QCP:
        obs[0] = s[0]; obs[1] = sin(s[1]); obs[2] = cos(s[1]); obs[3] = s[2]; obs[4] = s[3];
        break;
    case ORC_PEND: /* pendulum.py:93-94 */
        obs[0] = sin(s[0]); obs[1] = cos(s[0]); obs[2] = s[1];
        break;
    default:
        memcpy(obs, s, sizeof(double) * k_ds[e->kind]);
    }
}

/* ---------------------------------------------------------------------------------------------
 * QBallBalancerKin: quanser_ball_balancer.py:375-444.  100 iterations of torch SGD(lr=0.01,
 * momentum=0.9) in float32 on the two-constraint loss (:423-436), analytic gradient written out
 * (the reference uses autograd), then plate_ang (:438-444).
 * torch SGD with momentum (dampening 0, no nesterov): buf = g at the first step, afterwards
 * buf = 0.9*buf + g ; p -= lr*buf.
 * ------------------------------------------------------------------------------------------- */
double orc_qbb_kin(double arm_radius, double plate_length, double th_in) {
    const float r = (float)arm_radius;
    const float l = (float)(plate_length / 2.0);
    const float d = 0.10f;
    const float th = (float)th_in;
    float tip0 = r, tip1 = l;
    float buf0 = 0.f, buf1 = 0.f;
    const float cx = r * cosf(th), cy = r * sinf(th);
    for (int i = 0; i < 100; i++) {
        float ax = tip0 - cx, ay = tip1 - cy;
        float rod_len = sqrtf(ax * ax + ay * ay);
        float bx = tip0 - r - l, by = tip1 - d;
        float half_plate = sqrtf(bx * bx + by * by);
        float e1 = rod_len - d, e2 = half_plate - l;
        /* autograd order: pow' -> 2*e ; sqrt' -> /(2*len) ; pow' -> *(2*a)  (scalings by 2 are exact) */
        float t1 = e1 / rod_len, t2 = e2 / half_plate;
        float g0 = t1 * (2.f * ax) + t2 * (2.f * bx);
        float g1 = t1 * (2.f * ay) + t2 * (2.f * by);
        if (i == 0) { buf0 = g0; buf1 = g1; }
        else { buf0 = 0.9f * buf0 + g0; buf1 = 0.9f * buf1 + g1; }
        tip0 = tip0 - 0.01f * buf0;
        tip1 = tip1 - 0.01f * buf1;
    }
    /* np.pi/2.0 - to.atan2(...) : python float minus float32 tensor -> float32 result */
    float rl = (float)(arm_radius + plate_length / 2.0); /* python-float sum, then cast (:443) */
    float ang = (float)(M_PI / 2.0) - atan2f(rl - tip0, tip1 - d);
    return (double)ang;
}

/* ---------------------------------------------------------------------------------------------
 * reset: environments/pysim/base.py:166-203 with a GIVEN init state (sampling is in orc_sample_init).
 * n_init == ds: full state (:184-188); else element of the init space -> _state_from_init (:193).
 * QCP zeroes _th_ddot (quanser_cartpole.py:101-106); QBB runs the IK (quanser_ball_balancer.py:231-245).
 * ------------------------------------------------------------------------------------------- */
void orc_reset(orc_env *e, const double *init, int n_init, double *obs) {
    int ds = k_ds[e->kind];
    e->curr_step = 0;
    e->yielded_final = 0;
    if (n_init == ds) {
        memcpy(e->state, init, sizeof(double) * ds);
    } else {
        /* only QBB has an init space smaller than its state space: quanser_ball_balancer.py:225-229 */
        memset(e->state, 0, sizeof(double) * ds);
        e->state[2] = init[0]; e->state[3] = init[1]; e->state[6] = init[2]; e->state[7] = init[3];
    }
    e->hidden[0] = e->hidden[1] = 0.0;
    if (e->kind == ORC_QBB && !(e->flags & ORC_F_SIMPLE_DYN)) {
        e->hidden[0] = orc_qbb_kin(e->dp[4], e->dp[3], e->state[0] + e->dp[18]);
        e->hidden[1] = orc_qbb_kin(e->dp[4], e->dp[3], e->state[1] + e->dp[19]);
    }
    orc_update_cmax(e); /* task.reset -> rew_fcn.reset, tasks/desired_state.py:90-104 */
    if (obs) orc_observe(e, e->state, obs);
}

/* ---------------------------------------------------------------------------------------------
 * Task.step_rew: tasks/desired_state.py:107-110 (DesStateTask), :146-155 (RadiallySymm...),
 * reward functions tasks/reward_functions.py:202-222, :237-244, :276-282.
 * ------------------------------------------------------------------------------------------- */
double orc_step_rew(const orc_env *e, const double *state, const double *act) {
    int ds = k_ds[e->kind], da = k_da[e->kind];
    double err[ORC_MAX_DS], ea[ORC_MAX_DA];
    for (int i = 0; i < ds; i++) err[i] = e->state_des[i] - state[i];
    if (e->radial_mask) {
        for (int i = 0; i < ds; i++)
            if (e->radial_mask & (1u << i)) err[i] = fmod(err[i], e->radial_mod[i]);
        /* applied to EVERY element, sequentially (desired_state.py:152-153) */
        for (int i = 0; i < ds; i++)
            if (err[i] > M_PI) err[i] = 2 * M_PI - err[i];
        for (int i = 0; i < ds; i++)
            if (err[i] < -M_PI) err[i] = -2 * M_PI - err[i];
    }
    for (int i = 0; i < da; i++) ea[i] = -act[i];
    double cost = quad_form(e->Q, err, ds) + quad_form(e->R, ea, da);
    switch (e->rew_kind) {
    case ORC_REW_QUADR: return -cost;
    case ORC_REW_EXP: return exp(-cost);
    default: return exp(-e->c_max * cost);
    }
}

/* BoxSpace.contains, spaces/box.py:138-167 (inclusive bounds; NaN is an error -> reported as -1) */
static int orc_contains(const double *x, const double *lo, const double *up, int n) {
    for (int i = 0; i < n; i++)
        if (isnan(x[i])) return -1;
    for (int i = 0; i < n; i++)
        if (!(x[i] >= lo[i] && x[i] <= up[i])) return 0;
    return 1;
}

/* FinalRewTask.compute_final_rew, tasks/final_reward.py:111-275, with success_fcn = never_succeeded
 * (tasks/utils.py:32-34): only the has_failed branches can fire. */
static double orc_final_rew(orc_env *e, const double *state, long remaining_steps) {
    if (!(e->final_flags & ORC_FR_ENABLED)) return 0.0;
    if (e->yielded_final) return 0.0;
    e->yielded_final = 1;
    int failed = orc_contains(state, e->st_lo, e->st_up, k_ds[e->kind]) != 1;
    if (!failed) return 0.0;
    int sd = !!(e->final_flags & ORC_FR_STATE_DEP), td = !!(e->final_flags & ORC_FR_TIME_DEP);
    int ap = !!(e->final_flags & ORC_FR_ALWAYS_POS), an = !!(e->final_flags & ORC_FR_ALWAYS_NEG);
    if (ap) return 0.0;
    (void)an;
    double scale = td ? (double)remaining_steps : e->final_factor;
    if (sd) {
        double zero_act[ORC_MAX_DA] = {0};
        return -1.0 * scale * fabs(orc_step_rew(e, state, zero_act));
    }
    return -1.0 * scale;
}

/* ---------------------------------------------------------------------------------------------
 * Dynamics
 * ------------------------------------------------------------------------------------------- */
/* quanser_qube.py:89-125 */
static void qq_dyn(const orc_env *e, const double *x, double u, double *out) {
    double km = e->dp[2], Rm = e->dp[1], Dr = e->dp[5], Dp = e->dp[8];
    const double *c = e->c;
    double thd = x[2], ald = x[3], al = x[1];
    double sin_al = sin(al), sin_2al = sin(2 * al);
    double a = c[0] + c[1] * (sin_al * sin_al);
    double b = c[2] * cos(al);
    double cc = c[3];
    double det = a * cc - b * b;
    double trq = km * (u - km * thd) / Rm;
    double c0 = c[1] * sin_2al * thd * ald - c[2] * sin_al * ald * ald;
    double c1 = -0.5 * c[1] * sin_2al * thd * thd + c[4] * sin_al;
    double xx = trq - Dr * thd - c0;
    double yy = -Dp * ald - c1;
    out[0] = thd; out[1] = ald;
    out[2] = (cc * xx - b * yy) / det;
    out[3] = (a * yy - b * xx) / det;
}

/* quanser_qube.py:127-146 -- the fork's pseudo-RK4: _dyn is always evaluated at self.state */
static void qq_step(orc_env *e, double act) {
    if (e->dp[9] <= act && act <= e->dp[10]) act = 0.0;
    double k[4][4], d[4], s[4];
    qq_dyn(e, e->state, act, d);
    for (int i = 0; i < 4; i++) k[0][i] = d[i];
    for (int j = 1; j < 4; j++) {
        double h = (j <= 2) ? e->dt / 2.0 : e->dt;
        for (int i = 0; i < 4; i++) s[i] = e->state[i] + h * k[j - 1][i];
        qq_dyn(e, e->state, act, d);
        k[j][0] = s[2]; k[j][1] = s[3]; k[j][2] = d[2]; k[j][3] = d[3];
    }
    for (int i = 0; i < 4; i++) e->state[i] += e->dt / 6 * (k[0][i] + 2 * k[1][i] + 2 * k[2][i] + k[3][i]);
}

/* quanser_cartpole.py:166-230 (_dynamics, numpy) == :360-431 (_dynamics_diff, torch) */
static void qcp_dynamics(const orc_env *e, const double *sa, double th_ddot_in, double *deriv, double *th_ddot_out) {
    const double *p = e->dp;
    double g = p[0], m_c = p[1], eta_m = p[3], eta_g = p[4], K_g = p[5], r_mp = p[7], R_m = p[8], k_m = p[9],
           B_p = p[10], B_eq = p[11], m_p = p[12], l_p = p[13], mu_c = p[14];
    double J_pole = e->c[0], J_eq = e->c[1];
    int simple = e->flags & ORC_F_SIMPLE_DYN;
    double th = sa[1], x_dot = sa[2], th_dot = sa[3], u = sa[4];
    double sin_th = sin(th), cos_th = cos(th);
    double m_tot = m_c + m_p;
    if (!simple && p[15] <= u && u <= p[16]) u = 0.0;
    double f_act = (eta_g * K_g * eta_m * k_m) / (R_m * r_mp) * (eta_m * u - K_g * k_m * x_dot / r_mp);
    double f_tot;
    if (simple) {
        f_tot = f_act;
    } else {
        double f_normal = m_tot * g - m_p * l_p / 2 * (sin_th * th_ddot_in + cos_th * (th_dot * th_dot));
        double f_c;
        if (f_normal < 0) f_c = 0.0;
        else {
            double sgn = (x_dot > 0) - (x_dot < 0);
            f_c = mu_c * f_normal * sgn;
        }
        f_tot = f_act - f_c;
    }
    double M00 = m_p + J_eq, M01 = m_p * l_p * cos_th, M10 = M01, M11 = J_pole + m_p * (l_p * l_p);
    double r0 = f_tot - B_eq * x_dot - m_p * l_p * sin_th * (th_dot * th_dot);
    double r1 = -B_p * th_dot - m_p * l_p * g * sin_th;
    /* np.linalg.solve -> LAPACK dgesv: LU with partial pivoting (:224) */
    double x_ddot, th_ddot;
    if (fabs(M00) >= fabs(M10)) {
        double l = M10 / M00;
        double u11 = M11 - l * M01;
        double y1 = r1 - l * r0;
        th_ddot = y1 / u11;
        x_ddot = (r0 - M01 * th_ddot) / M00;
    } else {
        double l = M00 / M10;
        double u11 = M01 - l * M11;
        double y1 = r0 - l * r1;
        th_ddot = y1 / u11;
        x_ddot = (r1 - M11 * th_ddot) / M10;
    }
    /* "symplectic Euler" inside the derivative (:226-230) */
    double th_dot_n = th_dot + th_ddot * e->dt;
    double x_dot_n = x_dot + x_ddot * e->dt;
    deriv[0] = x_dot_n; deriv[1] = th_dot_n; deriv[2] = x_ddot; deriv[3] = th_ddot; deriv[4] = u * 0;
    *th_ddot_out = th_ddot;
}

/* rk4(): quanser_cartpole.py:591-655, t = [0, dt]; _step_dynamics :157-164 */
static void qcp_step(orc_env *e, double u) {
    double y0[5] = {e->state[0], e->state[1], e->state[2], e->state[3], u};
    double dt = e->dt - 0.0, dt2 = dt / 2.0;
    double k1[5], k2[5], k3[5], k4[5], y[5], t1, t2, t3, t4;
    qcp_dynamics(e, y0, e->hidden[0], k1, &t1);
    for (int i = 0; i < 5; i++) y[i] = y0[i] + dt2 * k1[i];
    qcp_dynamics(e, y, t1, k2, &t2);
    for (int i = 0; i < 5; i++) y[i] = y0[i] + dt2 * k2[i];
    qcp_dynamics(e, y, t2, k3, &t3);
    for (int i = 0; i < 5; i++) y[i] = y0[i] + dt * k3[i];
    qcp_dynamics(e, y, t3, k4, &t4);
    for (int i = 0; i < 4; i++) e->state[i] = y0[i] + dt / 6.0 * (k1[i] + 2 * k2[i] + 2 * k3[i] + k4[i]);
    e->hidden[0] = (t1 + t2 + t3 + t4) / 4;
}

/* quanser_ball_balancer.py:247-331; act is modified in place by the dead zone (:261-264) */
static void qbb_step(orc_env *e, double *act) {
    const double *p = e->dp;
    double g = p[0], m_ball = p[1], r_ball = p[2], ball_damping = p[13];
    double Vxp = p[14], Vxn = p[15], Vyp = p[16], Vyn = p[17], off_x = p[18], off_y = p[19];
    double J_ball = e->c[0], J_eq = e->c[1], c_kin = e->c[2], A_m = e->c[3], B_eq_v = e->c[4], zeta = e->c[5];
    int simple = e->flags & ORC_F_SIMPLE_DYN;
    double *s = e->state;
    if (!simple && Vxn <= act[0] && act[0] <= Vxp) act[0] = 0;
    if (!simple && Vyn <= act[1] && act[1] <= Vyp) act[1] = 0;
    double th_x = s[0] + off_x, th_y = s[1] + off_y, x = s[2], y = s[3];
    double th_x_dot = s[4], th_y_dot = s[5], x_dot = s[6], y_dot = s[7];
    double th_x_ddot = (A_m * act[0] - B_eq_v * th_x_dot) / J_eq;
    double th_y_ddot = (A_m * act[1] - B_eq_v * th_y_dot) / J_eq;
    double a = e->hidden[0], b = e->hidden[1];
    double a_dot = c_kin * th_x_dot * cos(th_x) / cos(a);
    double b_dot = c_kin * -th_y_dot * cos(-th_y) / cos(b);
    double a_ddot = 1.0 / cos(a) * (c_kin * (th_x_ddot * cos(th_x) - (th_x_dot * th_x_dot) * sin(th_x)) + (a_dot * a_dot) * sin(a));
    double b_ddot = 1.0 / cos(b) * (c_kin * (-th_y_ddot * cos(th_y) - ((-th_y_dot) * (-th_y_dot)) * sin(-th_y)) + (b_dot * b_dot) * sin(b));
    double x_ddot, y_ddot;
    double r2 = r_ball * r_ball;
    if (simple) {
        x_ddot = c_kin * m_ball * g * r2 * sin(th_x) / zeta;
        y_ddot = c_kin * m_ball * g * r2 * sin(th_y) / zeta;
    } else {
        x_ddot = (-ball_damping * x_dot * r2 - J_ball * r_ball * a_ddot + m_ball * x * (a_dot * a_dot) * r2 +
                  c_kin * m_ball * g * r2 * sin(th_x)) / zeta;
        y_ddot = (-ball_damping * y_dot * r2 - J_ball * r_ball * b_ddot + m_ball * y * ((-b_dot) * (-b_dot)) * r2 +
                  c_kin * m_ball * g * r2 * sin(th_y)) / zeta;
    }
    s[4] += th_x_ddot * e->dt; s[5] += th_y_ddot * e->dt; s[6] += x_ddot * e->dt; s[7] += y_ddot * e->dt;
    for (int i = 0; i < 4; i++) s[i] += s[4 + i] * e->dt;
    e->hidden[0] += a_dot * e->dt; e->hidden[1] += b_dot * e->dt;
}

/* ball_on_beam.py:110-130 */
static void bob_step(orc_env *e, double act) {
    const double *p = e->dp;
    double g = p[0], m_ball = p[1], c_frict = p[6], ang_offset = p[7];
    double J_beam = e->c[1], zeta_ball = e->c[2];
    double *s = e->state;
    double x = s[0], a = s[1] + ang_offset, x_dot = s[2], a_dot = s[3];
    double zeta_beam = m_ball * (x * x) + J_beam;
    double x_ddot = (-c_frict * x_dot + m_ball * x * (a_dot * a_dot) - m_ball * g * sin(a)) / zeta_ball;
    double a_ddot = (act - 2.0 * m_ball * x * x_dot * a_dot - m_ball * g * cos(a) * x) / zeta_beam;
    s[2] += x_ddot * e->dt; s[3] += a_ddot * e->dt;
    s[0] += s[2] * e->dt; s[1] += s[3] * e->dt;
}

/* one_mass_oscillator.py:105-116 */
static void omo_step(orc_env *e, double act) {
    double m = e->dp[0], omega = e->c[0], zeta = e->c[1];
    double A10 = -(omega * omega), A11 = -2.0 * zeta * omega;
    double *s = e->state;
    double sd0 = 0 * s[0] + 1 * s[1] + 0 * act;
    double sd1 = A10 * s[0] + A11 * s[1] + (1.0 / m) * act;
    s[0] = s[0] + sd0 * e->dt;
    s[1] = s[1] + sd1 * e->dt;
}

/* pendulum.py:106-122: rod pendulum, symplectic Euler (velocity first) */
static void pend_step(orc_env *e, double act) {
    double g = e->dp[0], m = e->dp[1], l = e->dp[2], d = e->dp[3];
    double *s = e->state;
    double th = s[0], th_dot = s[1];
    double th_ddot = (act - m * g * l / 2.0 * sin(th) - d * th_dot) / (m * (l * l) / 3.0);
    s[1] += th_ddot * e->dt;
    s[0] += s[1] * e->dt;
}

/* Env.limit_act -> BoxSpace.project_to, environments/base.py:215-222, spaces/box.py:180-184 */
void orc_limit_act(const orc_env *e, const double *act, double *out) {
    if (e->flags & ORC_F_DISC_ACT) {
        /* DiscreteSpace.project_to (spaces/discrete.py:119-127) over eles = linspace(lo, up, 3): an action that
         * np.isclose()s an element (|ele - a| <= 1e-8 + 1e-5 |a|) passes unchanged, else the closest element (first
         * one on ties, np.argmin) */
        double lo = e->act_lo[0], up = e->act_up[0], a = act[0];
        double eles[3] = {lo, lo + (up - lo) / 2.0, up};
        int close = 0, best = 0;
        for (int k = 0; k < 3; k++) {
            if (fabs(eles[k] - a) <= 1e-8 + 1e-5 * fabs(a)) close = 1;
            if (fabs(a - eles[k]) < fabs(a - eles[best])) best = k;
        }
        out[0] = close ? a : eles[best];
        return;
    }
    for (int i = 0; i < k_da[e->kind]; i++) out[i] = fmin(fmax(act[i], e->act_lo[i]), e->act_up[i]);
}

/* ---------------------------------------------------------------------------------------------
 * SimPyEnv.step: environments/pysim/base.py:217-241.  Returns done (0/1) or -1 if a NaN state was
 * met (the reference raises ValueErr in BoxSpace.contains, spaces/box.py:142).
 * ------------------------------------------------------------------------------------------- */
int orc_step(orc_env *e, const double *act, double *obs, double *rew) {
    long remaining = e->max_steps > 0 ? e->max_steps - (e->curr_step + 1) : 0;
    double r = orc_step_rew(e, e->state, act);
    double a[ORC_MAX_DA];
    orc_limit_act(e, act, a);
    switch (e->kind) {
    case ORC_QQ: qq_step(e, a[0]); break;
    case ORC_QCP: qcp_step(e, a[0]); break;
    case ORC_QBB: qbb_step(e, a); break;
    case ORC_BOB: bob_step(e, a[0]); break;
    case ORC_OMO: omo_step(e, a[0]); break;
    case ORC_PEND: pend_step(e, a[0]); break;
    }
    e->curr_step += 1;
    int inside = orc_contains(e->state, e->st_lo, e->st_up, k_ds[e->kind]);
    int done = (inside != 1); /* never_succeeded -> done == has_failed (tasks/base.py:159-172) */
    if (e->max_steps > 0 && e->curr_step >= e->max_steps) done = 1;
    if (done) r += orc_final_rew(e, e->state, remaining);
    if (obs) orc_observe(e, e->state, obs);
    *rew = r;
    return inside < 0 ? -1 : done;
}

/* =============================================================================================
 * Differentiable cart-pole step (value only): quanser_cartpole.py:257-358 with state_based=True.
 * act/act_diff already de-normalised by the wrappers (environment_wrappers/base.py:324-327).
 * Differences to orc_step that are reproduced: reward wrap (:295-321), STRICT done test (:291-293),
 * no step counter, no final reward, th_ddot passed explicitly.
 * ============================================================================================= */
int orc_qcp_step_diff_state(orc_env *e, const double *state, double act, double act_diff, double th_ddot_in,
                            double *next_state, double *rew, double *th_ddot_out) {
    double act_unclamped = act + act_diff;
    double a = fmin(fmax(act + act_diff, e->act_lo[0]), e->act_up[0]);
    /* rew_fn */
    double err[4];
    for (int i = 0; i < 4; i++) err[i] = e->state_des[i] - state[i];
    double two_pi = 2 * M_PI;
    double theta = err[1] + (-(floor(err[1] / two_pi)) * two_pi + (err[1] < 0 ? 1.0 : 0.0) * two_pi);
    if (theta > M_PI) theta = theta - two_pi;
    err[1] = theta;
    double cost = quad_form(e->Q, err, 4) + (-act_unclamped) * e->R[0] * (-act_unclamped);
    *rew = exp(-cost);
    memcpy(e->state, state, sizeof(double) * 4);
    e->hidden[0] = th_ddot_in;
    qcp_step(e, a);
    memcpy(next_state, e->state, sizeof(double) * 4);
    *th_ddot_out = e->hidden[0];
    int inside = 1;
    for (int i = 0; i < 4; i++) inside = inside && (e->st_lo[i] < next_state[i]) && (e->st_up[i] > next_state[i]);
    return !inside;
}

/* =============================================================================================
 * Counter-based RNG: Philox4x32-10 (Salmon, Moraes, Dror, Shaw 2011; Random123 philox.h constants
 * M0=0xD2511F53, M1=0xCD9E8D57, W0=0x9E3779B9, W1=0xBB67AE85).
 * ============================================================================================= */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
        uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
        uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

/* Engine RNG addressing: key = (seed_lo, seed_hi); counter = (env_lo, env_hi, episode, stream<<24 | block). */
#define ORC_STREAM_PARAMS 0u
#define ORC_STREAM_INIT 1u
#define ORC_STREAM_OBS_NOISE 2u
#define ORC_STREAM_ACT_NOISE 3u
#define ORC_STREAM_ENV_ACT_NOISE 4u

static void orc_rng(uint64_t seed, uint64_t env, uint32_t episode, uint32_t stream, uint32_t block, uint32_t out[4]) {
    uint32_t ctr[4] = {(uint32_t)env, (uint32_t)(env >> 32), episode, (stream << 24) | (block & 0xFFFFFFu)};
    uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    orc_philox4x32_10(ctr, key, out);
}
void orc_rng_block(uint64_t seed, uint64_t env, uint32_t episode, uint32_t stream, uint32_t block, uint32_t *out) {
    orc_rng(seed, env, episode, stream, block, out);
}

static inline double u53(uint32_t a, uint32_t b) {
    /* numpy-style 53-bit uniform in [0,1) */
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) * (1.0 / 9007199254740992.0);
}
static inline float u24(uint32_t a) { return (float)(a >> 8) * (1.0f / 16777216.0f); }
static inline double normal_d(uint32_t a, uint32_t b) {
    /* Box-Muller, engine RNG specification: u1 = (2 (a >> 9) + 1) / 2^24 in (0,1), x = ((int32) b >> 8) / 2^23 in
     * [-1,1), z = sqrt(-2 ln u1) cos(pi x).  Evaluated in fp64 here; the engine evaluates the same expression in
     * float32 on the (exactly representable) inputs. */
    double u1 = (double)(2u * (a >> 9) + 1u) * (1.0 / 16777216.0);
    double x = (double)((int32_t)b >> 8) * (1.0 / 8388608.0);
    return sqrt(-2.0 * log(u1)) * cos(M_PI * x);
}

/* Distribution table entry (engine spec, mirrors DomainParam fields: domain_parameter.py:46-76,
 * Uniform :135-169, Normal :171-205, Bernoulli :246-312).  Samples are float32 like the reference's
 * torch.distributions draws (torch default dtype), then clamp (:126) and optional round (:129-130). */
typedef struct {
    int32_t param_index;
    int32_t kind; /* 0 normal(a=mean,b=std) 1 uniform(a=lo,b=hi) 2 bernoulli(a=val_0,b=val_1,c=prob_1) */
    float a, b, c;
    float clip_lo, clip_up;
    int32_t roundint;
} orc_distr;

float orc_sample_entry(const orc_distr *d, uint64_t seed, uint64_t env, uint32_t episode, uint32_t entry) {
    uint32_t x[4];
    orc_rng(seed, env, episode, ORC_STREAM_PARAMS, entry, x);
    float s;
    if (d->kind == 0) {
        float z = (float)normal_d(x[0], x[1]);
        float t = d->b * z;      /* separate roundings: loc + scale * eps */
        s = d->a + t;
    } else if (d->kind == 1) {
        float u = u24(x[0]);
        float t = u * (d->b - d->a); /* low + rand * (high - low) */
        s = d->a + t;
    } else {
        float u = u24(x[0]);
        float bern = (u < d->c) ? 1.0f : 0.0f;
        s = (1.0f - bern) * d->a + bern * d->b;
    }
    s = fminf(fmaxf(s, d->clip_lo), d->clip_up);
    if (d->roundint) s = rintf(s);
    return s;
}

/* DomainRandomizer.randomize(1) + get_params + domain_param setter for one (env, episode):
 * domain_randomizer.py:123-157, environment_wrappers/domain_randomization.py:141-148. */
void orc_sample_params(orc_env *e, const orc_distr *table, int n_entries, uint64_t seed, uint64_t env, uint32_t episode) {
    for (int i = 0; i < n_entries; i++)
        e->dp[table[i].param_index] = (double)orc_sample_entry(&table[i], seed, env, episode, (uint32_t)i);
    orc_derive(e);
}

/* init_space.sample_uniform(): spaces/box.py:169-178, spaces/polar.py:108-113, spaces/compound.py:83-87.
 * Default init spaces: QQ quanser_qube.py:169,177; QCP quanser_cartpole.py:552-560 (wild_init selects
 * the box: 0 -> 'True', 1 -> 'False', 2 -> else); QBB quanser_ball_balancer.py:108-117;
 * BOB ball_on_beam.py:60-73; OMO one_mass_oscillator.py:57-58.
 * If box_lo/box_up are given they override the default space (init_space setter, base.py:99-103). */
void orc_sample_init(const orc_env *e, int wild_init, const double *box_lo, const double *box_up, uint64_t seed,
                     uint64_t env, uint32_t episode, double *init, int *n_init) {
    uint32_t x[4], y[4];
    orc_rng(seed, env, episode, ORC_STREAM_INIT, 0, x);
    orc_rng(seed, env, episode, ORC_STREAM_INIT, 1, y);
    double u[4] = {u53(x[0], x[1]), u53(x[2], x[3]), u53(y[0], y[1]), u53(y[2], y[3])};
    double lo[4], up[4];
    int n = k_dinit[e->kind];
    *n_init = n;
    const double pi = M_PI;
    if (box_lo && box_up) {
        for (int i = 0; i < n; i++) { lo[i] = box_lo[i]; up[i] = box_up[i]; }
    } else {
        switch (e->kind) {
        case ORC_QQ: {
            const double m[4] = {2.0, 1.0, 0.5, 0.5};
            for (int i = 0; i < 4; i++) { up[i] = m[i] / 180 * pi; lo[i] = -up[i]; }
        } break;
        case ORC_QCP: {
            double m[4];
            if (wild_init == 0) { m[0] = 0.25; m[1] = pi; m[2] = 0.8; m[3] = pi; }
            else if (wild_init == 1) { m[0] = 0.02; m[1] = 2 / 180.0 * pi; m[2] = 0.0; m[3] = 1 / 180.0 * pi; }
            else { m[0] = 0.02; m[1] = pi; m[2] = 0.0; m[3] = 1 / 180.0 * pi; }
            for (int i = 0; i < 4; i++) { up[i] = m[i]; lo[i] = -m[i]; }
        } break;
        case ORC_QBB: {
            double l_plate = e->dp[3];
            lo[0] = 0.75 * l_plate / 2; lo[1] = -pi; lo[2] = -0.05 * 0.5; lo[3] = -0.05 * 0.5;
            up[0] = 0.8 * l_plate / 2; up[1] = pi; up[2] = 0.05 * 0.5; up[3] = 0.05 * 0.5;
        } break;
        case ORC_BOB: {
            double l_beam = e->dp[4];
            /* CompoundSpace of two boxes: pick one (np.random.randint(2)) from a separate RNG block */
            uint32_t z[4];
            orc_rng(seed, env, episode, ORC_STREAM_INIT, 2, z);
            int right = (int)(z[0] >> 31);
            if (!right) { lo[0] = -0.8 * l_beam / 2.0; up[0] = -0.7 * l_beam / 2.0; }
            else { lo[0] = 0.7 * l_beam / 2.0; up[0] = 0.8 * l_beam / 2.0; }
            lo[1] = -5 / 180.0 * pi; up[1] = +5 / 180.0 * pi;
            lo[2] = -0.02 * 10.0; up[2] = +0.02 * 10.0;
            lo[3] = -0.02 * pi; up[3] = +0.02 * pi;
        } break;
        case ORC_OMO:
            lo[0] = -0.75 * 1.0; up[0] = -0.65 * 1.0; lo[1] = -0.01 * 10.0; up[1] = +0.01 * 10.0;
            break;
        case ORC_PEND: /* SingularStateSpace(init_state = zeros), pendulum.py:75 */
            lo[0] = up[0] = 0.0; lo[1] = up[1] = 0.0;
            break;
        }
    }
    for (int i = 0; i < n; i++) init[i] = lo[i] + (up[i] - lo[i]) * u[i];
    if (e->kind == ORC_QBB) {
        /* Polar2DPosVelSpace.sample_uniform, spaces/polar.py:108-113 */
        double r = init[0], phi = init[1];
        init[0] = r * cos(phi); init[1] = r * sin(phi);
    }
}

/* =============================================================================================
 * Policies (float32, like torch default dtype).
 *   FNN: policies/feed_back/fnn.py:138-160 -- Linear -> tanh per hidden layer, Linear output.
 *   Linear: policies/feed_back/linear.py:95-110 + elementwise features policies/features.py:127-175.
 * Weights: torch layout W[out][in] row-major, bias[out]; layers concatenated W0,b0,W1,b1,...
 * ============================================================================================= */
#define ORC_POL_ACTIONS 0 /* injected actions */
#define ORC_POL_FNN 1
#define ORC_POL_LINEAR 2
#define ORC_POL_IDLE 3

/* feature kinds for the Linear policy */
enum { F_CONST = 0, F_IDENT, F_SIGN, F_ABS, F_SQUARED, F_CUBIC, F_SIN, F_COS, F_SINSIN, F_SINCOS };

typedef struct {
    int kind;
    int n_layers;          /* FNN: number of Linear layers (hidden + output) */
    int sizes[9];          /* FNN: in, h1, ..., out */
    int n_feats;           /* Linear: number of feature functions */
    int feats[12];
    int hidden_nonlin;     /* 0 tanh, 1 relu, 2 identity */
    int output_nonlin;     /* 0 none, 1 tanh */
    int cartpole_feats;    /* fork's FNNPolicy input transform, fnn.py:214-218 */
    const float *weights;
    /* StochasticActionExplStrat.forward (exploration/stochastic_action.py:68-94): 0 none, 1 NormalActNoiseExplStrat
     * (act_noise_scale = std, normal_noise.py:155-163), 2 UniformActNoiseExplStrat (scale = halfspan,
     * uniform_noise.py:160-168); act_noise_mean = the learnable noise mean (train_mean) */
    int act_noise;
    float act_noise_scale[2];
    float act_noise_mean[2];
    /* parameter exploration (sampling/parameter_exploration_sampler.py:267-345): lane n evaluates parameter set
     * n / group_size of the n_weights-float sets stored back to back in `weights` (0: one set for all) */
    long group_size, n_weights;
} orc_policy;

static float nonlin_f(int kind, float x) {
    switch (kind) {
    case 0: return tanhf(x);
    case 1: return x > 0.f ? x : 0.f;
    default: return x;
    }
}

void orc_policy_forward(const orc_policy *pol, const float *obs, int n_obs, float *act, int n_act) {
    if (pol->kind == ORC_POL_IDLE) { for (int i = 0; i < n_act; i++) act[i] = 0.f; return; }
    if (pol->kind == ORC_POL_FNN) {
        float bufa[256], bufb[256];
        const float *in = obs; int n_in = n_obs;
        float feat[16];
        if (pol->cartpole_feats) {
            feat[0] = obs[0]; feat[1] = sinf(obs[1]); feat[2] = cosf(obs[1]); feat[3] = obs[2]; feat[4] = obs[3];
            in = feat; n_in = 5;
        }
        const float *w = pol->weights;
        float *cur = bufa, *nxt = bufb;
        for (int i = 0; i < n_in; i++) cur[i] = in[i];
        for (int L = 0; L < pol->n_layers; L++) {
            int ni = pol->sizes[L], no = pol->sizes[L + 1];
            const float *W = w, *b = w + (size_t)ni * no;
            for (int o = 0; o < no; o++) {
                float s = b[o];
                for (int i = 0; i < ni; i++) s += W[o * ni + i] * cur[i];
                if (L < pol->n_layers - 1) s = nonlin_f(pol->hidden_nonlin, s);
                else if (pol->output_nonlin == 1) s = tanhf(s);
                nxt[o] = s;
            }
            w += (size_t)ni * no + no;
            float *t = cur; cur = nxt; nxt = t;
        }
        for (int i = 0; i < n_act; i++) act[i] = cur[i];
        return;
    }
    if (pol->kind == ORC_POL_LINEAR) {
        float f[128]; int nf = 0;
        for (int k = 0; k < pol->n_feats; k++) {
            if (pol->feats[k] == F_CONST) { f[nf++] = 1.0f; continue; }
            for (int i = 0; i < n_obs; i++) {
                float x = obs[i], v;
                switch (pol->feats[k]) {
                case F_IDENT: v = x; break;
                case F_SIGN: v = (float)((x > 0) - (x < 0)); break;
                case F_ABS: v = fabsf(x); break;
                case F_SQUARED: v = x * x; break;
                case F_CUBIC: v = x * x * x; break;
                case F_SIN: v = sinf(x); break;
                case F_COS: v = cosf(x); break;
                case F_SINSIN: v = sinf(x) * sinf(x); break;
                default: v = sinf(x) * cosf(x); break;
                }
                f[nf++] = v;
            }
        }
        for (int o = 0; o < n_act; o++) {
            float s = 0.f;
            for (int i = 0; i < nf; i++) s += pol->weights[o * nf + i] * f[i];
            act[o] = s;
        }
    }
}

/* =============================================================================================
 * Batched rollout == rollout() per env (sampling/rollout.py:63-343) for N independent envs, as the
 * ParallelRolloutSampler would run them (sampling/parallel_rollout_sampler.py:248-323):
 *   per env: [DR sample] -> reset(init) -> loop { obs -> policy (float32) -> [ActNorm] -> step }.
 * Observation noise (GaussianObsNoiseWrapper, environment_wrappers/observation_noise.py:68-73) is
 * applied to the observation the policy sees and that is recorded, at reset and at every step.
 * Outputs (any pointer may be NULL): trajectories in [T, N, d] layout (float64), returns, lengths.
 * ============================================================================================= */
typedef struct {
    int kind, flags, long_pole, wild_init;
    double dt;
    long max_steps;
    /* task override (if has_task) */
    int has_task;
    double state_des[ORC_MAX_DS], Q[ORC_MAX_DS * ORC_MAX_DS], R[ORC_MAX_DA * ORC_MAX_DA];
    int rew_kind; double min_rew; unsigned radial_mask; int final_flags; double final_factor;
    /* wrappers */
    int act_norm;              /* ActNormWrapper, action_normalization.py:69-75 */
    int obs_noise;             /* GaussianObsNoiseWrapper */
    double noise_std[ORC_MAX_DO], noise_mean[ORC_MAX_DO];
    /* sources */
    const double *params;      /* [N, dp] or NULL (nominal / DR) */
    const orc_distr *distr; int n_distr; /* DR table or NULL */
    const double *init;        /* [N, n_init_cols] or NULL -> sample */
    int n_init_cols;
    const double *init_box_lo, *init_box_up;
    const double *actions;     /* [H, N, da] when policy.kind == ACTIONS */
    uint64_t seed; uint64_t env_offset; uint32_t episode;
    int stop_on_done;
    /* ObsNormWrapper (observation_normalization.py:119-122, after the noise) and GaussianActNoiseWrapper
     * (action_noise.py:73-78; 1 = outside ActNormWrapper, 2 = inside) */
    int obs_norm, env_act_noise;
    double obs_lb[ORC_MAX_DO], obs_ub[ORC_MAX_DO];
    double an_std[ORC_MAX_DA], an_mean[ORC_MAX_DA];
} orc_rollout_cfg;

typedef struct {
    double *obs;      /* [H+1, N, do] */
    double *act;      /* [H, N, da]   policy output */
    double *act_app;  /* [H, N, da]   clipped, inner-env units */
    double *rew;      /* [H, N] */
    double *state;    /* [H+1, N, ds] */
    double *hidden;   /* [H+1, N, nh] */
    uint8_t *done;    /* [H, N] */
    double *ret;      /* [N] undiscounted return */
    int32_t *length;  /* [N] */
    double *params_out; /* [N, dp] */
} orc_rollout_out;

long orc_rollout_batch(const orc_rollout_cfg *cfg, const orc_policy *pol, long N, long H, orc_rollout_out *out,
                       int n_threads) {
    int kind = cfg->kind;
    int ds = k_ds[kind], dob = k_do[kind], da = k_da[kind], dp = k_dp[kind], nh = k_nh[kind];
    long total_steps = 0;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#pragma omp parallel for schedule(static) reduction(+ : total_steps)
#endif
    for (long n = 0; n < N; n++) {
        orc_env e;
        orc_init(&e, kind, cfg->dt, cfg->max_steps, cfg->flags, cfg->long_pole);
        if (cfg->has_task)
            orc_set_task(&e, cfg->state_des, cfg->Q, cfg->R, cfg->rew_kind, cfg->min_rew, cfg->radial_mask,
                         cfg->final_flags, cfg->final_factor);
        uint64_t genv = cfg->env_offset + (uint64_t)n;
        if (cfg->params) orc_set_params(&e, cfg->params + n * dp);
        if (cfg->distr) orc_sample_params(&e, cfg->distr, cfg->n_distr, cfg->seed, genv, cfg->episode);
        if (out->params_out) memcpy(out->params_out + n * dp, e.dp, sizeof(double) * dp);
        double init[ORC_MAX_DS]; int n_init;
        if (cfg->init) { n_init = cfg->n_init_cols; memcpy(init, cfg->init + n * n_init, sizeof(double) * n_init); }
        else orc_sample_init(&e, cfg->wild_init, cfg->init_box_lo, cfg->init_box_up, cfg->seed, genv, cfg->episode, init, &n_init);
        double obs[ORC_MAX_DO];
        orc_reset(&e, init, n_init, obs);
        double ret = 0.0; long t = 0; int done = 0;
        for (; t < H; t++) {
            if (cfg->obs_noise) {
                for (int i = 0; i < dob; i += 4) {
                    uint32_t x[4];
                    orc_rng(cfg->seed, genv, cfg->episode, ORC_STREAM_OBS_NOISE, (uint32_t)(t * 2 + i / 4), x);
                    /* 4 normals per block via two Box-Muller pairs (float32 math on the engine, fp64 here) */
                    double u1 = ((double)x[0] + 1.0) / 4294967296.0, u2 = (double)x[1] / 4294967296.0;
                    double u3 = ((double)x[2] + 1.0) / 4294967296.0, u4 = (double)x[3] / 4294967296.0;
                    double z[4] = {sqrt(-2 * log(u1)) * cos(2 * M_PI * u2), sqrt(-2 * log(u1)) * sin(2 * M_PI * u2),
                                   sqrt(-2 * log(u3)) * cos(2 * M_PI * u4), sqrt(-2 * log(u3)) * sin(2 * M_PI * u4)};
                    for (int j = 0; j < 4 && i + j < dob; j++) obs[i + j] += z[j] * cfg->noise_std[i + j] + cfg->noise_mean[i + j];
                }
            }
            if (cfg->obs_norm)
                for (int i = 0; i < dob; i++) obs[i] = (obs[i] - cfg->obs_lb[i]) / (cfg->obs_ub[i] - cfg->obs_lb[i]) * 2 - 1;
            if (out->obs) memcpy(out->obs + (t * N + n) * dob, obs, sizeof(double) * dob);
            if (out->state) memcpy(out->state + (t * N + n) * ds, e.state, sizeof(double) * ds);
            if (out->hidden && nh) memcpy(out->hidden + (t * N + n) * nh, e.hidden, sizeof(double) * nh);
            double act[ORC_MAX_DA];
            if (pol->kind == ORC_POL_ACTIONS) {
                for (int i = 0; i < da; i++) act[i] = cfg->actions[(t * N + n) * da + i];
            } else {
                float of[ORC_MAX_DO], af[ORC_MAX_DA];
                for (int i = 0; i < dob; i++) of[i] = (float)obs[i];
                if (pol->group_size > 0) {
                    orc_policy set = *pol;
                    set.weights = pol->weights + (n / pol->group_size) * pol->n_weights;
                    orc_policy_forward(&set, of, dob, af, da);
                } else {
                    orc_policy_forward(pol, of, dob, af, da);
                }
                if (pol->act_noise) {
                    /* act = dist(policy(obs) + mean).rsample(), float32 with torch's roundings; one Philox block per
                     * step (engine RNG specification): normal pair from words 0,1 with the exactly representable
                     * inputs of normal_d (evaluated in fp64 here, float32 on the engine); uniforms from words 0,1 */
                    uint32_t x[4];
                    orc_rng(cfg->seed, genv, cfg->episode, ORC_STREAM_ACT_NOISE, (uint32_t)t, x);
                    if (pol->act_noise == 1) {
                        double u1 = (double)(2u * (x[0] >> 9) + 1u) * (1.0 / 16777216.0);
                        double ang = (double)((int32_t)x[1] >> 8) * (1.0 / 8388608.0);
                        double r = sqrt(-2.0 * log(u1));
                        float z[2] = {(float)(r * cos(M_PI * ang)), (float)(r * sin(M_PI * ang))};
                        for (int i = 0; i < da; i++) {
                            float loc = af[i] + pol->act_noise_mean[i];
                            float sc = z[i] * pol->act_noise_scale[i];
                            af[i] = loc + sc;
                        }
                    } else {
                        float u[2] = {u24(x[0]), u24(x[1])};
                        for (int i = 0; i < da; i++) {
                            float loc = af[i] + pol->act_noise_mean[i];
                            float lo = loc - pol->act_noise_scale[i], hi = loc + pol->act_noise_scale[i];
                            float w = hi - lo;
                            float sc = u[i] * w;
                            af[i] = lo + sc;
                        }
                    }
                }
                for (int i = 0; i < da; i++) act[i] = (double)af[i];
            }
            double act_env[ORC_MAX_DA], an[ORC_MAX_DA] = {0};
            if (cfg->env_act_noise) {
                /* noise = randn * std + mean; act + noise (action_noise.py:73-78).  Engine RNG: one block per step,
                 * the exact-input Box-Muller pair of normal_d */
                uint32_t x[4];
                orc_rng(cfg->seed, genv, cfg->episode, ORC_STREAM_ENV_ACT_NOISE, (uint32_t)t, x);
                double u1 = (double)(2u * (x[0] >> 9) + 1u) * (1.0 / 16777216.0);
                double ang = (double)((int32_t)x[1] >> 8) * (1.0 / 8388608.0);
                double r = sqrt(-2.0 * log(u1));
                double z[2] = {r * cos(M_PI * ang), r * sin(M_PI * ang)};
                for (int i = 0; i < da; i++) an[i] = z[i] * cfg->an_std[i] + cfg->an_mean[i];
            }
            for (int i = 0; i < da; i++) {
                double a = act[i];
                if (cfg->env_act_noise == 1) a = a + an[i];
                if (cfg->act_norm) a = e.act_lo[i] + (a + 1) * (e.act_up[i] - e.act_lo[i]) / 2;
                if (cfg->env_act_noise == 2) a = a + an[i];
                act_env[i] = a;
            }
            double rew;
            if (out->act) memcpy(out->act + (t * N + n) * da, act, sizeof(double) * da);
            if (out->act_app) {
                /* rollout.py:244 env.limit_act(act): through ActNormWrapper this clips the NORMALISED action
                 * to [-1, 1] (act_space of the wrapper) */
                for (int i = 0; i < da; i++) {
                    double lo = cfg->act_norm ? -1.0 : e.act_lo[i], up = cfg->act_norm ? 1.0 : e.act_up[i];
                    out->act_app[(t * N + n) * da + i] = fmin(fmax(act[i], lo), up);
                }
                if ((e.flags & ORC_F_DISC_ACT) && !cfg->act_norm) orc_limit_act(&e, act, out->act_app + (t * N + n) * da);
            }
            int d = orc_step(&e, act_env, obs, &rew);
            done = (d != 0);
            ret += rew;
            if (out->rew) out->rew[t * N + n] = rew;
            if (out->done) out->done[t * N + n] = (uint8_t)done;
            if (done && cfg->stop_on_done) { t++; break; }
        }
        /* final observation / state rows (rollout.py:309-311) */
        if (cfg->obs_noise) {
            for (int i = 0; i < dob; i += 4) {
                uint32_t x[4];
                orc_rng(cfg->seed, genv, cfg->episode, ORC_STREAM_OBS_NOISE, (uint32_t)(t * 2 + i / 4), x);
                double u1 = ((double)x[0] + 1.0) / 4294967296.0, u2 = (double)x[1] / 4294967296.0;
                double u3 = ((double)x[2] + 1.0) / 4294967296.0, u4 = (double)x[3] / 4294967296.0;
                double z[4] = {sqrt(-2 * log(u1)) * cos(2 * M_PI * u2), sqrt(-2 * log(u1)) * sin(2 * M_PI * u2),
                               sqrt(-2 * log(u3)) * cos(2 * M_PI * u4), sqrt(-2 * log(u3)) * sin(2 * M_PI * u4)};
                for (int j = 0; j < 4 && i + j < dob; j++) obs[i + j] += z[j] * cfg->noise_std[i + j] + cfg->noise_mean[i + j];
            }
        }
        if (cfg->obs_norm)
            for (int i = 0; i < dob; i++) obs[i] = (obs[i] - cfg->obs_lb[i]) / (cfg->obs_ub[i] - cfg->obs_lb[i]) * 2 - 1;
        if (out->obs) memcpy(out->obs + (t * N + n) * dob, obs, sizeof(double) * dob);
        if (out->state) memcpy(out->state + (t * N + n) * ds, e.state, sizeof(double) * ds);
        if (out->hidden && nh) memcpy(out->hidden + (t * N + n) * nh, e.hidden, sizeof(double) * nh);
        if (out->ret) out->ret[n] = ret;
        if (out->length) out->length[n] = (int32_t)t;
        total_steps += t;
    }
    return total_steps;
}

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* =============================================================================================
 * Post-processing of rollouts: discounted_reverse_cumsum (sampling/step_sequence.py:1049-1060),
 * discounted_value (:1063-1072) and gae_returns (:1093-1116), applied to every episode of every lane of a
 * time-major buffer.  Row flags as written by the engine: 0 step, 1 step with done, 2 closing row of a finished
 * episode, 3 closing row of a truncated episode, 4 idle.  The reference only knows complete rollouts; a truncated
 * episode (engine extension) bootstraps from the value estimate of its closing row.
 * ============================================================================================= */
void orc_returns(const double *rew, const uint8_t *flag, const double *val, double gamma, double lamb, double *out_ret,
                 double *out_adv, long T, long N) {
    for (long n = 0; n < N; n++) {
        for (long t = 0; t < T; t++) {
            out_ret[t * N + n] = 0.0;
            if (out_adv) out_adv[t * N + n] = 0.0;
        }
        long t = 0;
        while (t < T) {
            if (flag[t * N + n] > 1) { t++; continue; }
            /* episode = rows a .. b - 1 (steps), row b = closing row (or b == T) */
            long a = t, b = t;
            while (b < T && flag[b * N + n] <= 1) b++;
            int truncated = (b < T && flag[b * N + n] == 3);
            /* discounted_reverse_cumsum of the rewards: y[i] = x[i] + gamma y[i+1], y[end] = 0 (lfilter zero init) */
            double y = (truncated && val) ? val[b * N + n] : 0.0;
            for (long i = b - 1; i >= a; i--) {
                y = rew[i * N + n] + gamma * y;
                out_ret[i * N + n] = y;
            }
            if (val && out_adv) {
                /* deltas[i] = r_i + gamma * next_value_i - value_i with next_value = 0 for a done step (:1109-1113) */
                double c = 0.0;
                for (long i = b - 1; i >= a; i--) {
                    int done = flag[i * N + n] == 1;
                    double nv = done ? 0.0 : ((i + 1 < T) ? val[(i + 1) * N + n] : 0.0);
                    double delta = rew[i * N + n] + gamma * nv - val[i * N + n];
                    c = delta + gamma * lamb * c; /* discounted_reverse_cumsum(deltas, gamma * lamb) */
                    out_adv[i * N + n] = c;
                }
            }
            t = b;
        }
    }
}

