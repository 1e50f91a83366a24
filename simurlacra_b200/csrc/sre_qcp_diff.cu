// sre_qcp_diff.cu -- differentiable cart-pole step: forward over a horizon with a tape, and the hand-derived
// adjoint (reverse mode) through the RK4 step.  Replaces torch autograd through
//   QCartPoleSim.step_diff_state -> _step_dynamics_diff -> rk4(_dynamics_diff)   (quanser_cartpole.py:257-431, :591-655)
// as used by sac_wgrad (sac_wgrad/sac.py:97-110) and rollout_tensor_full (sampling/rollout.py:832-837).
// Gradient conventions reproduced from the torch graph of the reference:
//   * torch.clamp on the action: gradient 1 inside [lo, up] (inclusive), 0 outside (:264, :282-284)
//   * dead-zone mask, normal-force mask and sign(x_dot) are constants of the graph (:376-377, :394-396)
//   * the reward's angle wrap is detached: d theta_err / d th = -1 (:299-306)
//   * th_ddot carried from stage to stage and from step to step stays attached (:645-652)
#include <mutex>

#include "sre_host.h"

namespace sre {

// one state row (4 reals) as a single 128-bit access (float) or two (double); rows are 16-byte aligned because the
// tape is a contiguous [H+1, N, 4] buffer whose base the C-ABI checks
SRE_DEV void row4_store(float *p, const float *v) { *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]); }
SRE_DEV void row4_store(double *p, const double *v) {
    reinterpret_cast<double2 *>(p)[0] = make_double2(v[0], v[1]);
    reinterpret_cast<double2 *>(p)[1] = make_double2(v[2], v[3]);
}
SRE_DEV void row4_load(const float *p, float *v) {
    const float4 q = *reinterpret_cast<const float4 *>(p);
    v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
}
SRE_DEV void row4_load(const double *p, double *v) {
    const double2 a = reinterpret_cast<const double2 *>(p)[0], b = reinterpret_cast<const double2 *>(p)[1];
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}

template <typename R> struct QcpDiff {
    using E = Env<SRE_QCP, R>;
    using Cn = typename E::Consts;

    struct Stage {
        R xd, thd, cin, sn, cs;  // stage inputs; (sin, cos) of the stage angle are kept so the reverse pass reuses them
    };

    // reverse of E::accel: adjoints (gax, gat) of (xdd, thdd) -> adjoints of (th, xd, thd, u, cin), accumulated
    SRE_DEV static void accel_rev(const Cn &c, int flags, const Stage &st, R u, R gax, R gat, R &g_th, R &g_xd,
                                  R &g_thd, R &g_u, R &g_cin) {
        const R sn = st.sn, cs = st.cs;
        const bool simple = flags & SRE_F_SIMPLE_DYN;
        const bool dead = !simple && c.thold_neg <= u && u <= c.thold_pos;
        const R ue = dead ? (R)0 : u;
        const R thd2 = st.thd * st.thd;
        R f_tot = c.k_u * ue - c.k_xd * st.xd;
        R fn = (R)0, sgn = (R)0;
        if (!simple) {
            fn = c.mtg - c.mpl2 * (sn * st.cin + cs * thd2);
            sgn = m_sign(st.xd);
            if (!(fn < (R)0)) f_tot -= c.mu_c * fn * sgn;
        }
        const R M01 = c.mpl * cs;
        const R r0 = f_tot - c.B_eq * st.xd - c.mpl * sn * thd2;
        const R r1 = -c.B_p * st.thd - c.mplg * sn;
        const R inv = m_rcp(c.M00 * c.M11 - M01 * M01);
        const R ax = (c.M11 * r0 - M01 * r1) * inv;
        const R at = (c.M00 * r1 - M01 * r0) * inv;
        // ---- reverse ----
        const R g_r0 = (c.M11 * gax - M01 * gat) * inv;
        const R g_r1 = (c.M00 * gat - M01 * gax) * inv;
        const R g_D = -(ax * gax + at * gat) * inv;
        const R g_M01 = -(r1 * gax + r0 * gat) * inv - (R)2 * M01 * g_D;
        R g_cs = g_M01 * c.mpl;
        R g_sn = -g_r0 * c.mpl * thd2 - g_r1 * c.mplg;
        R gthd = -g_r0 * c.mpl * sn * (R)2 * st.thd - g_r1 * c.B_p;
        R gxd = -g_r0 * c.B_eq - g_r0 * c.k_xd;
        g_u += dead ? (R)0 : g_r0 * c.k_u;
        if (!simple && !(fn < (R)0)) {
            const R g_fn = -g_r0 * c.mu_c * sgn;
            g_sn -= g_fn * c.mpl2 * st.cin;
            g_cs -= g_fn * c.mpl2 * thd2;
            gthd -= g_fn * c.mpl2 * cs * (R)2 * st.thd;
            g_cin += -g_fn * c.mpl2 * sn;
        }
        g_th += g_sn * cs - g_cs * sn;
        g_xd += gxd;
        g_thd += gthd;
    }

    // forward RK4 step that also returns the four stage inputs (needed by the reverse pass)
    SRE_DEV static void step_fwd(const Cn &c, R dt, int flags, const R *s, R c0, R u, R *sn_out, R &c_out,
                                 Stage *stg, R *ax, R *at) {
        const R dt2 = dt * (R)0.5;
        stg[0].xd = s[2]; stg[0].thd = s[3]; stg[0].cin = c0;
        m_sincos(s[1], stg[0].sn, stg[0].cs);
        E::accel_sc(c, flags, stg[0].sn, stg[0].cs, stg[0].xd, stg[0].thd, u, stg[0].cin, ax[0], at[0]);
        R kx[4], kt[4];
        kx[0] = stg[0].xd + ax[0] * dt; kt[0] = stg[0].thd + at[0] * dt;
#pragma unroll
        for (int i = 1; i < 4; i++) {
            const R h = (i < 3) ? dt2 : dt;
            stg[i].xd = s[2] + h * ax[i - 1]; stg[i].thd = s[3] + h * at[i - 1]; stg[i].cin = at[i - 1];
            m_sincos(s[1] + h * kt[i - 1], stg[i].sn, stg[i].cs);
            E::accel_sc(c, flags, stg[i].sn, stg[i].cs, stg[i].xd, stg[i].thd, u, stg[i].cin, ax[i], at[i]);
            kx[i] = stg[i].xd + ax[i] * dt;
            kt[i] = stg[i].thd + at[i] * dt;
        }
        const R w = dt / (R)6;
        sn_out[0] = s[0] + w * (kx[0] + (R)2 * kx[1] + (R)2 * kx[2] + kx[3]);
        sn_out[1] = s[1] + w * (kt[0] + (R)2 * kt[1] + (R)2 * kt[2] + kt[3]);
        sn_out[2] = s[2] + w * (ax[0] + (R)2 * ax[1] + (R)2 * ax[2] + ax[3]);
        sn_out[3] = s[3] + w * (at[0] + (R)2 * at[1] + (R)2 * at[2] + at[3]);
        c_out = (at[0] + at[1] + at[2] + at[3]) / (R)4;
    }

    // reverse of step_fwd: g_out[0..3] adjoint of the next state, g_cout adjoint of the next carry.
    // Returns adjoints of (s[0..3], u, c0).
    SRE_DEV static void step_rev(const Cn &c, R dt, int flags, const Stage *stg, R u, const R *g_out, R g_cout,
                                 R *g_s, R &g_u, R &g_c0) {
        const R dt2 = dt * (R)0.5, w = dt / (R)6;
        const R bw[4] = {w, (R)2 * w, (R)2 * w, w};
        g_s[0] = g_out[0]; g_s[1] = g_out[1]; g_s[2] = g_out[2]; g_s[3] = g_out[3];
        g_u = (R)0;
        R G_ax = (R)0, G_at = (R)0, G_kt = (R)0;  // adjoints flowing from stage i+1 into stage i's outputs
#pragma unroll
        for (int i = 3; i >= 0; i--) {
            const R g_kx = bw[i] * g_out[0];
            const R g_kt = bw[i] * g_out[1] + G_kt;
            const R g_ax = bw[i] * g_out[2] + G_ax + g_kx * dt;
            const R g_at = bw[i] * g_out[3] + g_cout * (R)0.25 + G_at + g_kt * dt;
            R g_th = (R)0, g_xd = g_kx, g_thd = g_kt, g_cin = (R)0;
            accel_rev(c, flags, stg[i], u, g_ax, g_at, g_th, g_xd, g_thd, g_u, g_cin);
            g_s[1] += g_th; g_s[2] += g_xd; g_s[3] += g_thd;
            if (i > 0) {
                const R h = (i < 3) ? dt2 : dt;
                G_kt = h * g_th; G_ax = h * g_xd; G_at = h * g_thd + g_cin;
            } else {
                g_c0 = g_cin;
            }
        }
    }

    // rew_fn (:295-321): returns the reward and d rew / d(state, act_unclamped)
    SRE_DEV static R reward(const KTask<R> &tk, const R *s, R a_un, R *d_s, R &d_a) {
        R e[4];
#pragma unroll
        for (int i = 0; i < 4; i++) e[i] = tk.state_des[i] - s[i];
        const R two_pi = (R)(2 * SRE_PI);
        R theta = e[1] + (-(m_floor(e[1] / two_pi)) * two_pi + (e[1] < (R)0 ? two_pi : (R)0));
        if (theta > (R)SRE_PI) theta -= two_pi;
        e[1] = theta;
        R cost = (R)0, qe[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            R t = (R)0, tt = (R)0;
#pragma unroll
            for (int j = 0; j < 4; j++) { t += tk.Q[i * 4 + j] * e[j]; tt += tk.Q[j * 4 + i] * e[j]; }
            cost += e[i] * t;
            qe[i] = t + tt;  // (Q + Q^T) e
        }
        cost += a_un * tk.Rm[0] * a_un;
        const R r = m_exp(-cost);
#pragma unroll
        for (int i = 0; i < 4; i++) d_s[i] = r * qe[i];  // d/ds exp(-c) = -r * dc/de * de/ds, de/ds = -1
        d_a = -r * (R)2 * tk.Rm[0] * a_un;
        return r;
    }
};

template <typename R>
SRE_DEV void qcp_load_consts(const KCfg<R> &cfg, const R *params, long long n, typename Env<SRE_QCP, R>::Consts &c,
                             R *st_lo, R *st_up, R *act_lo, R *act_up) {
    constexpr int DP = Env<SRE_QCP, R>::DP;
    R dp[DP];
#pragma unroll
    for (int i = 0; i < DP; i++) dp[i] = params ? params[n * DP + i] : cfg.base_params[i];
    Env<SRE_QCP, R>::derive(dp, cfg.flags, c);
    Env<SRE_QCP, R>::bounds(c, cfg.flags, st_lo, st_up, act_lo, act_up);
}

// forward over the horizon; writes the tape (states, carries) the adjoint reads
template <typename R>
__global__ void __launch_bounds__(128)
qcp_diff_forward_kernel(const __grid_constant__ KCfg<R> cfg, const R *params, const R *s0, const R *thdd0,
                        const R *act, long long N, long long H, R *states, R *thdd, R *rew, unsigned char *done) {
    using D = QcpDiff<R>;
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    typename D::Cn c;
    R st_lo[4], st_up[4], act_lo[1], act_up[1];
    qcp_load_consts<R>(cfg, params, n, c, st_lo, st_up, act_lo, act_up);
    R s[4];
    row4_load(s0 + n * 4, s);
    R carry = thdd0 ? thdd0[n] : (R)0;
    for (long long t = 0; t < H; t++) {
        row4_store(states + (t * N + n) * 4, s);
        thdd[t * N + n] = carry;
        const R a_un = act[t * N + n];
        const R u = m_min(m_max(a_un, act_lo[0]), act_up[0]);  // limit_act_clamp :282-284
        R d_s[4], d_a;
        const R r = D::reward(cfg.task, s, a_un, d_s, d_a);
        typename D::Stage stg[4];
        R ax[4], at[4], sn[4], cn;
        D::step_fwd(c, cfg.dt, cfg.flags, s, carry, u, sn, cn, stg, ax, at);
#pragma unroll
        for (int i = 0; i < 4; i++) s[i] = sn[i];
        carry = cn;
        if (rew) rew[t * N + n] = r;
        if (done) {  // is_done :291-293: STRICT bounds
            bool inside = true;
#pragma unroll
            for (int i = 0; i < 4; i++) inside = inside && (st_lo[i] < s[i]) && (st_up[i] > s[i]);
            done[t * N + n] = inside ? 0 : 1;
        }
    }
    row4_store(states + (H * N + n) * 4, s);
    thdd[H * N + n] = carry;
}

// adjoint through the horizon: one thread walks its environment's tape backwards, recomputing the RK4 stages
template <typename R>
__global__ void __launch_bounds__(128)
qcp_diff_backward_kernel(const __grid_constant__ KCfg<R> cfg, const R *params, const R *states, const R *thdd,
                         const R *act, const R *g_states, const R *g_rew, const R *g_thdd, long long N, long long H,
                         R *g_s0, R *g_act, R *g_thdd0) {
    using D = QcpDiff<R>;
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    typename D::Cn c;
    R st_lo[4], st_up[4], act_lo[1], act_up[1];
    qcp_load_consts<R>(cfg, params, n, c, st_lo, st_up, act_lo, act_up);
    R lam[4], lam_c;
#pragma unroll
    for (int i = 0; i < 4; i++) lam[i] = g_states ? g_states[(H * N + n) * 4 + i] : (R)0;
    lam_c = g_thdd ? g_thdd[H * N + n] : (R)0;
    for (long long t = H - 1; t >= 0; t--) {
        R s[4];
        row4_load(states + (t * N + n) * 4, s);
        const R carry = thdd[t * N + n];
        const R a_un = act[t * N + n];
        const bool in_range = (a_un >= act_lo[0]) && (a_un <= act_up[0]);
        const R u = m_min(m_max(a_un, act_lo[0]), act_up[0]);
        typename D::Stage stg[4];
        R ax[4], at[4], sn[4], cn;
        D::step_fwd(c, cfg.dt, cfg.flags, s, carry, u, sn, cn, stg, ax, at);
        R g_s[4], g_u, g_c0;
        D::step_rev(c, cfg.dt, cfg.flags, stg, u, lam, lam_c, g_s, g_u, g_c0);
        R g_a = in_range ? g_u : (R)0;
        const R gr = g_rew ? g_rew[t * N + n] : (R)0;
        if (gr != (R)0) {
            R d_s[4], d_a;
            D::reward(cfg.task, s, a_un, d_s, d_a);
#pragma unroll
            for (int i = 0; i < 4; i++) g_s[i] += gr * d_s[i];
            g_a += gr * d_a;
        }
        g_act[t * N + n] = g_a;
#pragma unroll
        for (int i = 0; i < 4; i++) lam[i] = g_s[i] + (g_states ? g_states[(t * N + n) * 4 + i] : (R)0);
        lam_c = g_c0 + (g_thdd ? g_thdd[t * N + n] : (R)0);
    }
#pragma unroll
    for (int i = 0; i < 4; i++) g_s0[n * 4 + i] = lam[i];
    if (g_thdd0) g_thdd0[n] = lam_c;
}

// per-step Jacobian d(next_state, rew) / d(state, act): five reverse passes with unit adjoints
template <typename R>
__global__ void __launch_bounds__(128)
qcp_diff_jacobian_kernel(const __grid_constant__ KCfg<R> cfg, const R *params, const R *state, const R *thdd,
                         const R *act, long long N, R *jac) {
    using D = QcpDiff<R>;
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    typename D::Cn c;
    R st_lo[4], st_up[4], act_lo[1], act_up[1];
    qcp_load_consts<R>(cfg, params, n, c, st_lo, st_up, act_lo, act_up);
    R s[4];
#pragma unroll
    for (int i = 0; i < 4; i++) s[i] = state[n * 4 + i];
    const R carry = thdd ? thdd[n] : (R)0;
    const R a_un = act[n];
    const bool in_range = (a_un >= act_lo[0]) && (a_un <= act_up[0]);
    const R u = m_min(m_max(a_un, act_lo[0]), act_up[0]);
    typename D::Stage stg[4];
    R ax[4], at[4], sn[4], cn;
    D::step_fwd(c, cfg.dt, cfg.flags, s, carry, u, sn, cn, stg, ax, at);
#pragma unroll
    for (int row = 0; row < 4; row++) {
        R g_out[4] = {(R)0, (R)0, (R)0, (R)0};
        g_out[row] = (R)1;
        R g_s[4], g_u, g_c0;
        D::step_rev(c, cfg.dt, cfg.flags, stg, u, g_out, (R)0, g_s, g_u, g_c0);
#pragma unroll
        for (int i = 0; i < 4; i++) jac[(n * 5 + row) * 5 + i] = g_s[i];
        jac[(n * 5 + row) * 5 + 4] = in_range ? g_u : (R)0;
    }
    R d_s[4], d_a;
    D::reward(cfg.task, s, a_un, d_s, d_a);
#pragma unroll
    for (int i = 0; i < 4; i++) jac[(n * 5 + 4) * 5 + i] = d_s[i];
    jac[(n * 5 + 4) * 5 + 4] = d_a;
}

}  // namespace sre

// ---- tape of sre_qcp_diff_grad_static: no allocation at call time ----------------------------------------------
// File-scope device memory: every GPU that loads the library holds these SRE_STATIC_TAPE_BYTES until the process exits.
// In host code `g_static_tape` is only a stand-in for the symbol, NOT a device address: the address comes from
// cudaGetSymbolAddress, once per device with that device current (static_tape_ptr below).
static __device__ __align__(256) unsigned char g_static_tape[SRE_STATIC_TAPE_BYTES];

using namespace sre;

static int qcp_check(const sre_env_cfg *cfg) {
    int e = check_cfg(cfg);
    if (e) return e;
    SRE_CHECK_ARG(cfg->kind == SRE_QCP, SRE_E_KIND, "the differentiable step exists for the cart-pole only (kind %d)",
                  cfg->kind);
    return 0;
}

extern "C" {

int sre_qcp_diff_forward(const sre_env_cfg *cfg, const void *params, const void *s0, const void *thdd0,
                         const void *act, int64_t n, int64_t horizon, void *states, void *thdd, void *rew,
                         uint8_t *done, void *stream) {
    int e = qcp_check(cfg);
    if (e) return e;
    SRE_CHECK_ARG(s0 && act && states && thdd && n >= 0 && horizon >= 0, SRE_E_ARG, "NULL pointer or negative size");
    SRE_CHECK_ARG(((uintptr_t)s0 | (uintptr_t)states) % 16 == 0, SRE_E_ARG, "s0 and states must be 16-byte aligned");
    if (n == 0) return 0;
    const unsigned blocks = (unsigned)((n + 127) / 128);
    cudaStream_t s = (cudaStream_t)stream;
    if (cfg->dtype == SRE_F64) {
        KCfg<double> k;
        make_kcfg<double>(*cfg, k);
        qcp_diff_forward_kernel<double><<<blocks, 128, 0, s>>>(k, (const double *)params, (const double *)s0,
                                                               (const double *)thdd0, (const double *)act, n, horizon,
                                                               (double *)states, (double *)thdd, (double *)rew, done);
    } else {
        KCfg<float> k;
        make_kcfg<float>(*cfg, k);
        qcp_diff_forward_kernel<float><<<blocks, 128, 0, s>>>(k, (const float *)params, (const float *)s0,
                                                              (const float *)thdd0, (const float *)act, n, horizon,
                                                              (float *)states, (float *)thdd, (float *)rew, done);
    }
    SRE_CUDA(cudaGetLastError());
    return 0;
}

int sre_qcp_diff_backward(const sre_env_cfg *cfg, const void *params, const void *states, const void *thdd,
                          const void *act, const void *g_states, const void *g_rew, const void *g_thdd,
                          int64_t n, int64_t horizon, void *g_s0, void *g_act, void *g_thdd0, void *stream) {
    int e = qcp_check(cfg);
    if (e) return e;
    SRE_CHECK_ARG(states && thdd && act && g_s0 && g_act && n >= 0 && horizon >= 0, SRE_E_ARG,
                  "NULL pointer or negative size");
    SRE_CHECK_ARG((uintptr_t)states % 16 == 0, SRE_E_ARG, "states must be 16-byte aligned");
    if (n == 0) return 0;
    const unsigned blocks = (unsigned)((n + 127) / 128);
    cudaStream_t s = (cudaStream_t)stream;
    if (cfg->dtype == SRE_F64) {
        KCfg<double> k;
        make_kcfg<double>(*cfg, k);
        qcp_diff_backward_kernel<double><<<blocks, 128, 0, s>>>(
            k, (const double *)params, (const double *)states, (const double *)thdd, (const double *)act,
            (const double *)g_states, (const double *)g_rew, (const double *)g_thdd, n, horizon, (double *)g_s0,
            (double *)g_act, (double *)g_thdd0);
    } else {
        KCfg<float> k;
        make_kcfg<float>(*cfg, k);
        qcp_diff_backward_kernel<float><<<blocks, 128, 0, s>>>(
            k, (const float *)params, (const float *)states, (const float *)thdd, (const float *)act,
            (const float *)g_states, (const float *)g_rew, (const float *)g_thdd, n, horizon, (float *)g_s0,
            (float *)g_act, (float *)g_thdd0);
    }
    SRE_CUDA(cudaGetLastError());
    return 0;
}

int sre_qcp_diff_jacobian(const sre_env_cfg *cfg, const void *params, const void *state, const void *thdd,
                          const void *act, int64_t n, void *jac, void *stream) {
    int e = qcp_check(cfg);
    if (e) return e;
    SRE_CHECK_ARG(state && act && jac && n >= 0, SRE_E_ARG, "NULL pointer or negative size");
    if (n == 0) return 0;
    const unsigned blocks = (unsigned)((n + 127) / 128);
    cudaStream_t s = (cudaStream_t)stream;
    if (cfg->dtype == SRE_F64) {
        KCfg<double> k;
        make_kcfg<double>(*cfg, k);
        qcp_diff_jacobian_kernel<double><<<blocks, 128, 0, s>>>(k, (const double *)params, (const double *)state,
                                                                (const double *)thdd, (const double *)act, n,
                                                                (double *)jac);
    } else {
        KCfg<float> k;
        make_kcfg<float>(*cfg, k);
        qcp_diff_jacobian_kernel<float><<<blocks, 128, 0, s>>>(k, (const float *)params, (const float *)state,
                                                               (const float *)thdd, (const float *)act, n, (float *)jac);
    }
    SRE_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"

// Device address of g_static_tape on the CURRENT device, looked up once per device.  cudaGetSymbolAddress allocates
// nothing (the array is part of the loaded module); the cache is a fixed table, so the steady state takes one lock.
static int static_tape_ptr(unsigned char **out) {
    static std::mutex mtx;
    static unsigned char *ptrs[SRE_STATIC_MAX_DEVICES] = {nullptr};
    int dev = -1;
    SRE_CUDA(cudaGetDevice(&dev));
    SRE_CHECK_ARG(dev >= 0 && dev < SRE_STATIC_MAX_DEVICES, SRE_E_UNSUPPORTED,
                  "sre_qcp_diff_grad_static: device ordinal %d beyond the %d the pointer table holds", dev,
                  SRE_STATIC_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(mtx);
    if (!ptrs[dev]) {
        void *p = nullptr;
        SRE_CUDA(cudaGetSymbolAddress(&p, g_static_tape));
        SRE_CHECK_ARG(p != nullptr && (uintptr_t)p % 256 == 0, SRE_E_UNSUPPORTED,
                      "sre_qcp_diff_grad_static: cudaGetSymbolAddress gave %p", p);
        ptrs[dev] = (unsigned char *)p;
    }
    *out = ptrs[dev];
    return 0;
}

template <typename R>
static int diff_grad_static(const sre_env_cfg *cfg, const void *params, const void *s0, const void *thdd0,
                            const void *act, const void *g_states, const void *g_rew, const void *g_thdd, int64_t n,
                            int64_t horizon, void *rew, uint8_t *done, void *g_s0, void *g_act, void *g_thdd0,
                            cudaStream_t s) {
    // tape layout inside the scratch: states [H+1, N, 4] | (256-byte aligned) thdd [H+1, N]
    const unsigned long long rows = (unsigned long long)horizon + 1ull;
    const unsigned long long st_bytes = (rows * (unsigned long long)n * 4ull * sizeof(R) + 255ull) & ~255ull;
    const unsigned long long need = st_bytes + rows * (unsigned long long)n * sizeof(R);
    SRE_CHECK_ARG(need <= (unsigned long long)SRE_STATIC_TAPE_BYTES, SRE_E_ARG,
                  "sre_qcp_diff_grad_static: the tape of %lld envs x %lld steps needs %llu bytes, the static scratch "
                  "holds %llu (SRE_STATIC_TAPE_BYTES); use sre_qcp_diff_forward / _backward with a caller-owned tape",
                  (long long)n, (long long)horizon, need, (unsigned long long)SRE_STATIC_TAPE_BYTES);
    unsigned char *base = nullptr;
    int e = static_tape_ptr(&base);
    if (e) return e;
    R *states = (R *)base;
    R *thdd = (R *)(base + st_bytes);
    const unsigned blocks = (unsigned)((n + 127) / 128);
    KCfg<R> k;
    make_kcfg<R>(*cfg, k);
    // Nothing in the tape has to start at zero: the forward pass writes every (row, env) element the adjoint reads
    // (rows 0..H of the n envs of THIS call), so the previous call's contents are dead.
    qcp_diff_forward_kernel<R><<<blocks, 128, 0, s>>>(k, (const R *)params, (const R *)s0, (const R *)thdd0,
                                                      (const R *)act, n, horizon, states, thdd, (R *)rew, done);
    SRE_CUDA(cudaGetLastError());
    qcp_diff_backward_kernel<R><<<blocks, 128, 0, s>>>(k, (const R *)params, states, thdd, (const R *)act,
                                                       (const R *)g_states, (const R *)g_rew, (const R *)g_thdd, n,
                                                       horizon, (R *)g_s0, (R *)g_act, (R *)g_thdd0);
    SRE_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int sre_qcp_diff_grad_static(const sre_env_cfg *cfg, const void *params, const void *s0, const void *thdd0,
                                        const void *act, const void *g_states, const void *g_rew, const void *g_thdd,
                                        int64_t n, int64_t horizon, void *rew, uint8_t *done, void *g_s0, void *g_act,
                                        void *g_thdd0, void *stream) {
    int e = qcp_check(cfg);
    if (e) return e;
    SRE_CHECK_ARG(s0 && act && g_s0 && g_act && n >= 0 && horizon >= 0, SRE_E_ARG, "NULL pointer or negative size");
    SRE_CHECK_ARG((uintptr_t)s0 % 16 == 0, SRE_E_ARG, "s0 must be 16-byte aligned");
    if (n == 0) return 0;
    cudaStream_t s = (cudaStream_t)stream;
    if (cfg->dtype == SRE_F64)
        return diff_grad_static<double>(cfg, params, s0, thdd0, act, g_states, g_rew, g_thdd, n, horizon, rew, done,
                                        g_s0, g_act, g_thdd0, s);
    return diff_grad_static<float>(cfg, params, s0, thdd0, act, g_states, g_rew, g_thdd, n, horizon, rew, done, g_s0,
                                   g_act, g_thdd0, s);
}
