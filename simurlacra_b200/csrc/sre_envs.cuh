// sre_envs.cuh -- per-environment physics of the pysim family, one struct per env kind, templated on the real
// type.  Every function cites the reference lines whose arithmetic it reproduces (paths relative to
// /root/reference/Pyrado/pyrado/environments/pysim/).  The formulas are written in algebraically reduced form
// (per-episode constants hoisted into `Consts`, common sub-expressions shared); the CPU oracle executes the
// reference's own operation order, so the parity tests compare two independent formulations.
//
// Layout: a thread owns one environment.  Its domain parameters live only as the derived `Consts` (registers),
// computed once per episode in fp64 and rounded to the working type.
#pragma once
#include "sre_device.cuh"

namespace sre {

template <int KIND, typename R> struct Env;

// ======================================================================================================
// QQubeSwingUpSim (Furuta pendulum) -- quanser_qube.py
// state [th, al, thd, ald]; obs [sin th, cos th, sin al, cos al, thd, ald]; act [V]
// params: gravity_const, motor_resistance, motor_back_emf, mass_rot_pole, length_rot_pole, damping_rot_pole,
//         mass_pend_pole, length_pend_pole, damping_pend_pole, voltage_thold_neg, voltage_thold_pos (:54-68)
// ======================================================================================================
template <typename R> struct Env<SRE_QQ, R> {
    static constexpr int DS = 4, DO = 6, DA = 1, DP = 11, NH = 0, DINIT = 4;
    struct Consts {
        R c1, c2, c3, c4, c0;  // _calc_constants :70-87
        R k_u, k_thd;          // motor torque km/Rm * u - (km^2/Rm + Dr) * thd
        R Dp;
        R thold_neg, thold_pos;
    };
    SRE_DEV static void derive(const R *dp, int flags, Consts &c) {
        const double g = dp[0], Rm = dp[1], km = dp[2], mr = dp[3], lr = dp[4], Dr = dp[5], mp = dp[6], lp = dp[7];
        const double Jr = mr * lr * lr / 12, Jp = mp * lp * lp / 12;
        const double c1 = 0.25 * mp * lp * lp;
        c.c0 = (R)(Jr + mp * lr * lr);
        c.c1 = (R)c1;
        c.c2 = (R)(0.5 * mp * lp * lr);
        c.c3 = (R)(Jp + c1);
        c.c4 = (R)(0.5 * mp * lp * g);
        c.k_u = (R)(km / Rm);
        c.k_thd = (R)(km * km / Rm + Dr);
        c.Dp = dp[8];
        c.thold_neg = dp[9];
        c.thold_pos = dp[10];
    }
    // _create_spaces :167-181: no bound depends on a domain parameter
    SRE_DEV static void bounds(const Consts &c, int flags, R *st_lo, R *st_up, R *act_lo, R *act_up) {
        const double ms[4] = {115.0 / 180 * SRE_PI, 4 * SRE_PI, 20 * SRE_PI, 20 * SRE_PI};
#pragma unroll
        for (int i = 0; i < 4; i++) { st_lo[i] = (R)(-ms[i]); st_up[i] = (R)ms[i]; }
        act_lo[0] = (R)-4.5; act_up[0] = (R)4.5;  // MAX_ACT_QQ, environments/quanser/__init__.py:34
    }
    // _dyn :89-125 evaluated once; the fork's pseudo-RK4 (:127-146) evaluates _dyn at self.state on every stage,
    // so the update is exactly v += dt*a0, p += dt*v0 + dt^2/2*a0.
    // trig = (sin al, cos al) of the CURRENT state, produced by observe(): the dynamics reuse it
    SRE_DEV static void step(const Consts &c, R dt, int flags, R *s, R *h, R *act, const R *trig) {
        R u = act[0];
        if (c.thold_neg <= u && u <= c.thold_pos) u = (R)0;
        const R thd = s[2], ald = s[3];
        const R sa = trig[0], ca = trig[1];
        const R s2a = (R)2 * sa * ca;  // sin(2 al)
        const R a = c.c0 + c.c1 * (sa * sa);
        const R b = c.c2 * ca;
        const R det = a * c.c3 - b * b;
        const R t1 = c.c1 * s2a * thd;  // shared by both Coriolis terms
        const R xx = c.k_u * u - c.k_thd * thd - (t1 * ald - c.c2 * sa * (ald * ald));
        const R yy = -c.Dp * ald - ((R)-0.5 * t1 * thd + c.c4 * sa);
        const R inv = m_rcp(det);
        const R thdd = (c.c3 * xx - b * yy) * inv;
        const R aldd = (a * yy - b * xx) * inv;
        const R hdt2 = (R)0.5 * dt * dt;
        s[0] += dt * thd + hdt2 * thdd;
        s[1] += dt * ald + hdt2 * aldd;
        s[2] += dt * thdd;
        s[3] += dt * aldd;
    }
    SRE_DEV static void observe(const R *s, R *obs, R *trig) {  // :148-149
        if constexpr (sizeof(R) == 4) {  // float build: the two angles as one packed pair (15 packed + 16 scalar instructions for both)
            P2 sn, cs;
            sincos2(p2((float)s[0], (float)s[1]), sn, cs);
            float a, b;
            pun(sn, a, b); obs[0] = (R)a; obs[2] = (R)b;
            pun(cs, a, b); obs[1] = (R)a; obs[3] = (R)b;
        } else {
            m_sincos(s[0], obs[0], obs[1]);
            m_sincos(s[1], obs[2], obs[3]);
        }
        obs[4] = s[2];
        obs[5] = s[3];
        trig[0] = obs[2];
        trig[1] = obs[3];
    }
    SRE_DEV static void reset_hidden(const R *dp, int flags, const R *s, R *h) {}
    SRE_DEV static void state_from_init(const R *init, int n_cols, R *s) {
#pragma unroll
        for (int i = 0; i < 4; i++) s[i] = init[i];
    }
    // default init space :169,177
    SRE_DEV static void init_box(const R *dp, int wild_init, bool right, double *lo, double *up) {
        const double m[4] = {2.0, 1.0, 0.5, 0.5};
#pragma unroll
        for (int i = 0; i < 4; i++) { up[i] = m[i] / 180 * SRE_PI; lo[i] = -up[i]; }
    }
};

// ======================================================================================================
// QCartPoleSwingUpSim -- quanser_cartpole.py
// state [x, th, xd, thd]; hidden [th_ddot carried between RK4 stages and steps]; obs [x, sin th, cos th, xd, thd]
// params (:111-143): gravity_const, cart_mass, rail_length, motor_efficiency, gear_efficiency, gear_ratio,
//   motor_inertia, pinion_radius, motor_resistance, motor_back_emf, pole_damping, combined_damping, pole_mass,
//   pole_length, cart_friction_coeff, voltage_thold_neg, voltage_thold_pos
// ======================================================================================================
template <typename R> struct Env<SRE_QCP, R> {
    static constexpr int DS = 4, DO = 5, DA = 1, DP = 17, NH = 1, DINIT = 4;
    struct Consts {
        R M00, M11, mpl;       // mass matrix [[M00, mpl cos],[mpl cos, M11]]
        R k_u, k_xd;           // f_act = k_u * u - k_xd * xd
        R B_eq, B_p, mplg;     // rhs terms
        R mtg, mpl2, mu_c;     // normal force: mtg - mpl2 * (sin*thdd_in + cos*thd^2)
        R thold_neg, thold_pos;
        R x_max, xd_max;       // the two state bounds that depend on rail_length
    };
    SRE_HD static void derive(const R *dp, int flags, Consts &c) {
        const double g = dp[0], m_c = dp[1], l_rail = dp[2], eta_m = dp[3], eta_g = dp[4], K_g = dp[5], J_m = dp[6],
                     r_mp = dp[7], R_m = dp[8], k_m = dp[9], B_p = dp[10], B_eq = dp[11], m_p = dp[12], l_p = dp[13],
                     mu_c = dp[14];
        const double J_pole = l_p * l_p * m_p / 3.0;                           // :145-155
        const double J_eq = m_c + (eta_g * K_g * K_g * J_m) / (r_mp * r_mp);
        const double kf = (eta_g * K_g * eta_m * k_m) / (R_m * r_mp);         // :196
        c.M00 = (R)(m_p + J_eq);
        c.M11 = (R)(J_pole + m_p * l_p * l_p);
        c.mpl = (R)(m_p * l_p);
        c.k_u = (R)(kf * eta_m);
        c.k_xd = (R)(kf * K_g * k_m / r_mp);
        c.B_eq = (R)B_eq; c.B_p = (R)B_p;
        c.mplg = (R)(m_p * l_p * g);
        c.mtg = (R)((m_c + m_p) * g);
        c.mpl2 = (R)(m_p * l_p / 2);
        c.mu_c = (R)mu_c;
        c.thold_neg = dp[15]; c.thold_pos = dp[16];
        c.x_max = (R)(l_rail / 2.0 - 0.15);  // _create_spaces :541-549, x_buffer = 0.15 (:79)
        c.xd_max = (R)l_rail;
    }
    SRE_HD static void bounds(const Consts &c, int flags, R *st_lo, R *st_up, R *act_lo, R *act_up) {
        st_up[0] = c.x_max; st_up[1] = (R)(4 * SRE_PI); st_up[2] = c.xd_max; st_up[3] = (R)(20 * SRE_PI);
#pragma unroll
        for (int i = 0; i < 4; i++) st_lo[i] = -st_up[i];
        if (flags & SRE_F_QCP_STAB) {  // QCartPoleStabSim._create_spaces :473-482, stab_thold = 15 deg (:467)
            const double thold = 15 / 180.0 * SRE_PI;
            st_lo[1] = (R)(SRE_PI - thold); st_up[1] = (R)(SRE_PI + thold);
            st_lo[3] = (R)(-2 * SRE_PI); st_up[3] = (R)(2 * SRE_PI);
        }
        act_lo[0] = (R)-6; act_up[0] = (R)6;  // MAX_ACT_QCP
    }
    // _dynamics :166-230 == _dynamics_diff :360-431.  Returns x_ddot, th_ddot at (th, xd, thd, u) with the carried
    // th_ddot_in in the Coulomb friction.  The caller forms the "post-Euler velocity" derivative (:226-230).
    SRE_DEV static void accel(const Consts &c, int flags, R th, R xd, R thd, R u, R thdd_in, R &xdd, R &thdd) {
        R sn, cs;
        m_sincos(th, sn, cs);
        accel_sc(c, flags, sn, cs, xd, thd, u, thdd_in, xdd, thdd);
    }
    SRE_DEV static void accel_sc(const Consts &c, int flags, R sn, R cs, R xd, R thd, R u, R thdd_in, R &xdd, R &thdd) {
        const bool simple = flags & SRE_F_SIMPLE_DYN;
        if (!simple && c.thold_neg <= u && u <= c.thold_pos) u = (R)0;
        R f_tot = c.k_u * u - c.k_xd * xd;
        const R thd2 = thd * thd;
        if (!simple) {
            const R f_normal = c.mtg - c.mpl2 * (sn * thdd_in + cs * thd2);
            const R f_c = (f_normal < (R)0) ? (R)0 : c.mu_c * f_normal * m_sign(xd);
            f_tot -= f_c;
        }
        const R M01 = c.mpl * cs;
        const R r0 = f_tot - c.B_eq * xd - c.mpl * sn * thd2;
        const R r1 = -c.B_p * thd - c.mplg * sn;
        // 2x2 solve (np.linalg.solve :224 / torch.solve :422) by Cramer's rule
        const R inv = m_rcp(c.M00 * c.M11 - M01 * M01);
        xdd = (c.M11 * r0 - M01 * r1) * inv;
        thdd = (c.M00 * r1 - M01 * r0) * inv;
    }
    // rk4() :591-655 over [0, dt] on [x, th, xd, thd, u] with derivative [xd + xdd dt, thd + thdd dt, xdd, thdd, 0];
    // th_ddot of stage i feeds stage i+1, the mean of the four becomes the next step's carry (:652).
    // trig = (sin th, cos th) of the current state (from observe()), used by the first RK4 stage
    SRE_DEV static void step(const Consts &c, R dt, int flags, R *s, R *h, R *act, const R *trig) {
        const R u = act[0];
        const R dt2 = dt * (R)0.5;
        R a1x, a1t, a2x, a2t, a3x, a3t, a4x, a4t;
        accel_sc(c, flags, trig[0], trig[1], s[2], s[3], u, h[0], a1x, a1t);
        const R k1x = s[2] + a1x * dt, k1t = s[3] + a1t * dt;
        accel(c, flags, s[1] + dt2 * k1t, s[2] + dt2 * a1x, s[3] + dt2 * a1t, u, a1t, a2x, a2t);
        const R k2x = (s[2] + dt2 * a1x) + a2x * dt, k2t = (s[3] + dt2 * a1t) + a2t * dt;
        accel(c, flags, s[1] + dt2 * k2t, s[2] + dt2 * a2x, s[3] + dt2 * a2t, u, a2t, a3x, a3t);
        const R k3x = (s[2] + dt2 * a2x) + a3x * dt, k3t = (s[3] + dt2 * a2t) + a3t * dt;
        accel(c, flags, s[1] + dt * k3t, s[2] + dt * a3x, s[3] + dt * a3t, u, a3t, a4x, a4t);
        const R k4x = (s[2] + dt * a3x) + a4x * dt, k4t = (s[3] + dt * a3t) + a4t * dt;
        const R w = dt / (R)6;
        s[0] += w * (k1x + (R)2 * k2x + (R)2 * k3x + k4x);
        s[1] += w * (k1t + (R)2 * k2t + (R)2 * k3t + k4t);
        s[2] += w * (a1x + (R)2 * a2x + (R)2 * a3x + a4x);
        s[3] += w * (a1t + (R)2 * a2t + (R)2 * a3t + a4t);
        h[0] = (a1t + a2t + a3t + a4t) / (R)4;
    }
    SRE_DEV static void observe(const R *s, R *obs, R *trig) {  // :108-109
        obs[0] = s[0];
        m_sincos(s[1], obs[1], obs[2]);
        obs[3] = s[2];
        obs[4] = s[3];
        trig[0] = obs[1];
        trig[1] = obs[2];
    }
    SRE_DEV static void reset_hidden(const R *dp, int flags, const R *s, R *h) { h[0] = (R)0; }  // :101-106
    SRE_DEV static void state_from_init(const R *init, int n_cols, R *s) {
#pragma unroll
        for (int i = 0; i < 4; i++) s[i] = init[i];
    }
    // :552-560; wild_init selects the box: 0 -> 'True', 1 -> 'False', 2 -> else
    SRE_DEV static void init_box(const R *dp, int wild_init, bool right, double *lo, double *up) {
        double m[4];
        if (wild_init == 0) { m[0] = 0.25; m[1] = SRE_PI; m[2] = 0.8; m[3] = SRE_PI; }
        else if (wild_init == 1) { m[0] = 0.02; m[1] = 2 / 180.0 * SRE_PI; m[2] = 0.0; m[3] = 1 / 180.0 * SRE_PI; }
        else { m[0] = 0.02; m[1] = SRE_PI; m[2] = 0.0; m[3] = 1 / 180.0 * SRE_PI; }
#pragma unroll
        for (int i = 0; i < 4; i++) { up[i] = m[i]; lo[i] = -m[i]; }
    }
};

// ======================================================================================================
// QBallBalancerSim -- quanser_ball_balancer.py
// state [th_x, th_y, x, y, th_xd, th_yd, xd, yd]; hidden plate_angs [a, b]; obs = state; act [Vx, Vy]
// params (:169-202): gravity_const, ball_mass, ball_radius, plate_length, arm_radius, gear_ratio, gear_efficiency,
//   load_inertia, motor_inertia, motor_back_emf, motor_resistance, motor_efficiency, combined_damping,
//   ball_damping, voltage_thold_x_pos, voltage_thold_x_neg, voltage_thold_y_pos, voltage_thold_y_neg,
//   offset_th_x, offset_th_y
// ======================================================================================================
template <typename R> struct Env<SRE_QBB, R> {
    static constexpr int DS = 8, DO = 8, DA = 2, DP = 20, NH = 2, DINIT = 4;
    struct Consts {
        R k_act, k_vel;          // servo accel: (A_m * u - B_eq_v * thd) / J_eq = k_act*u - k_vel*thd
        R c_kin;
        R kx_damp, kx_J, kx_m, kx_g;  // ball accel terms divided by zeta (kx_g = c_kin m g r^2 / zeta)
        R Vxp, Vxn, Vyp, Vyn, off_x, off_y;
        R half_plate;  // the state bound that depends on plate_length
    };
    SRE_DEV static void derive(const R *dp, int flags, Consts &c) {
        const double g = dp[0], m_ball = dp[1], r_ball = dp[2], l_plate = dp[3], r_arm = dp[4], K_g = dp[5],
                     eta_g = dp[6], J_l = dp[7], J_m = dp[8], k_m = dp[9], R_m = dp[10], eta_m = dp[11],
                     B_eq = dp[12], c_frict = dp[13];
        // _calc_constants :204-223
        const double J_ball = 2.0 / 5 * m_ball * r_ball * r_ball;
        const double J_eq = eta_g * K_g * K_g * J_m + J_l;
        const double c_kin = 2.0 * r_arm / l_plate;
        const double A_m = eta_g * K_g * eta_m * k_m / R_m;
        const double B_eq_v = eta_g * K_g * K_g * eta_m * k_m * k_m / R_m + B_eq;
        const double zeta = m_ball * r_ball * r_ball + J_ball;
        const double r2 = r_ball * r_ball;
        c.k_act = (R)(A_m / J_eq);
        c.k_vel = (R)(B_eq_v / J_eq);
        c.c_kin = (R)c_kin;
        c.kx_damp = (R)(c_frict * r2 / zeta);
        c.kx_J = (R)(J_ball * r_ball / zeta);
        c.kx_m = (R)(m_ball * r2 / zeta);
        c.kx_g = (R)(c_kin * m_ball * g * r2 / zeta);
        c.Vxp = dp[14]; c.Vxn = dp[15]; c.Vyp = dp[16]; c.Vyn = dp[17]; c.off_x = dp[18]; c.off_y = dp[19];
        c.half_plate = (R)(l_plate / 2.0);
    }
    // _create_spaces :93-117
    SRE_DEV static void bounds(const Consts &c, int flags, R *st_lo, R *st_up, R *act_lo, R *act_up) {
        const R ms[8] = {(R)(SRE_PI / 4.0), (R)(SRE_PI / 4.0), c.half_plate, c.half_plate, (R)(5 * SRE_PI), (R)(5 * SRE_PI), (R)0.5, (R)0.5};
#pragma unroll
        for (int i = 0; i < 8; i++) { st_lo[i] = -ms[i]; st_up[i] = ms[i]; }
        act_lo[0] = act_lo[1] = (R)-3; act_up[0] = act_up[1] = (R)3;  // MAX_ACT_QBB
    }
    // _step_dynamics :247-331
    SRE_DEV static void step(const Consts &c, R dt, int flags, R *s, R *h, R *act, const R *trig) {
        const bool simple = flags & SRE_F_SIMPLE_DYN;
        if (!simple && c.Vxn <= act[0] && act[0] <= c.Vxp) act[0] = (R)0;  // dead zones :261-264 (in place)
        if (!simple && c.Vyn <= act[1] && act[1] <= c.Vyp) act[1] = (R)0;
        if constexpr (sizeof(R) == 4) {
            // Production (float) build: the two axes of the plate are the same formulas on (x, y) pairs -- servo angles,
            // plate angles, ball positions and all their derivatives -- so the whole step runs on packed f32x2 arithmetic
            // with the pair {x axis, y axis} in one register pair: two sincos2 instead of four sincos, one packed Newton
            // step for the two reciprocals, every update one instruction for both axes.  Same operation order as the scalar
            // code below (which the fp64 verification build runs); the y axis' sign flips sit in the constants.
            const P2 pdt = p2s((float)dt);
            const P2 th = p2((float)s[0], (float)s[1]) + p2((float)c.off_x, (float)c.off_y);
            const P2 thd = p2((float)s[4], (float)s[5]);
            const P2 thdd = pfma(p2s(-(float)c.k_vel), thd, p2s((float)c.k_act) * p2((float)act[0], (float)act[1]));
            P2 sxy, cxy, sab, cab;
            sincos2(th, sxy, cxy);
            sincos2(p2((float)h[0], (float)h[1]), sab, cab);
            // 1 / cos of the plate angles (m_rcp): MUFU seeds + one packed Newton step
            float ca, cb, ra, rb;
            pun(cab, ca, cb);
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ra) : "f"(ca));
            asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rb) : "f"(cb));
            const P2 r0 = p2(ra, rb);
            const P2 icab = psub(r0, r0 * pfma(cab, r0, p2s(-1.0f)));
            const P2 ck = p2((float)c.c_kin, -(float)c.c_kin);  // a_dot = c_kin thxd cx / ca, b_dot = -c_kin thyd cy / cb
            const P2 abd = ((ck * thd) * cxy) * icab;
            const P2 abd2 = abd * abd;
            P2 acc;
            if (simple) {
                acc = p2s((float)c.kx_g) * sxy;
            } else {
                const P2 t = psub(thdd * cxy, (thd * thd) * sxy);
                const P2 abdd = icab * pfma(ck, t, abd2 * sab);
                acc = pfma(p2s(-(float)c.kx_J), abdd, p2s(-(float)c.kx_damp) * p2((float)s[6], (float)s[7]));
                acc = pfma(p2s((float)c.kx_m) * p2((float)s[2], (float)s[3]), abd2, acc);
                acc = pfma(p2s((float)c.kx_g), sxy, acc);
            }
            const P2 hn = pfma(abd, pdt, p2((float)h[0], (float)h[1]));
            const P2 s45 = pfma(thdd, pdt, thd);
            const P2 s67 = pfma(acc, pdt, p2((float)s[6], (float)s[7]));
            const P2 s01 = pfma(s45, pdt, p2((float)s[0], (float)s[1]));
            const P2 s23 = pfma(s67, pdt, p2((float)s[2], (float)s[3]));
            float o0, o1;
            pun(hn, o0, o1); h[0] = (R)o0; h[1] = (R)o1;
            pun(s01, o0, o1); s[0] = (R)o0; s[1] = (R)o1;
            pun(s23, o0, o1); s[2] = (R)o0; s[3] = (R)o1;
            pun(s45, o0, o1); s[4] = (R)o0; s[5] = (R)o1;
            pun(s67, o0, o1); s[6] = (R)o0; s[7] = (R)o1;
            return;
        }
        const R th_x = s[0] + c.off_x, th_y = s[1] + c.off_y;
        const R thxd = s[4], thyd = s[5];
        const R thxdd = c.k_act * act[0] - c.k_vel * thxd;  // :276-277
        const R thydd = c.k_act * act[1] - c.k_vel * thyd;
        R sx, cx, sy, cy;
        m_sincos(th_x, sx, cx);
        m_sincos(th_y, sy, cy);
        // plate angular velocities :289-290 (cos(-x) = cos x); integrated in both modes (:330)
        R sa, ca, sb, cb;
        m_sincos(h[0], sa, ca);
        m_sincos(h[1], sb, cb);
        const R ica = m_rcp(ca), icb = m_rcp(cb);
        const R a_dot = c.c_kin * thxd * cx * ica;
        const R b_dot = -c.c_kin * thyd * cy * icb;
        R xdd, ydd;
        if (simple) {
            xdd = c.kx_g * sx;  // :306-309
            ydd = c.kx_g * sy;
        } else {
            // plate angular accelerations :292-303 (sin(-x) = -sin x)
            const R a_ddot = ica * (c.c_kin * (thxdd * cx - thxd * thxd * sx) + a_dot * a_dot * sa);
            const R b_ddot = icb * (c.c_kin * (-thydd * cy + thyd * thyd * sy) + b_dot * b_dot * sb);
            // ball accelerations :312-323
            xdd = -c.kx_damp * s[6] - c.kx_J * a_ddot + c.kx_m * s[2] * (a_dot * a_dot) + c.kx_g * sx;
            ydd = -c.kx_damp * s[7] - c.kx_J * b_ddot + c.kx_m * s[3] * (b_dot * b_dot) + c.kx_g * sy;
        }
        h[0] += a_dot * dt;  // :330 forward Euler on the plate angles (pre-step velocities)
        h[1] += b_dot * dt;
        // symplectic Euler :326-327
        s[4] += thxdd * dt; s[5] += thydd * dt; s[6] += xdd * dt; s[7] += ydd * dt;
#pragma unroll
        for (int i = 0; i < 4; i++) s[i] += s[4 + i] * dt;
    }
    SRE_DEV static void observe(const R *s, R *obs, R *trig) {
#pragma unroll
        for (int i = 0; i < 8; i++) obs[i] = s[i];
    }
    // reset :231-245: plate angles from the inverse kinematics of the servo angles
    SRE_DEV static void reset_hidden(const R *dp, int flags, const R *s, R *h) {
        h[0] = h[1] = (R)0;
        if (flags & SRE_F_SIMPLE_DYN) return;
        h[0] = (R)plate_ang(dp[4], dp[3], s[0] + dp[18]);
        h[1] = (R)plate_ang(dp[4], dp[3], s[1] + dp[19]);
    }
    // QBallBalancerKin.__call__ :375-444: 100 iterations of torch SGD(lr=0.01, momentum=0.9) in float32 on the
    // two-constraint loss (:423-436) with the analytic gradient (the reference uses autograd), then the plate
    // angle (:438-444).  One rounding per operation like the eager torch loop: no FMA contraction.
    SRE_DEV static float plate_ang(R arm_radius, R plate_length, R th_in) {
        const float r = (float)arm_radius;
        const float l = (float)((double)plate_length / 2.0);
        const float d = 0.10f;
        const float th = (float)th_in;
        float tip0 = r, tip1 = l, buf0 = 0.f, buf1 = 0.f;
        const float cx = __fmul_rn(r, cosf(th)), cy = __fmul_rn(r, sinf(th));
        if (sizeof(R) == 4) {
            // production (float) build: the same 100 momentum-SGD iterations with one rsqrt per constraint instead of an
            // IEEE sqrt and an IEEE division, and contracted FMAs.  The iteration is a contraction, so the 1-2 ulp of
            // the approximations do not accumulate: the plate angle agrees with the exact-rounding loop below to a few
            // float32 ulp (the float build's parity budget is 1e-4), at a third of the instructions -- the two solves
            // are the most expensive part of an episode start.
            for (int i = 0; i < 100; i++) {
                const float ax = tip0 - cx, ay = tip1 - cy;
                const float bx = (tip0 - r) - l, by = tip1 - d;
                const float q1 = ax * ax + ay * ay, q2 = bx * bx + by * by;
                const float i1 = rsqrtf(q1), i2 = rsqrtf(q2);
                const float t1 = (q1 * i1 - d) * i1, t2 = (q2 * i2 - l) * i2;  // (len - target) / len
                const float g0 = 2.f * (t1 * ax + t2 * bx), g1 = 2.f * (t1 * ay + t2 * by);
                buf0 = i == 0 ? g0 : 0.9f * buf0 + g0;
                buf1 = i == 0 ? g1 : 0.9f * buf1 + g1;
                tip0 -= 0.01f * buf0;
                tip1 -= 0.01f * buf1;
            }
        } else
        for (int i = 0; i < 100; i++) {
            const float ax = __fsub_rn(tip0, cx), ay = __fsub_rn(tip1, cy);
            const float rod_len = __fsqrt_rn(__fadd_rn(__fmul_rn(ax, ax), __fmul_rn(ay, ay)));
            const float bx = __fsub_rn(__fsub_rn(tip0, r), l), by = __fsub_rn(tip1, d);
            const float half_plate = __fsqrt_rn(__fadd_rn(__fmul_rn(bx, bx), __fmul_rn(by, by)));
            const float e1 = __fsub_rn(rod_len, d), e2 = __fsub_rn(half_plate, l);
            const float t1 = __fdiv_rn(e1, rod_len), t2 = __fdiv_rn(e2, half_plate);
            const float g0 = __fadd_rn(__fmul_rn(t1, __fmul_rn(2.f, ax)), __fmul_rn(t2, __fmul_rn(2.f, bx)));
            const float g1 = __fadd_rn(__fmul_rn(t1, __fmul_rn(2.f, ay)), __fmul_rn(t2, __fmul_rn(2.f, by)));
            if (i == 0) { buf0 = g0; buf1 = g1; }
            else { buf0 = __fadd_rn(__fmul_rn(0.9f, buf0), g0); buf1 = __fadd_rn(__fmul_rn(0.9f, buf1), g1); }
            tip0 = __fsub_rn(tip0, __fmul_rn(0.01f, buf0));
            tip1 = __fsub_rn(tip1, __fmul_rn(0.01f, buf1));
        }
        const float rl = (float)((double)arm_radius + (double)plate_length / 2.0);
        return __fsub_rn((float)(SRE_PI / 2.0), atan2f(__fsub_rn(rl, tip0), __fsub_rn(tip1, d)));
    }
    SRE_DEV static void state_from_init(const R *init, int n_cols, R *s) {
        if (n_cols == 8) {
#pragma unroll
            for (int i = 0; i < 8; i++) s[i] = init[i];
        } else {  // _state_from_init :225-229
#pragma unroll
            for (int i = 0; i < 8; i++) s[i] = (R)0;
            s[2] = init[0]; s[3] = init[1]; s[6] = init[2]; s[7] = init[3];
        }
    }
    // Polar2DPosVelSpace box :108-117 (r, phi, xd, yd); the polar -> cartesian map is applied by the caller
    SRE_DEV static void init_box(const R *dp, int wild_init, bool right, double *lo, double *up) {
        const double l_plate = dp[3];
        lo[0] = 0.75 * l_plate / 2; lo[1] = -SRE_PI; lo[2] = -0.05 * 0.5; lo[3] = -0.05 * 0.5;
        up[0] = 0.8 * l_plate / 2; up[1] = SRE_PI; up[2] = 0.05 * 0.5; up[3] = 0.05 * 0.5;
    }
};

// ======================================================================================================
// BallOnBeamSim -- ball_on_beam.py
// state [x, a, xd, ad]; obs = state; act [torque]
// params (:75-87): gravity_const, ball_mass, ball_radius, beam_mass, beam_length, beam_thickness, friction_coeff,
//   ang_offset
// ======================================================================================================
template <typename R> struct Env<SRE_BOB, R> {
    static constexpr int DS = 4, DO = 4, DA = 1, DP = 8, NH = 0, DINIT = 4;
    struct Consts {
        R m_ball, mg, J_beam, k_frict, k_m, k_g, ang_offset;
        R half_beam, act_max;  // param-dependent bounds
    };
    SRE_DEV static void derive(const R *dp, int flags, Consts &c) {
        const double g = dp[0], m_ball = dp[1], r_ball = dp[2], m_beam = dp[3], l_beam = dp[4], d_beam = dp[5],
                     c_frict = dp[6];
        // _calc_constants :89-98
        const double J_ball = 2.0 / 5 * m_ball * r_ball * r_ball;
        const double J_beam = 1.0 / 12 * m_beam * (l_beam * l_beam + d_beam * d_beam);
        const double zeta_ball = m_ball + J_ball / (r_ball * r_ball);
        c.m_ball = (R)m_ball;
        c.mg = (R)(m_ball * g);
        c.J_beam = (R)J_beam;
        c.k_frict = (R)(c_frict / zeta_ball);
        c.k_m = (R)(m_ball / zeta_ball);
        c.k_g = (R)(m_ball * g / zeta_ball);
        c.ang_offset = dp[7];
        c.half_beam = (R)(l_beam / 2.0);
        c.act_max = (R)(l_beam / 2.0 * g * 3.0);
    }
    // _create_spaces :48-58
    SRE_DEV static void bounds(const Consts &c, int flags, R *st_lo, R *st_up, R *act_lo, R *act_up) {
        const R ms[4] = {c.half_beam, (R)(SRE_PI / 4.0), (R)10.0, (R)SRE_PI};
#pragma unroll
        for (int i = 0; i < 4; i++) { st_lo[i] = -ms[i]; st_up[i] = ms[i]; }
        act_lo[0] = -c.act_max; act_up[0] = c.act_max;
    }
    // _step_dynamics :110-130
    SRE_DEV static void step(const Consts &c, R dt, int flags, R *s, R *h, R *act, const R *trig) {
        const R x = s[0], xd = s[2], ad = s[3];
        R sa, ca;
        m_sincos(s[1] + c.ang_offset, sa, ca);
        const R zeta_beam = c.m_ball * (x * x) + c.J_beam;
        const R xdd = -c.k_frict * xd + c.k_m * x * (ad * ad) - c.k_g * sa;
        const R add = (act[0] - (R)2 * c.m_ball * x * xd * ad - c.mg * ca * x) / zeta_beam;
        s[2] += xdd * dt; s[3] += add * dt;
        s[0] += s[2] * dt; s[1] += s[3] * dt;
    }
    SRE_DEV static void observe(const R *s, R *obs, R *trig) {
#pragma unroll
        for (int i = 0; i < 4; i++) obs[i] = s[i];
    }
    SRE_DEV static void reset_hidden(const R *dp, int flags, const R *s, R *h) {}
    SRE_DEV static void state_from_init(const R *init, int n_cols, R *s) {
#pragma unroll
        for (int i = 0; i < 4; i++) s[i] = init[i];
    }
    // CompoundSpace of two boxes :60-73, `right` picks the box
    SRE_DEV static void init_box(const R *dp, int wild_init, bool right, double *lo, double *up) {
        const double l_beam = dp[4];
        if (!right) { lo[0] = -0.8 * l_beam / 2.0; up[0] = -0.7 * l_beam / 2.0; }
        else { lo[0] = 0.7 * l_beam / 2.0; up[0] = 0.8 * l_beam / 2.0; }
        lo[1] = -5 / 180.0 * SRE_PI; up[1] = +5 / 180.0 * SRE_PI;
        lo[2] = -0.02 * 10.0; up[2] = +0.02 * 10.0;
        lo[3] = -0.02 * SRE_PI; up[3] = +0.02 * SRE_PI;
    }
};

// ======================================================================================================
// OneMassOscillatorSim -- one_mass_oscillator.py
// state [x, xd]; obs = state; act [force]; params (:81-87): mass, stiffness, damping
// ======================================================================================================
template <typename R> struct Env<SRE_OMO, R> {
    static constexpr int DS = 2, DO = 2, DA = 1, DP = 3, NH = 0, DINIT = 2;
    struct Consts {
        R A10, A11, B1;
        R act_max;
    };
    SRE_DEV static void derive(const R *dp, int flags, Consts &c) {
        const double m = dp[0], k = dp[1], d = dp[2];
        const double omega = sqrt(k / m), zeta = d / (2.0 * sqrt(m * k));  // :89-103
        c.A10 = (R)(-(omega * omega));
        c.A11 = (R)(-2.0 * zeta * omega);
        c.B1 = (R)(1.0 / m);
        c.act_max = (R)k;
    }
    SRE_DEV static void bounds(const Consts &c, int flags, R *st_lo, R *st_up, R *act_lo, R *act_up) {
        st_lo[0] = (R)-1; st_up[0] = (R)1; st_lo[1] = (R)-10; st_up[1] = (R)10;  // :54-66
        act_lo[0] = -c.act_max; act_up[0] = c.act_max;
    }
    // _step_dynamics :105-116: forward Euler on s' = A s + B a
    SRE_DEV static void step(const Consts &c, R dt, int flags, R *s, R *h, R *act, const R *trig) {
        const R sd0 = s[1];
        const R sd1 = c.A10 * s[0] + c.A11 * s[1] + c.B1 * act[0];
        s[0] += sd0 * dt;
        s[1] += sd1 * dt;
    }
    SRE_DEV static void observe(const R *s, R *obs, R *trig) { obs[0] = s[0]; obs[1] = s[1]; }
    SRE_DEV static void reset_hidden(const R *dp, int flags, const R *s, R *h) {}
    SRE_DEV static void state_from_init(const R *init, int n_cols, R *s) { s[0] = init[0]; s[1] = init[1]; }
    SRE_DEV static void init_box(const R *dp, int wild_init, bool right, double *lo, double *up) {  // :57-58
        lo[0] = -0.75; up[0] = -0.65; lo[1] = -0.1; up[1] = 0.1;
    }
};

// ======================================================================================================
// PendulumSim -- pendulum.py:43-140 (under-actuated pendulum modelled as a rod)
// state [th, thd]; obs [sin th, cos th, thd]; act [torque]; params (:96-104): gravity_const, pole_mass, pole_length,
// pole_damping, torque_thold.  The init space is a SingularStateSpace (:75): the host passes it as a box with lo == up.
// ======================================================================================================
template <typename R> struct Env<SRE_PEND, R> {
    static constexpr int DS = 2, DO = 3, DA = 1, DP = 5, NH = 0, DINIT = 2;
    struct Consts {
        R mgl2, d, inv_J;  // m g l / 2, damping, 1 / (m l^2 / 3)
        R act_max;
    };
    SRE_DEV static void derive(const R *dp, int flags, Consts &c) {
        const double g = dp[0], m = dp[1], l = dp[2];
        c.mgl2 = (R)(m * g * l / 2.0);
        c.d = dp[3];
        c.inv_J = (R)(1.0 / (m * (l * l) / 3.0));
        c.act_max = dp[4];
    }
    SRE_DEV static void bounds(const Consts &c, int flags, R *st_lo, R *st_up, R *act_lo, R *act_up) {
        st_lo[0] = (R)(-4 * SRE_PI); st_up[0] = (R)(4 * SRE_PI);  // :66-74
        st_lo[1] = (R)(-4 * SRE_PI); st_up[1] = (R)(4 * SRE_PI);
        act_lo[0] = -c.act_max; act_up[0] = c.act_max;
    }
    // _step_dynamics :106-122: th_ddot = (tau - m g l / 2 sin th - d thd) / (m l^2 / 3); symplectic Euler
    SRE_DEV static void step(const Consts &c, R dt, int flags, R *s, R *h, R *act, const R *trig) {
        const R thdd = (act[0] - c.mgl2 * trig[0] - c.d * s[1]) * c.inv_J;
        s[1] += thdd * dt;
        s[0] += s[1] * dt;
    }
    SRE_DEV static void observe(const R *s, R *obs, R *trig) {  // :93-94
        m_sincos(s[0], obs[0], obs[1]);
        obs[2] = s[1];
        trig[0] = obs[0];
        trig[1] = obs[1];
    }
    SRE_DEV static void reset_hidden(const R *dp, int flags, const R *s, R *h) {}
    SRE_DEV static void state_from_init(const R *init, int n_cols, R *s) { s[0] = init[0]; s[1] = init[1]; }
    SRE_DEV static void init_box(const R *dp, int wild_init, bool right, double *lo, double *up) {  // init_state = zeros
        lo[0] = up[0] = 0.0; lo[1] = up[1] = 0.0;
    }
};

}  // namespace sre
