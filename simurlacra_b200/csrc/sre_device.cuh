// sre_device.cuh -- device-side building blocks shared by all kernels: math wrappers templated on the real
// type (float = production build, double = verification build), Philox4x32-10 and the sample mappings of the
// engine's RNG specification (restated independently by the CPU checker under oracle/, pinned by the Random123 vectors).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sre.h"

#define SRE_DEV __device__ __forceinline__
#define SRE_HD __host__ __device__ __forceinline__  // also evaluated on the host (constants passed as launch arguments)
#define SRE_PI 3.14159265358979323846

namespace sre {

// ------------------------------------------------------------------------------------------------------
// math: one overload set per real type, so that the env code reads like the reference's formulas
// ------------------------------------------------------------------------------------------------------
SRE_DEV float m_sin(float x) { return sinf(x); }
SRE_DEV double m_sin(double x) { return sin(x); }
SRE_DEV float m_cos(float x) { return cosf(x); }
SRE_DEV double m_cos(double x) { return cos(x); }
// float sincos of the production build: three-term Cody-Waite reduction by pi/2 with FMAs, then the two degree-7 / degree-8
// minimax polynomials on [-pi/4, pi/4] evaluated with CONSTANT coefficients and swapped by quadrant parity (sincosf()
// selects coefficients per call and carries a Payne-Hanek slow path: ~48 instructions and three divergence regions
// against 26 straight-line ones here).  Max error 1.6 ulp for |x| <= 1e5 (checked against float64 on 1e7 samples),
// growing beyond ~1e6 rad, a range no simulated angle reaches; non-finite x gives NaN like sincosf.
#ifndef SRE_LIBM_SINCOS
SRE_DEV void m_sincos(float x, float &s, float &c) {
    const float q = rintf(x * 0.63661977236758138f);
    const int i = (int)q;
    float r = fmaf(q, -0x1.921fb6p+0f, x);
    r = fmaf(q, 0x1.777a5cp-25f, r);
    r = fmaf(q, 0x1.ee59dap-50f, r);
    const float z = r * r;
    float ps = fmaf(fmaf(-1.9515295891e-4f, z, 8.3321608736e-3f), z, -1.6666654611e-1f);
    ps = fmaf(ps * z, r, r);
    float pc = fmaf(fmaf(2.443315711809948e-5f, z, -1.388731625493765e-3f), z, 4.166664568298827e-2f);
    pc = fmaf(pc * z, z, fmaf(-0.5f, z, 1.0f));
    const bool odd = i & 1;
    s = __int_as_float(__float_as_int(odd ? pc : ps) ^ ((i & 2) << 30));
    c = __int_as_float(__float_as_int(odd ? ps : pc) ^ (((i + 1) & 2) << 30));
}
#else
SRE_DEV void m_sincos(float x, float &s, float &c) { sincosf(x, &s, &c); }
#endif
SRE_DEV void m_sincos(double x, double &s, double &c) { sincos(x, &s, &c); }
// reciprocal of a well-scaled positive quantity (mass-matrix determinants, cos of a plate angle): MUFU.RCP + one Newton
// step, <= 1 ulp, without the range check / slow path that `1.0f / x` carries (5 instructions against ~10 and a
// divergence region).  The verification (double) build divides.
SRE_DEV float m_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    const float e = fmaf(-x, r, 1.0f);
    return (e == e) ? fmaf(r, e, r) : r;  // x = 0 or inf: keep rcp's inf / 0 like the division
}
SRE_DEV double m_rcp(double x) { return 1.0 / x; }
SRE_DEV float m_exp(float x) { return expf(x); }
SRE_DEV double m_exp(double x) { return exp(x); }
SRE_DEV float m_log(float x) { return logf(x); }
SRE_DEV double m_log(double x) { return log(x); }
SRE_DEV float m_sqrt(float x) { return sqrtf(x); }
SRE_DEV double m_sqrt(double x) { return sqrt(x); }
SRE_DEV float m_fmod(float x, float y) { return fmodf(x, y); }
SRE_DEV double m_fmod(double x, double y) { return fmod(x, y); }
// fmod for quotients of a few units (angle errors against a 2 pi period): q = trunc(x / y) from one multiplication by
// the host-computed reciprocal, corrected by at most one, then r = x - q y in ONE fma.  The true remainder is exactly
// representable whenever |r| < |y| (it is a multiple of the smaller ulp and shorter than 24 bits), so the fma returns
// fmodf's value bit for bit (the sign of a zero result aside) for |x / y| < 2^22 -- at a dozen straight-line
// instructions instead of fmodf's division loop and its branches.  y > 0.
SRE_DEV float m_fmod_pos(float x, float y, float inv_y) {
    float q = truncf(x * inv_y);
    const float r0 = fmaf(-q, y, x);
    const bool neg = x < 0.0f;
    // wrong side of zero: |q| was one too large; not below |y|: one too small (r0 may be rounded there, the test is not)
    const bool too_big = neg ? (r0 > 0.0f) : (r0 < 0.0f);
    const bool too_small = neg ? (r0 <= -y) : (r0 >= y);
    const float one = neg ? -1.0f : 1.0f;
    q = too_big ? q - one : (too_small ? q + one : q);
    return fmaf(-q, y, x);
}
SRE_DEV double m_fmod_pos(double x, double y, double inv_y) { return fmod(x, y); }
SRE_DEV float m_abs(float x) { return fabsf(x); }
SRE_DEV double m_abs(double x) { return fabs(x); }
SRE_DEV float m_min(float x, float y) { return fminf(x, y); }
SRE_DEV double m_min(double x, double y) { return fmin(x, y); }
SRE_DEV float m_max(float x, float y) { return fmaxf(x, y); }
SRE_DEV double m_max(double x, double y) { return fmax(x, y); }
SRE_DEV float m_floor(float x) { return floorf(x); }
SRE_DEV double m_floor(double x) { return floor(x); }
SRE_DEV float shfl_r(float v, int src) { return __shfl_sync(0xffffffffu, v, src); }
SRE_DEV double shfl_r(double v, int src) {
    const int lo = __shfl_sync(0xffffffffu, __double2loint(v), src), hi = __shfl_sync(0xffffffffu, __double2hiint(v), src);
    return __hiloint2double(hi, lo);
}
template <typename R> SRE_DEV R m_sign(R x) { return (R)((x > (R)0) - (x < (R)0)); }
template <typename R> SRE_DEV bool m_isnan(R x) { return x != x; }

// ------------------------------------------------------------------------------------------------------
// packed FP32 pairs (Blackwell add / mul / fma .f32x2: two values per issue slot).  Cost on a scheduler's FMA pipe
// (profiles/r02_pipe_microbench.json): 2.0 cycles per warp instruction with an immediate / uniform / two-register form, 2.3
// with one reused register, 3.1 with three changing registers.
// ------------------------------------------------------------------------------------------------------
SRE_DEV unsigned long long pack2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
SRE_DEV void unpack2(unsigned long long v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
SRE_DEV unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// packed FP32 pairs (Blackwell add/mul/fma .f32x2: two lanes per issue slot)
SRE_DEV unsigned long long add2(unsigned long long a, unsigned long long b) {
    unsigned long long d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
SRE_DEV unsigned long long mul2(unsigned long long a, unsigned long long b) {
    unsigned long long d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
struct P2 {
    unsigned long long v;
};
SRE_DEV P2 p2(float a, float b) { return P2{pack2(a, b)}; }
SRE_DEV P2 p2s(float a) { return P2{pack2(a, a)}; }
SRE_DEV P2 operator+(P2 a, P2 b) { return P2{add2(a.v, b.v)}; }
SRE_DEV P2 operator*(P2 a, P2 b) { return P2{mul2(a.v, b.v)}; }
SRE_DEV P2 pfma(P2 a, P2 b, P2 c) { return P2{fma2(a.v, b.v, c.v)}; }
SRE_DEV P2 psub(P2 a, P2 b) { return P2{fma2(b.v, pack2(-1.0f, -1.0f), a.v)}; }  // a - b, exact product
SRE_DEV void pun(P2 a, float &x, float &y) { unpack2(a.v, x, y); }

// m_sincos for a pair.  q = rint(x 2/pi) comes from the magic-number addition (the integer sits in the low mantissa bits
// of t, two's complement, so its parity / bit 1 are read from there): no FRND / F2I.
SRE_DEV void sincos2(P2 x, P2 &s, P2 &c) {
    const P2 magic = p2s(12582912.0f);
    const P2 t = pfma(x, p2s(0.63661977236758138f), magic);
    const P2 q = psub(t, magic);
    P2 r = pfma(q, p2s(-0x1.921fb6p+0f), x);
    r = pfma(q, p2s(0x1.777a5cp-25f), r);
    r = pfma(q, p2s(0x1.ee59dap-50f), r);
    const P2 z = r * r;
    P2 ps = pfma(pfma(p2s(-1.9515295891e-4f), z, p2s(8.3321608736e-3f)), z, p2s(-1.6666654611e-1f));
    ps = pfma(ps * z, r, r);
    P2 pc = pfma(pfma(p2s(2.443315711809948e-5f), z, p2s(-1.388731625493765e-3f)), z, p2s(4.166664568298827e-2f));
    pc = pfma(pc * z, z, pfma(p2s(-0.5f), z, p2s(1.0f)));
    float ps0, ps1, pc0, pc1;
    pun(ps, ps0, ps1);
    pun(pc, pc0, pc1);
    const int i0 = (int)(unsigned)t.v, i1 = (int)(unsigned)(t.v >> 32);
    const bool o0 = i0 & 1, o1 = i1 & 1;
    const float s0 = __int_as_float(__float_as_int(o0 ? pc0 : ps0) ^ ((i0 & 2) << 30));
    const float c0 = __int_as_float(__float_as_int(o0 ? ps0 : pc0) ^ (((i0 + 1) & 2) << 30));
    const float s1 = __int_as_float(__float_as_int(o1 ? pc1 : ps1) ^ ((i1 & 2) << 30));
    const float c1 = __int_as_float(__float_as_int(o1 ? ps1 : pc1) ^ (((i1 + 1) & 2) << 30));
    s = p2(s0, s1);
    c = p2(c0, c1);
}
SRE_DEV P2 rcp2_neg(P2 nx) {  // 1 / (-nx) for a pair (m_rcp): two MUFU seeds (operand negation is free there), one packed Newton step
    float x0, x1, r0, r1;
    pun(nx, x0, x1);
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(-x0));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r1) : "f"(-x1));
    const P2 r = p2(r0, r1);
    return pfma(r, pfma(nx, r, p2s(1.0f)), r);
}


// ------------------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11; Random123 constants)
// Engine addressing: key = (seed_lo, seed_hi); counter = (env_lo, env_hi, episode, stream << 24 | block)
// ------------------------------------------------------------------------------------------------------
#define SRE_STREAM_PARAMS 0u
#define SRE_STREAM_INIT 1u
#define SRE_STREAM_OBS_NOISE 2u
#define SRE_STREAM_ACT_NOISE 3u
#define SRE_STREAM_ENV_ACT_NOISE 4u

struct u32x4 {
    uint32_t x, y, z, w;
};

SRE_DEV u32x4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return u32x4{c0, c1, c2, c3};
}

SRE_DEV u32x4 rng_block(uint64_t seed, uint64_t env, uint32_t episode, uint32_t stream, uint32_t block) {
    return philox4x32_10((uint32_t)env, (uint32_t)(env >> 32), episode, (stream << 24) | (block & 0xFFFFFFu),
                         (uint32_t)seed, (uint32_t)(seed >> 32));
}

// numpy-style 53-bit uniform in [0, 1)
SRE_DEV double u53(uint32_t a, uint32_t b) {
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) * (1.0 / 9007199254740992.0);
}
SRE_DEV float u24(uint32_t a) { return (float)(a >> 8) * (1.0f / 16777216.0f); }
// Standard normal for a domain parameter (engine RNG specification, restated in fp64 by the CPU checker):
//   u1 = (2 (a >> 9) + 1) / 2^24 in (0, 1),  x = ((int32) b >> 8) / 2^23 in [-1, 1),  z = sqrt(-2 ln u1) cos(pi x).
// Both inputs are EXACT float32 numbers (odd 24-bit integer, signed 24-bit integer), so float32 logf / sqrtf / cospif
// reproduce the fp64 evaluation to a few float32 ulp -- an episode start costs no fp64 instructions (the fp64 pipe
// of sm_100 made resets the most expensive rows of the kernel).  |z| <= sqrt(48 ln 2) = 5.77.
SRE_DEV float normal_f(uint32_t a, uint32_t b) {
    const float u1 = (float)(2u * (a >> 9) + 1u) * (1.0f / 16777216.0f);
    const float x = (float)((int32_t)b >> 8) * (1.0f / 8388608.0f);
    return __fmul_rn(sqrtf(-2.0f * logf(u1)), cospif(x));
}

// Two standard normals from one (a, b) word pair, same exactly-representable inputs as normal_f: r cos(pi x), r sin(pi x)
SRE_DEV void normal2_f(uint32_t a, uint32_t b, float &z0, float &z1) {
    const float u1 = (float)(2u * (a >> 9) + 1u) * (1.0f / 16777216.0f);
    const float x = (float)((int32_t)b >> 8) * (1.0f / 8388608.0f);
    const float r = sqrtf(-2.0f * logf(u1));
    float sn, cs;
    sincospif(x, &sn, &cs);
    z0 = __fmul_rn(r, cs);
    z1 = __fmul_rn(r, sn);
}

// the same pair in the working type: float32 math in the production build, fp64 in the verification build
SRE_DEV void normal2_r(uint32_t a, uint32_t b, float &z0, float &z1) { normal2_f(a, b, z0, z1); }
SRE_DEV void normal2_r(uint32_t a, uint32_t b, double &z0, double &z1) {
    const double u1 = (double)(2u * (a >> 9) + 1u) * (1.0 / 16777216.0);
    const double x = (double)((int32_t)b >> 8) * (1.0 / 8388608.0);
    const double r = sqrt(-2.0 * log(u1));
    double sn, cs;
    sincospi(x, &sn, &cs);
    z0 = r * cs;
    z1 = r * sn;
}

// DomainParam.sample for one table entry (domain_parameter.py:104-132): float32 draw like torch's default
// dtype -> clamp -> optional round.  Separate roundings (no FMA contraction): loc + scale * eps.
SRE_DEV float sample_entry(const sre_distr &d, uint64_t seed, uint64_t env, uint32_t episode, uint32_t entry) {
    u32x4 x = rng_block(seed, env, episode, SRE_STREAM_PARAMS, entry);
    float s;
    if (d.kind == SRE_DISTR_NORMAL) {
        float z = normal_f(x.x, x.y);
        s = __fadd_rn(d.a, __fmul_rn(d.b, z));
    } else if (d.kind == SRE_DISTR_UNIFORM) {
        float u = u24(x.x);
        s = __fadd_rn(d.a, __fmul_rn(u, __fsub_rn(d.b, d.a)));
    } else {
        float u = u24(x.x);
        float bern = (u < d.c) ? 1.0f : 0.0f;
        s = __fadd_rn(__fmul_rn(1.0f - bern, d.a), __fmul_rn(bern, d.b));
    }
    s = fminf(fmaxf(s, d.clip_lo), d.clip_up);
    if (d.roundint) s = rintf(s);
    return s;
}

// four standard normals from one Philox block (two Box-Muller pairs), observation / action noise
template <typename R> SRE_DEV void normal4(const u32x4 &x, R z[4]);
template <> SRE_DEV void normal4<double>(const u32x4 &x, double z[4]) {
    double u1 = ((double)x.x + 1.0) / 4294967296.0, u2 = (double)x.y / 4294967296.0;
    double u3 = ((double)x.z + 1.0) / 4294967296.0, u4 = (double)x.w / 4294967296.0;
    double r1 = sqrt(-2 * log(u1)), r2 = sqrt(-2 * log(u3));
    double s1, c1, s2, c2;
    sincospi(2.0 * u2, &s1, &c1);
    sincospi(2.0 * u4, &s2, &c2);
    z[0] = r1 * c1; z[1] = r1 * s1; z[2] = r2 * c2; z[3] = r2 * s2;
}
template <> SRE_DEV void normal4<float>(const u32x4 &x, float z[4]) {
    // same uniforms as the fp64 form, rounded to float (u1 in (0, 1], u2 in [0, 1])
    const float k = 1.0f / 4294967296.0f;
    float u1 = ((float)x.x + 1.0f) * k, u2 = (float)x.y * k;
    float u3 = ((float)x.z + 1.0f) * k, u4 = (float)x.w * k;
    float r1 = sqrtf(-2.0f * logf(u1)), r2 = sqrtf(-2.0f * logf(u3));
    float s1, c1, s2, c2;
    sincospif(2.0f * u2, &s1, &c1);
    sincospif(2.0f * u4, &s2, &c2);
    z[0] = r1 * c1; z[1] = r1 * s1; z[2] = r2 * c2; z[3] = r2 * s2;
}

}  // namespace sre
