// sre_policy_tc.cuh -- FNN [IN, 64, 64, OUT] (tanh, linear output) with the 64x64 hidden layer on the 5th-generation
// tensor cores (tcgen05 + TMEM), hand-written PTX for sm_100a.  Policy of policies/feed_back/fnn.py:138-160.
//
// A CTA of 128 threads owns 128 environments = the M = 128 rows of one UMMA tile:
//   layer 1  (IN <= 8 inputs)   thread-per-env FFMA, weights broadcast from shared memory
//   tanh, split h1 = hi + lo    hi = h1 with the low 13 mantissa bits cleared (a TF32 number), lo = h1 - hi (exact)
//   A tiles  [128 x 64] tf32    each thread writes ITS row into the canonical K-major SWIZZLE_128B layout (conflict-
//                               free 128-bit stores: the swizzle XORs the 16-byte chunk index with row % 8)
//   layer 2  tcgen05.mma kind::tf32, M = 128, N = 64, K = 8 per instruction, one elected thread issues
//            D  = A_lo * B_hi^T + A_hi * B_lo^T + A_hi * B_hi^T     (3 x TF32 split: fp32-level accuracy, the dropped
//            term lo*lo is < 2^-22 relative), accumulators fp32 in TMEM (64 columns), tcgen05.commit -> mbarrier
//   epilogue tcgen05.ld 32x32b: thread r reads row r (its own env): + b2, tanh, layer 3 as 64 FFMA per output
// B = W2 (torch layout W[out][in] IS the K-major [N x K] operand), split hi / lo once per CTA at staging.
// With SINGLE_PASS only A_hi * B_hi^T is issued (plain TF32, ~1e-3 relative): offered for throughput studies.
#pragma once
#include "sre_policy.cuh"

namespace sre {

SRE_DEV uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// tanh of a packed pair of prescaled arguments (see tanh_prescaled), ONE reciprocal for both lanes:
//   a_i = 1 + 2^(z_i);  r = 1 / (a_0 a_1);  1/a_0 = r a_1, 1/a_1 = r a_0;  tanh_i = 1 - 2 / a_i
// 3 MUFU per pair instead of 4 (the XU pipe, 16 lanes/clk/SM, is the scarcest unit of this kernel).  The arguments are
// clamped to 60 (tanh is already 1 - 2^-59 there, i.e. exactly 1.0f after rounding) so that a_0 a_1 <= 2^120 stays finite.
#ifndef SRE_TANH_SHARED_RCP
#define SRE_TANH_SHARED_RCP 1
#endif
SRE_DEV unsigned long long tanh2_prescaled(unsigned long long zs) {
    float z0, z1, t0, t1;
    unpack2(zs, z0, z1);
#if SRE_TANH_SHARED_RCP
    z0 = fminf(z0, 60.0f);
    z1 = fminf(z1, 60.0f);
#endif
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t0) : "f"(z0));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t1) : "f"(z1));
    float a0, a1;
    unpack2(add2(pack2(t0, t1), pack2(1.0f, 1.0f)), a0, a1);
#if SRE_TANH_SHARED_RCP
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a0 * a1));
    return fma2(pack2(r * a1, r * a0), pack2(-2.0f, -2.0f), pack2(1.0f, 1.0f));
#else
    float r0, r1;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(a0));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r1) : "f"(a1));
    return fma2(pack2(r0, r1), pack2(-2.0f, -2.0f), pack2(1.0f, 1.0f));
#endif
}

// Two packed pairs at once: the shared reciprocal pairs lane i of `za` with lane i of `zb`, so that the products and
// the two back-multiplications stay packed (a_i b_i, then 1/a_i = r_i b_i and 1/b_i = r_i a_i): per 4 tanh
// 4 FMNMX + 4 EX2 + 2 FADD2 + 3 FMUL2 + 2 RCP + 2 FFMA2 = 17 issue slots instead of 20.
#ifndef SRE_TANH_QUAD
#define SRE_TANH_QUAD 1
#endif
#ifndef SRE_TANH_RCP4
#define SRE_TANH_RCP4 1  // 1: ONE reciprocal for the four tanh (5 MUFU per quad instead of 6, three more multiplications)
#endif
SRE_DEV void tanh4_prescaled(unsigned long long &za, unsigned long long &zb) {
#if SRE_TANH_RCP4 && SRE_TANH_QUAD && SRE_TANH_SHARED_RCP
    // a_i = 1 + 2^(z_i) with z clamped to 30 (tanh is exactly 1.0f from z = 26 on), so a_0 a_1 a_2 a_3 <= 2^120 is finite:
    //   P = (a0 b0)(a1 b1),  r = 1 / P,  1/(a0 b0) = r (a1 b1),  1/(a1 b1) = r (a0 b0),  1/a_i = b_i / (a_i b_i), 1/b_i = a_i / (a_i b_i)
    float a0, a1, b0, b1, ea0, ea1, eb0, eb1, r, p0, p1;
    unpack2(za, a0, a1);
    unpack2(zb, b0, b1);
    a0 = fminf(a0, 30.0f); a1 = fminf(a1, 30.0f); b0 = fminf(b0, 30.0f); b1 = fminf(b1, 30.0f);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ea0) : "f"(a0));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ea1) : "f"(a1));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(eb0) : "f"(b0));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(eb1) : "f"(b1));
    const unsigned long long one = pack2(1.0f, 1.0f), m2 = pack2(-2.0f, -2.0f);
    const unsigned long long pa = add2(pack2(ea0, ea1), one), pb = add2(pack2(eb0, eb1), one);
    unpack2(mul2(pa, pb), p0, p1);  // p_i = a_i b_i
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(p0 * p1));
    const unsigned long long rp = mul2(pack2(r, r), pack2(p1, p0));  // {1 / p0, 1 / p1}
    za = fma2(mul2(pb, rp), m2, one);
    zb = fma2(mul2(pa, rp), m2, one);
#elif SRE_TANH_QUAD && SRE_TANH_SHARED_RCP
    float a0, a1, b0, b1, ea0, ea1, eb0, eb1, r0, r1, p0, p1;
    unpack2(za, a0, a1);
    unpack2(zb, b0, b1);
    a0 = fminf(a0, 60.0f); a1 = fminf(a1, 60.0f); b0 = fminf(b0, 60.0f); b1 = fminf(b1, 60.0f);
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ea0) : "f"(a0));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(ea1) : "f"(a1));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(eb0) : "f"(b0));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(eb1) : "f"(b1));
    const unsigned long long one = pack2(1.0f, 1.0f), m2 = pack2(-2.0f, -2.0f);
    const unsigned long long pa = add2(pack2(ea0, ea1), one), pb = add2(pack2(eb0, eb1), one);
    unpack2(mul2(pa, pb), p0, p1);
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r0) : "f"(p0));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r1) : "f"(p1));
    const unsigned long long r = pack2(r0, r1);
    za = fma2(mul2(pb, r), m2, one);
    zb = fma2(mul2(pa, r), m2, one);
#else
    za = tanh2_prescaled(za);
    zb = tanh2_prescaled(zb);
#endif
}

SRE_DEV void st_shared_v4(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// UMMA shared-memory matrix descriptor, K-major, SWIZZLE_128B (cute/arch/mma_sm100_desc.hpp SmemDescriptor):
// start address >> 4 in [0,14), LBO (ignored for swizzled K-major, 1) in [16,30), SBO = 8 rows * 128 B = 1024 >> 4
// in [32,46), version 1 in [46,48), layout type SWIZZLE_128B = 2 in [61,64)
SRE_DEV uint64_t umma_desc_k_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

// instruction descriptor (InstrDescriptor): D = F32 (1 << 4), A = B = TF32 (2 << 7, 2 << 10), both K-major,
// N >> 3 at [17,23), M >> 4 at [24,29)
constexpr uint32_t UMMA_IDESC_TF32_M128_N64 = (1u << 4) | (2u << 7) | (2u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);

SRE_DEV void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
SRE_DEV void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
SRE_DEV void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
// SRE_TC_WAIT: 0 = poll try_wait back to back, 1 = try_wait with a suspend-time hint of SRE_TC_WAIT_NS, 2 = sleep
// SRE_TC_WAIT_NS between polls (a polling warp takes issue slots from the warps that have work)
#ifndef SRE_TC_WAIT
#define SRE_TC_WAIT 0
#endif
#ifndef SRE_TC_WAIT_NS
#define SRE_TC_WAIT_NS 200
#endif
SRE_DEV void mbar_wait(uint32_t bar, uint32_t parity) {
#if SRE_TC_WAIT == 1
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}"
        ::"r"(bar), "r"(parity), "r"((uint32_t)SRE_TC_WAIT_NS)
        : "memory");
#elif SRE_TC_WAIT == 2
    uint32_t ok;
    for (;;) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (ok) break;
        __nanosleep(SRE_TC_WAIT_NS);
    }
#else
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}"
        ::"r"(bar), "r"(parity)
        : "memory");
#endif
}
SRE_DEV void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
SRE_DEV void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
SRE_DEV void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// 16 consecutive fp32 accumulator columns of this thread's TMEM lane
SRE_DEV void tmem_ld16(uint32_t taddr, float *v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}
SRE_DEV void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

template <int IN, int OUT, bool SINGLE_PASS = false> struct FnnTC {
    static constexpr int HID = 64;
    static constexpr int THREADS = 128;
    static constexpr int TMEM_COLS = 64;
    // shared memory (bytes from a 1024-aligned base)
    static constexpr int A_KB = 128 * 128;  // one K-block (32 tf32 = 128 B per row) of the A tile
    static constexpr int B_KB = 64 * 128;
    static constexpr int OFF_A_HI = 0;
    static constexpr int OFF_A_LO = OFF_A_HI + 2 * A_KB;
    static constexpr int OFF_B_HI = OFF_A_LO + 2 * A_KB;
    static constexpr int OFF_B_LO = OFF_B_HI + 2 * B_KB;
    static constexpr int OFF_W1T = OFF_B_LO + 2 * B_KB;          // float [IN][HID], prescaled
    static constexpr int OFF_B1 = OFF_W1T + IN * HID * 4;        // float [HID], prescaled
    static constexpr int OFF_B2 = OFF_B1 + HID * 4;              // float [HID], prescaled
    static constexpr int OFF_W3 = OFF_B2 + HID * 4;              // float [OUT][HID]
    static constexpr int OFF_B3 = OFF_W3 + OUT * HID * 4;        // float [OUT] (padded to 16 B)
    static constexpr int OFF_BAR = OFF_B3 + 16;                  // mbarrier (8 B) + TMEM base address (4 B)
    static constexpr int SMEM_BYTES = OFF_BAR + 16 + 1024;       // + slack for the manual 1024-byte alignment
    static constexpr float PRE = SRE_TANH_PRESCALE;              // folded into W1, b1, W2, b2 (tanh_prescaled)

    struct Ctx {
        uint32_t base;      // shared-memory address of the aligned block
        uint32_t tmem;      // TMEM base (lane 0, first column)
        uint32_t parity;    // phase of the mbarrier
        const float *gen;   // generic pointer to the aligned block (for ld.shared through C++)
    };

    SRE_DEV static uint32_t tf32_hi(float x) { return __float_as_uint(x) & 0xFFFFE000u; }

    // byte offset of element (row, k) inside an operand whose K-blocks are `kb_bytes` apart
    SRE_DEV static uint32_t sw_off(int row, int k, int kb_bytes) {
        return (uint32_t)((k >> 5) * kb_bytes + row * 128 + ((((k & 31) >> 2) ^ (row & 7)) << 4) + ((k & 3) << 2));
    }

    // Called by all THREADS threads of the CTA once: stage the weights, allocate TMEM, init the barrier.
    SRE_DEV static void setup(Ctx &c, unsigned char *smem_raw, const float *__restrict__ w) {
        const uint32_t raw = smem_u32(smem_raw);
        const uint32_t pad = ((raw + 1023u) & ~1023u) - raw;
        c.base = raw + pad;
        c.gen = reinterpret_cast<const float *>(smem_raw + pad);
        c.parity = 0;
        float *g = const_cast<float *>(c.gen);
        const float *W1 = w, *b1 = W1 + IN * HID, *W2 = b1 + HID, *b2 = W2 + HID * HID, *W3 = b2 + HID,
                    *b3 = W3 + OUT * HID;
        for (int i = threadIdx.x; i < IN * HID; i += blockDim.x) g[OFF_W1T / 4 + (i % IN) * HID + i / IN] = PRE * W1[i];
        for (int i = threadIdx.x; i < HID; i += blockDim.x) {
            g[OFF_B1 / 4 + i] = PRE * b1[i];
            g[OFF_B2 / 4 + i] = PRE * b2[i];
        }
        for (int i = threadIdx.x; i < OUT * HID; i += blockDim.x) g[OFF_W3 / 4 + i] = W3[i];
        for (int i = threadIdx.x; i < 4; i += blockDim.x) g[OFF_B3 / 4 + i] = i < OUT ? b3[i] : 0.f;
        for (int i = threadIdx.x; i < HID * HID; i += blockDim.x) {
            const int n = i / HID, k = i % HID;  // W2[n][k]: output n, input k
            const float v = PRE * W2[i];
            const float hi = __uint_as_float(tf32_hi(v));
            const uint32_t off = sw_off(n, k, B_KB);
            g[(OFF_B_HI + off) / 4] = hi;
            g[(OFF_B_LO + off) / 4] = v - hi;
        }
        if (threadIdx.x == 0) mbar_init(c.base + OFF_BAR, 1);
        if (threadIdx.x < 32) {  // one warp allocates TMEM and writes the base address to shared memory
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(c.base + OFF_BAR + 8),
                         "n"(TMEM_COLS)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
        fence_async_smem();  // weights (generic-proxy stores) become visible to the tensor core (async proxy)
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
        c.tmem = *reinterpret_cast<const volatile uint32_t *>(reinterpret_cast<const unsigned char *>(c.gen) + OFF_BAR + 8);
    }

    SRE_DEV static void teardown(Ctx &c) {
        tc_fence_before();
        __syncthreads();
        if (threadIdx.x < 32)
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(c.tmem), "n"(TMEM_COLS) : "memory");
    }

    // CTA-cooperative: EVERY thread of the CTA calls this each step (convergent).  x = observation of the thread's
    // own environment.  `active` is OR-reduced over the CTA by the barrier that hands the A tile to the tensor core;
    // the result lets the caller leave its loop CTA-uniformly.
    SRE_DEV static bool forward(Ctx &c, const float *x, float *out, bool active) {
        const int r = threadIdx.x;
        const float *sm = c.gen;
        // ---- layer 1 + tanh + hi/lo split, 16 outputs at a time -> this thread's row of A_hi / A_lo ----
#pragma unroll 1
        for (int j0 = 0; j0 < HID; j0 += 16) {
            float a[16];
            const float4 *bv = reinterpret_cast<const float4 *>(sm + OFF_B1 / 4 + j0);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const float4 b = bv[q];
                a[4 * q] = b.x; a[4 * q + 1] = b.y; a[4 * q + 2] = b.z; a[4 * q + 3] = b.w;
            }
#pragma unroll
            for (int i = 0; i < IN; i++) {
                const float4 *wv = reinterpret_cast<const float4 *>(sm + OFF_W1T / 4 + i * HID + j0);
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const float4 w = wv[q];
                    a[4 * q] = fmaf(w.x, x[i], a[4 * q]);
                    a[4 * q + 1] = fmaf(w.y, x[i], a[4 * q + 1]);
                    a[4 * q + 2] = fmaf(w.z, x[i], a[4 * q + 2]);
                    a[4 * q + 3] = fmaf(w.w, x[i], a[4 * q + 3]);
                }
            }
#pragma unroll
            for (int q = 0; q < 4; q++) {
                float h[4], hi[4];
#pragma unroll
                for (int e = 0; e < 4; e++) {
                    h[e] = tanh_prescaled(a[4 * q + e]);
                    hi[e] = __uint_as_float(tf32_hi(h[e]));
                }
                const uint32_t off = sw_off(r, j0 + 4 * q, A_KB);
                st_shared_v4(c.base + OFF_A_HI + off, hi[0], hi[1], hi[2], hi[3]);
                if (!SINGLE_PASS)
                    st_shared_v4(c.base + OFF_A_LO + off, h[0] - hi[0], h[1] - hi[1], h[2] - hi[2], h[3] - hi[3]);
            }
        }
        // ---- hand the tile to the tensor core ----
        fence_async_smem();
        tc_fence_before();
        const bool any = __syncthreads_or(active);
        if (threadIdx.x == 0) {
            tc_fence_after();
            uint32_t acc = 0;
#pragma unroll
            for (int pass = SINGLE_PASS ? 2 : 0; pass < 3; pass++) {
                const uint32_t a0 = c.base + (pass == 0 ? OFF_A_LO : OFF_A_HI);
                const uint32_t b0 = c.base + (pass == 1 ? OFF_B_LO : OFF_B_HI);
#pragma unroll
                for (int ks = 0; ks < 8; ks++) {  // K = 64 in slices of 8 tf32 (32 bytes inside the 128-byte atom)
                    const uint32_t ao = (ks >> 2) * A_KB + (ks & 3) * 32, bo = (ks >> 2) * B_KB + (ks & 3) * 32;
                    umma_tf32(c.tmem, umma_desc_k_sw128(a0 + ao), umma_desc_k_sw128(b0 + bo), UMMA_IDESC_TF32_M128_N64, acc);
                    acc = 1;
                }
            }
            umma_commit(c.base + OFF_BAR);
        }
        mbar_wait(c.base + OFF_BAR, c.parity);
        c.parity ^= 1;
        tc_fence_after();
        // ---- epilogue: this thread's row of D -> + b2, tanh, layer 3 ----
        float o[OUT];
#pragma unroll
        for (int q = 0; q < OUT; q++) o[q] = sm[OFF_B3 / 4 + q];
        const uint32_t trow = c.tmem + ((uint32_t)(r & ~31) << 16);  // warp w reads lanes 32w .. 32w + 31
#pragma unroll
        for (int j0 = 0; j0 < HID; j0 += 16) {
            float d[16];
            tmem_ld16(trow + j0, d);
            tmem_ld_wait();
            const float4 *bv = reinterpret_cast<const float4 *>(sm + OFF_B2 / 4 + j0);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const float4 b = bv[q];
                d[4 * q] = tanh_prescaled(d[4 * q] + b.x);
                d[4 * q + 1] = tanh_prescaled(d[4 * q + 1] + b.y);
                d[4 * q + 2] = tanh_prescaled(d[4 * q + 2] + b.z);
                d[4 * q + 3] = tanh_prescaled(d[4 * q + 3] + b.w);
            }
#pragma unroll
            for (int oo = 0; oo < OUT; oo++) {
                const float4 *wv = reinterpret_cast<const float4 *>(sm + OFF_W3 / 4 + oo * HID + j0);
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const float4 w = wv[q];
                    o[oo] = fmaf(w.x, d[4 * q], o[oo]);
                    o[oo] = fmaf(w.y, d[4 * q + 1], o[oo]);
                    o[oo] = fmaf(w.z, d[4 * q + 2], o[oo]);
                    o[oo] = fmaf(w.w, d[4 * q + 3], o[oo]);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < OUT; q++) out[q] = o[q];
        tc_fence_before();  // the next step's MMA overwrites the accumulators: order our TMEM reads before its barrier
        return any;
    }
};


// ---- variant 2: A operand in TENSOR MEMORY, K split in two halves ---------------------------------------------
// The shared-memory A tiles above cost 64 KB per CTA (two CTAs per SM = 8 warps: latency bound, profiles/r01_d).
// Here the activations never touch shared memory: every thread writes its row of h1 (hi and lo parts) straight into
// TMEM with tcgen05.st and the MMA reads A from TMEM (tcgen05.mma [d], [a], b_desc).  To keep a tile at 128 TMEM
// columns (D 64 + A_hi 32 + A_lo 32 -> four tiles per SM) the hidden layer is fed in two K halves of 32 units:
//   half 0: L1 + tanh + split of units 0..31 -> tcgen05.st -> barrier -> 12 MMAs (lo*hi, hi*lo, hi*hi; 4 K slices)
//   half 1: the same for units 32..63; its L1 / tanh arithmetic overlaps the MMAs of half 0, only the tcgen05.st
//           waits (mbarrier) until they have consumed the A columns
// Shared memory per CTA: W2 hi/lo (32 KB, canonical K-major SWIZZLE_128B) + W1/b1/b2/W3/b3 (2.3 KB).
SRE_DEV void tmem_st16(uint32_t taddr, const uint32_t *r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
          "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
SRE_DEV void tmem_st8(uint32_t taddr, const uint32_t *r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), "r"(r[0]),
                 "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
}
SRE_DEV void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
SRE_DEV void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

template <int IN, int OUT, bool SINGLE_PASS = false, bool CW = false> struct FnnTC2 {
    static constexpr int HID = 64;
    static constexpr int THREADS = 128;
    static constexpr int TMEM_COLS = 128;  // D [0,64), A_hi [64,96), A_lo [96,128)
    static constexpr int B_KB = 64 * 128;
    static constexpr int OFF_B_HI = 0;
    static constexpr int OFF_B_LO = OFF_B_HI + 2 * B_KB;
    static constexpr int OFF_W1T = OFF_B_LO + 2 * B_KB;
    static constexpr int OFF_B1 = OFF_W1T + IN * HID * 4;
    static constexpr int OFF_B2 = OFF_B1 + HID * 4;
    static constexpr int OFF_W3 = OFF_B2 + HID * 4;
    static constexpr int OFF_B3 = OFF_W3 + OUT * HID * 4;
    static constexpr int OFF_BAR = OFF_B3 + 16;             // two mbarriers (16 B) + TMEM base address (4 B)
    static constexpr int OFF_CNT = OFF_BAR + 24;            // two arrival counters (one per K half)
    static constexpr int SMEM_BYTES = OFF_BAR + 32 + 1024;  // + slack for the manual 1024-byte alignment
    static constexpr float PRE = SRE_TANH_PRESCALE;

    struct Ctx {
        uint32_t base, tmem, parity;
        const float *gen;
#ifdef SRE_TC_TRACE
        uint32_t *trace;  // debug build: per-warp clock stamps of the steps [SRE_TC_TRACE_S0, +8) of the first 592 CTAs
        uint32_t step;
#endif
    };
#ifdef SRE_TC_TRACE
#ifndef SRE_TC_TRACE_S0
#define SRE_TC_TRACE_S0 200
#endif
    SRE_DEV static void stamp(Ctx &c, int k) {
        if ((threadIdx.x & 31) == 0 && c.trace && blockIdx.x < 592 && c.step >= SRE_TC_TRACE_S0 && c.step < SRE_TC_TRACE_S0 + 8) {
            uint32_t v;
            if (k == 7) asm volatile("mov.u32 %0, %%smid;" : "=r"(v));
            else v = (uint32_t)clock();
            c.trace[((blockIdx.x * 8 + (c.step - SRE_TC_TRACE_S0)) * 4 + (threadIdx.x >> 5)) * 8 + k] = v;
        }
    }
#define SRE_STAMP(c, k) stamp(c, k)
#else
#define SRE_STAMP(c, k)
#endif

    SRE_DEV static uint32_t tf32_hi(float x) { return __float_as_uint(x) & 0xFFFFE000u; }
    SRE_DEV static uint32_t sw_off(int row, int k, int kb_bytes) {
        return (uint32_t)((k >> 5) * kb_bytes + row * 128 + ((((k & 31) >> 2) ^ (row & 7)) << 4) + ((k & 3) << 2));
    }

    SRE_DEV static void setup(Ctx &c, unsigned char *smem_raw, const float *__restrict__ w) {
        const uint32_t raw = smem_u32(smem_raw);
        const uint32_t pad = ((raw + 1023u) & ~1023u) - raw;
        c.base = raw + pad;
        c.gen = reinterpret_cast<const float *>(smem_raw + pad);
        c.parity = 0;
        float *g = const_cast<float *>(c.gen);
        const float *W1 = w, *b1 = W1 + IN * HID, *W2 = b1 + HID, *b2 = W2 + HID * HID, *W3 = b2 + HID,
                    *b3 = W3 + OUT * HID;
        for (int i = threadIdx.x; i < IN * HID; i += blockDim.x) g[OFF_W1T / 4 + (i % IN) * HID + i / IN] = PRE * W1[i];
        for (int i = threadIdx.x; i < HID; i += blockDim.x) {
            g[OFF_B1 / 4 + i] = PRE * b1[i];
            g[OFF_B2 / 4 + i] = PRE * b2[i];
        }
        for (int i = threadIdx.x; i < OUT * HID; i += blockDim.x) g[OFF_W3 / 4 + i] = W3[i];
        for (int i = threadIdx.x; i < 4; i += blockDim.x) g[OFF_B3 / 4 + i] = i < OUT ? b3[i] : 0.f;
        for (int i = threadIdx.x; i < HID * HID; i += blockDim.x) {
            const int n = i / HID, k = i % HID;
            const float v = PRE * W2[i];
            const float hi = __uint_as_float(tf32_hi(v));
            const uint32_t off = sw_off(n, k, B_KB);
            g[(OFF_B_HI + off) / 4] = hi;
            g[(OFF_B_LO + off) / 4] = v - hi;
        }
        if (threadIdx.x == 0) {
            mbar_init(c.base + OFF_BAR, 1);
            mbar_init(c.base + OFF_BAR + 8, 1);
            g[OFF_CNT / 4] = 0.f;  // bit pattern 0
            g[OFF_CNT / 4 + 1] = 0.f;
        }
        if (threadIdx.x < 32) {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(c.base + OFF_BAR + 16),
                         "n"(TMEM_COLS)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
        fence_async_smem();
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
        c.tmem = *reinterpret_cast<const volatile uint32_t *>(reinterpret_cast<const unsigned char *>(c.gen) + OFF_BAR + 16);
    }

    SRE_DEV static void teardown(Ctx &c) {
        tc_fence_before();
        __syncthreads();
        if (threadIdx.x < 32)
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(c.tmem), "n"(TMEM_COLS) : "memory");
    }

    // offsets (floats) of the small layers inside KPolicy.cw (launch arguments, SRE_TC_CW builds)
    static constexpr int CW_W1T = 0, CW_B1 = IN * HID, CW_B2 = CW_B1 + HID, CW_W3 = CW_B2 + HID, CW_B3 = CW_W3 + OUT * HID;
    // CW: layer 1, the biases and layer 3 are read from the LAUNCH ARGUMENTS (KPolicy.cw, constant bank) instead of shared
    // memory.  With compile-time offsets (the loops below are then fully unrolled) the weights become uniform-register
    // operands: LDCU.128 on the uniform datapath instead of 145 broadcast LDS.128 per env-step through the LSU, no vector
    // registers for them, the observation as the scalar broadcast operand of FFMA2 (R.F32, UR pair, acc) at 2.06 cycles of
    // the FMA pipe instead of 2.29.  Measured: headline 1.240e10 -> 1.506e10 env-steps/s, bit-identical results.
    using CwRef = const float (&)[SRE_TC_CW_FLOATS];
    SRE_DEV static float4 ld4(const float *sm, CwRef cw, int off_sm, int off_cw) {
        if constexpr (CW) return *reinterpret_cast<const float4 *>(&cw[off_cw]);
        else return *reinterpret_cast<const float4 *>(sm + off_sm);
    }

    // layer 1 pre-activations of 16 hidden units j0 .. j0 + 15 as 8 packed pairs; x2[i] = {x_i, x_i}
    SRE_DEV static void l1_chunk(const float *sm, CwRef cw, const unsigned long long *x2, int j0, unsigned long long *a2) {
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const float4 b = ld4(sm, cw, OFF_B1 / 4 + j0 + 4 * q, CW_B1 + j0 + 4 * q);
            a2[2 * q] = pack2(b.x, b.y);
            a2[2 * q + 1] = pack2(b.z, b.w);
        }
#pragma unroll
        for (int i = 0; i < IN; i++) {
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const float4 w = ld4(sm, cw, OFF_W1T / 4 + i * HID + j0 + 4 * q, CW_W1T + i * HID + j0 + 4 * q);
                a2[2 * q] = fma2(pack2(w.x, w.y), x2[i], a2[2 * q]);
                a2[2 * q + 1] = fma2(pack2(w.z, w.w), x2[i], a2[2 * q + 1]);
            }
        }
    }
    // tanh + hi/lo split of 4 packed pairs -> this thread's TMEM lane, 8 columns of A_hi / A_lo
    SRE_DEV static void act_store8(const unsigned long long *a2, uint32_t t_hi, uint32_t t_lo) {
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int e2 = 0; e2 < 4; e2 += 2) {
            unsigned long long hp[2] = {a2[e2], a2[e2 + 1]};
            tanh4_prescaled(hp[0], hp[1]);
#pragma unroll
            for (int k = 0; k < 2; k++) {
                const int e = e2 + k;
                const unsigned long long h = hp[k];
                const unsigned long long hh = h & 0xFFFFE000FFFFE000ull;
                const unsigned long long ll = fma2(hh, pack2(-1.0f, -1.0f), h);  // h - hi, exact
                hi[2 * e] = (uint32_t)hh; hi[2 * e + 1] = (uint32_t)(hh >> 32);
                lo[2 * e] = (uint32_t)ll; lo[2 * e + 1] = (uint32_t)(ll >> 32);
            }
        }
        tmem_st8(t_hi, hi);
        if (!SINGLE_PASS) tmem_st8(t_lo, lo);
    }

    // Hand this warp's 32 rows of the A half to the tensor core.  No CTA barrier: every warp bumps an arrival counter
    // and goes on; the warp that arrives LAST (fourth of the group) issues the 12 MMAs of the half and commits them to
    // the half's mbarrier.  Groups cannot interleave: nobody gets past the mbarrier of a half before all four arrived.
    // HALF is a compile-time constant: every descriptor is then (CTA-uniform base + constant), which the compiler keeps
    // in uniform registers.  With a run-time `half` it wrapped each of the 12 MMAs in an ELECT / R2UR.BROADCAST loop
    // (14 instructions and a branch per MMA): the tile waited ~1 240 cycles from its last arrival to the wake-up, three
    // times the 389 cycles the 12 MMAs occupy the tensor pipe (tools/tc_trace.py).
    template <int HALF> SRE_DEV static void publish_half(Ctx &c, uint32_t t_d, uint32_t t_ahi, uint32_t t_alo) {
        constexpr int half = HALF;
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        uint32_t old = 0;
        if ((threadIdx.x & 31) == 0)
            asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], 1;"
                         : "=r"(old)
                         : "r"(c.base + OFF_CNT + half * 4)
                         : "memory");
        old = __shfl_sync(0xffffffffu, old, 0);
        if ((old & 3u) == 3u) {  // warp-uniform: this warp arrived last; ONE elected lane issues (elect.sync keeps the
                                 // branch uniform for the compiler: plain UTCHMMA, no per-instruction election loop)
            uint32_t elected;
            asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(elected));
            if (elected) {
                tc_fence_after();
                uint32_t acc = half;
#pragma unroll
                for (int pass = SINGLE_PASS ? 2 : 0; pass < 3; pass++) {
                    const uint32_t a0 = pass == 0 ? t_alo : t_ahi;
                    const uint32_t b0 = c.base + (pass == 1 ? OFF_B_LO : OFF_B_HI) + half * B_KB;
#pragma unroll
                    for (int ks = 0; ks < 4; ks++) {
                        umma_tf32_ts(t_d, a0 + ks * 8, umma_desc_k_sw128(b0 + ks * 32), UMMA_IDESC_TF32_M128_N64, acc);
                        acc = 1;
                    }
                }
                umma_commit(c.base + OFF_BAR + half * 8);
            }
        }
        __syncwarp();
    }

    // CTA-cooperative: all 128 threads call this every step (whole warps; warps need not be in lockstep).
    SRE_DEV static void forward(Ctx &c, const float *x, float *out, CwRef cw) {
        const int r = threadIdx.x;
        const float *sm = c.gen;
        const uint32_t lane_base = (uint32_t)(r & ~31) << 16;
        const uint32_t t_d = c.tmem, t_ahi = c.tmem + 64, t_alo = c.tmem + 96;
        unsigned long long x2[IN];
        SRE_STAMP(c, 0);
        SRE_STAMP(c, 7);
#pragma unroll
        for (int i = 0; i < IN; i++) x2[i] = pack2(x[i], x[i]);
        // shared-memory weights: not unrolled (the hot loop stays small); launch-argument weights: unrolled, their offsets
        // have to be compile-time constants
#pragma unroll(CW ? 2 : 1)
        for (int half = 0; half < 2; half++) {
#pragma unroll(CW ? 2 : 1)
            for (int ch = 0; ch < 2; ch++) {
                unsigned long long a[8];
                l1_chunk(sm, cw, x2, half * 32 + ch * 16, a);
                if (half == 1 && ch == 0) {  // the MMAs of half 0 must have consumed the A columns
                    SRE_STAMP(c, 2);
                    mbar_wait(c.base + OFF_BAR, c.parity);
                    tc_fence_after();
                    SRE_STAMP(c, 3);
                }
                act_store8(a, lane_base + t_ahi + ch * 16, lane_base + t_alo + ch * 16);
                act_store8(a + 4, lane_base + t_ahi + ch * 16 + 8, lane_base + t_alo + ch * 16 + 8);
            }
            SRE_STAMP(c, half == 0 ? 1 : 4);
            if (half == 0) publish_half<0>(c, t_d, t_ahi, t_alo);
            else publish_half<1>(c, t_d, t_ahi, t_alo);
        }
        SRE_STAMP(c, 5);
        mbar_wait(c.base + OFF_BAR + 8, c.parity);
        c.parity ^= 1;
        tc_fence_after();
        SRE_STAMP(c, 6);
        unsigned long long o2[OUT];
#pragma unroll
        for (int q = 0; q < OUT; q++) o2[q] = pack2(CW ? cw[CW_B3 + q] : sm[OFF_B3 / 4 + q], 0.0f);
#pragma unroll(CW ? 4 : 1)
        for (int j0 = 0; j0 < HID; j0 += 16) {
            float d[16];
            tmem_ld16(lane_base + t_d + j0, d);
            tmem_ld_wait();
            unsigned long long h2[8];
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const float4 b = ld4(sm, cw, OFF_B2 / 4 + j0 + 4 * q, CW_B2 + j0 + 4 * q);
                h2[2 * q] = add2(pack2(d[4 * q], d[4 * q + 1]), pack2(b.x, b.y));
                h2[2 * q + 1] = add2(pack2(d[4 * q + 2], d[4 * q + 3]), pack2(b.z, b.w));
                tanh4_prescaled(h2[2 * q], h2[2 * q + 1]);
            }
#pragma unroll
            for (int oo = 0; oo < OUT; oo++) {
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const float4 w = ld4(sm, cw, OFF_W3 / 4 + oo * HID + j0 + 4 * q, CW_W3 + oo * HID + j0 + 4 * q);
                    o2[oo] = fma2(pack2(w.x, w.y), h2[2 * q], o2[oo]);
                    o2[oo] = fma2(pack2(w.z, w.w), h2[2 * q + 1], o2[oo]);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < OUT; q++) {
            float e, od;
            unpack2(o2[q], e, od);
            out[q] = e + od;
        }
        tc_fence_before();  // orders this step's TMEM reads before the arrival that lets the next MMAs overwrite them
#ifdef SRE_TC_TRACE
        c.step++;
#endif
    }
};

// ---- [IN, 16, 16, OUT] in the production (float) build: thread per environment, everything in registers ----------
// The warp-tiled layout above pays for its activation exchange through shared memory and the transpose-reduce with
// ~730 instructions per forward pass at HID = 16 (profiles/r02_c_c4_*: 57 % of the C4 step for 416 multiply-adds).  At
// this width a lane holds the whole network state itself: 8 packed accumulator pairs per layer, weights as broadcast
// LDS.128 of the transposed matrices (four per input unit), the activation as the scalar operand of FFMA2, tanh four at a
// time with one shared reciprocal.  64 + 128 + 8 OUT FFMA2, 2 x 16 tanh, ~110 LDS.128: ~450 instructions.
// CW: the weights are read from the launch arguments (KPolicy.cw, same layout) as uniform-register operands instead of
// broadcast LDS.128 (see FnnTC2).
template <int IN, int OUT, bool CW> struct FnnLane16 {
    static constexpr bool TAKES_CW = true;  // forward() takes KPolicy.cw (the tiled variants do not)
    static constexpr int HID = 16;
    static constexpr int OP = (OUT + 1) / 2;                 // output pairs
    static constexpr int OFF_W1T = 0;                        // [IN][16]
    static constexpr int OFF_B1 = OFF_W1T + IN * HID;        // [16]
    static constexpr int OFF_W2T = OFF_B1 + HID;             // [16 k][16 j]
    static constexpr int OFF_B2 = OFF_W2T + HID * HID;       // [16]
    static constexpr int OFF_W3T = OFF_B2 + HID;             // [16 k][2 OP]
    static constexpr int OFF_B3 = OFF_W3T + HID * 2 * OP;    // [2 OP]
    static constexpr int W_FLOATS = ((OFF_B3 + 2 * OP + 3) / 4) * 4;
    static constexpr int H1S_FLOATS = 0;
    static constexpr int WARPS = 8;
    static constexpr int SMEM_FLOATS = W_FLOATS;
    static constexpr float PRE = SRE_TANH_PRESCALE;

    SRE_DEV static void stage(const float *__restrict__ w, float *sm) {
        const float *W1 = w, *b1 = W1 + IN * HID, *W2 = b1 + HID, *b2 = W2 + HID * HID, *W3 = b2 + HID,
                    *b3 = W3 + OUT * HID;
        for (int i = threadIdx.x; i < IN * HID; i += blockDim.x) sm[OFF_W1T + (i % IN) * HID + i / IN] = PRE * W1[i];
        for (int i = threadIdx.x; i < HID * HID; i += blockDim.x) sm[OFF_W2T + (i % HID) * HID + i / HID] = PRE * W2[i];
        for (int i = threadIdx.x; i < HID; i += blockDim.x) { sm[OFF_B1 + i] = PRE * b1[i]; sm[OFF_B2 + i] = PRE * b2[i]; }
        for (int i = threadIdx.x; i < HID * 2 * OP; i += blockDim.x) {
            const int k = i / (2 * OP), o = i % (2 * OP);
            sm[OFF_W3T + i] = o < OUT ? W3[o * HID + k] : 0.f;
        }
        for (int i = threadIdx.x; i < 2 * OP; i += blockDim.x) sm[OFF_B3 + i] = i < OUT ? b3[i] : 0.f;
    }

    using CwRef = const float (&)[SRE_TC_CW_FLOATS];
    SRE_DEV static float4 ld4(const float *sm, CwRef cw, int off) {
        if constexpr (CW) return *reinterpret_cast<const float4 *>(&cw[off]);
        else return *reinterpret_cast<const float4 *>(sm + off);
    }
    // one layer into 16 units: acc[8] = {b[2p], b[2p+1]} + sum_i wT[i][2p..2p+1] * x_i  (offsets in floats)
    template <int NI>
    SRE_DEV static void layer16(const float *sm, CwRef cw, int off_w, int off_b, const float *x, unsigned long long *acc) {
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const float4 v = ld4(sm, cw, off_b + 4 * q);
            acc[2 * q] = pack2(v.x, v.y);
            acc[2 * q + 1] = pack2(v.z, v.w);
        }
#pragma unroll
        for (int i = 0; i < NI; i++) {
            const unsigned long long xi = pack2(x[i], x[i]);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const float4 w = ld4(sm, cw, off_w + i * HID + 4 * q);
                acc[2 * q] = fma2(pack2(w.x, w.y), xi, acc[2 * q]);
                acc[2 * q + 1] = fma2(pack2(w.z, w.w), xi, acc[2 * q + 1]);
            }
        }
    }
    SRE_DEV static void tanh16(unsigned long long *a, float *h) {
#pragma unroll
        for (int q = 0; q < 4; q++) {
            tanh4_prescaled(a[2 * q], a[2 * q + 1]);
            unpack2(a[2 * q], h[4 * q], h[4 * q + 1]);
            unpack2(a[2 * q + 1], h[4 * q + 2], h[4 * q + 3]);
        }
    }

    // same interface as the tiled variants (h1s / lane unused): x = observation of the lane's own environment
    SRE_DEV static void forward(const float *sm, float *, const float *x, float *out, int, CwRef cw) {
        unsigned long long a[8];
        float h1[16], h2[16];
        layer16<IN>(sm, cw, OFF_W1T, OFF_B1, x, a);
        tanh16(a, h1);
        layer16<HID>(sm, cw, OFF_W2T, OFF_B2, h1, a);
        tanh16(a, h2);
        unsigned long long o[OP];
#pragma unroll
        for (int p = 0; p < OP; p++) o[p] = CW ? pack2(cw[OFF_B3 + 2 * p], cw[OFF_B3 + 2 * p + 1]) : pack2(sm[OFF_B3 + 2 * p], sm[OFF_B3 + 2 * p + 1]);
#pragma unroll
        for (int k = 0; k < HID; k++) {
            const unsigned long long hk = pack2(h2[k], h2[k]);
#pragma unroll
            for (int p = 0; p < OP; p++) {
                const int off = OFF_W3T + k * 2 * OP + 2 * p;
                const unsigned long long w = CW ? pack2(cw[off], cw[off + 1]) : pack2(sm[off], sm[off + 1]);
                o[p] = fma2(w, hk, o[p]);
            }
        }
#pragma unroll
        for (int p = 0; p < OP; p++) {
            float lo, hi;
            unpack2(o[p], lo, hi);
            out[2 * p] = lo;
            if (2 * p + 1 < OUT) out[2 * p + 1] = hi;
        }
    }
};
template <int IN, int OUT> struct FnnTiled<IN, 16, OUT, true> : FnnLane16<IN, OUT, false> {};

}  // namespace sre
