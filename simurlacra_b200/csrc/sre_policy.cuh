// sre_policy.cuh -- policies fused into the rollout: one thread evaluates the policy of its own environment,
// weights are staged once per CTA in shared memory and read with warp-broadcast loads (all lanes read the same
// address, so a 128-bit load feeds four FMAs per lane).
//   FNN     policies/feed_back/fnn.py:138-160  (Linear -> nonlin per hidden layer, Linear output)
//   Linear  policies/feed_back/linear.py:95-110 + policies/features.py:127-175
// float32 throughout, like torch's default dtype; summation order = bias first, inputs ascending.
#pragma once
#include "sre_device.cuh"

namespace sre {

#define SRE_TC_CW_FLOATS (8 * 64 + 64 + 64 + 2 * 64 + 8)
struct KPolicy {
    int kind, n_layers;
    int sizes[SRE_MAX_LAYERS + 1];
    int n_feats;
    int feats[SRE_MAX_FEATS];
    int hidden_nonlin, output_nonlin, cartpole_feats, precision;
    const float *weights;  // device, torch layout
    long long n_weights;
    int smem_floats;  // size of the staged (re-laid-out) copy
    long long group_size;  // > 0: lane n uses weights + (n / group_size) * n_weights (parameter exploration)
    int cta_groups;        // group_size is a multiple of the CTA size: one parameter set per CTA, staged like a single set
    int act_noise;    // SRE_ACT_NOISE_*
    float noise_scale[SRE_MAX_DA], noise_mean[SRE_MAX_DA];
    int n_lin_feats;                       // Linear policy: total number of feature values (columns of W)
    int feat_arg[SRE_MAX_FEATS];           // offset of feature k's parameter block in fparams
    float fparams[SRE_MAX_FEAT_PARAMS];    // parameters of the parametrised features (host-validated copy)
    // tensor-core policy: W1^T / b1 / b2 (prescaled), W3, b3 as LAUNCH ARGUMENTS (constant bank -> uniform-register operands)
    // when the caller has the parameters on the host (sre_policy.weights_host); layout of FnnTC2's shared-memory block
    int has_cw;
    alignas(16) float cw[SRE_TC_CW_FLOATS];
};

SRE_DEV float nonlin_f(int kind, float x) {
    if (kind == 0) return tanhf(x);
    if (kind == 1) return x > 0.f ? x : 0.f;
    return x;
}

// ---- warp-tiled FNN [IN, HID, HID, OUT], tanh hidden, linear output -------------------------------------------
// The dominant cost of a rollout step is layer 2 (HID x HID MACs per env).  A thread-per-env dot product needs one
// warp-broadcast shared-memory load per 4 FMAs and a broadcast load only delivers 8 bytes per shared-memory
// wavefront (measured: profiles/r01_a_*), which caps the kernel well below the FMA roofline.  Here a warp treats
// its 32 environments as a [32 x HID] x [HID x HID] product with an 8 x (HID/8) register tile per lane:
//   lane = (r, c), r = lane / 8 owns environments 8r .. 8r+7, c = lane % 8 owns outputs {16m + 2c, 16m + 2c + 1}
//   per k: 2 x LDS.128 of activations (h1s[k][env], written once per step by the owner lanes)
//          HID/16 x LDS.128 of weights stored DUPLICATED {w, w} so that they feed packed FP32 FMAs directly
//          4 x (HID/8) FFMA2 (Blackwell fma.rn.f32x2: two environments per instruction, one issue slot)
// Layer 1 (IN <= 8 inputs) and the dynamics stay thread-per-env; layer 3 is folded into the epilogue of layer 2
// and reduced across the 8 lanes of a row group with a transpose-reduce (7 shuffles), which leaves the action of
// environment e on lane e.
// tanh(z) = 1 - 2 / (1 + e^(2z)).  The caller passes zs = 2 log2(e) z (the scale is folded into the weights and
// biases of the producing layer when they are staged), so this is MUFU.EX2, FADD, MUFU.RCP, FFMA: four
// instructions, absolute error < 3e-7; saturates correctly (ex2 -> inf gives rcp -> 0, ex2 -> 0 gives -1).
#define SRE_TANH_PRESCALE 2.8853900817779268f
SRE_DEV float tanh_prescaled(float zs) {
    float t, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(zs));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + t));
    return fmaf(-2.0f, r, 1.0f);
}

template <int IN, int HID, int OUT, bool FAST_TANH> struct FnnTiled {
    static constexpr bool TAKES_CW = false;
    static constexpr int TN = HID / 8;  // outputs per lane
    static constexpr int NM = TN / 2;   // 16-output chunks
    static_assert(HID % 16 == 0 && HID <= 64, "hidden width must be 16, 32 or 64");
    // block-shared weights (floats)
    static constexpr int OFF_W1T = 0;                      // [IN][HID]
    static constexpr int OFF_B1 = OFF_W1T + IN * HID;      // [HID]
    static constexpr int OFF_W2D = OFF_B1 + HID;           // [HID k][HID j][2] duplicated
    static constexpr int OFF_B2 = OFF_W2D + 2 * HID * HID; // [HID]
    static constexpr int OFF_W3D = OFF_B2 + HID;           // [OUT][HID][2] duplicated
    static constexpr int OFF_B3 = OFF_W3D + 2 * OUT * HID; // [OUT]
    static constexpr int W_FLOATS = ((OFF_B3 + OUT + 3) / 4) * 4;
    static constexpr int H1S_FLOATS = HID * 32;            // per warp: h1s[k][env]
    static constexpr int WARPS = FAST_TANH ? 8 : 4;        // production build: 256-thread CTAs, 2 per SM
    static constexpr int SMEM_FLOATS = W_FLOATS + WARPS * H1S_FLOATS;
    static constexpr float PRE = FAST_TANH ? SRE_TANH_PRESCALE : 1.0f;  // folded into W1, b1, W2, b2

    SRE_DEV static float act(float x) { return FAST_TANH ? tanh_prescaled(x) : tanhf(x); }

    SRE_DEV static void stage(const float *__restrict__ w, float *sm) {
        const float *W1 = w, *b1 = W1 + IN * HID, *W2 = b1 + HID, *b2 = W2 + HID * HID, *W3 = b2 + HID,
                    *b3 = W3 + OUT * HID;
        for (int i = threadIdx.x; i < IN * HID; i += blockDim.x) sm[OFF_W1T + (i % IN) * HID + i / IN] = PRE * W1[i];
        for (int i = threadIdx.x; i < HID * HID; i += blockDim.x) {
            const int j = i / HID, k = i % HID;  // W2[j][k]
            sm[OFF_W2D + 2 * (k * HID + j)] = PRE * W2[i];
            sm[OFF_W2D + 2 * (k * HID + j) + 1] = PRE * W2[i];
        }
        for (int i = threadIdx.x; i < HID; i += blockDim.x) { sm[OFF_B1 + i] = PRE * b1[i]; sm[OFF_B2 + i] = PRE * b2[i]; }
        for (int i = threadIdx.x; i < OUT * HID; i += blockDim.x) { sm[OFF_W3D + 2 * i] = W3[i]; sm[OFF_W3D + 2 * i + 1] = W3[i]; }
        for (int i = threadIdx.x; i < OUT; i += blockDim.x) sm[OFF_B3 + i] = b3[i];
    }

    // Warp-cooperative: every lane of the warp must call this (convergent), x = observation of the lane's own env.
    SRE_DEV static void forward(const float *sm, float *h1s, const float *x, float *out, int lane) {
        // ---- layer 1, thread-per-env, 16 outputs at a time -> h1s[j][lane] ----
        __syncwarp();  // previous step's readers are done with h1s
#pragma unroll 1
        for (int j0 = 0; j0 < HID; j0 += 16) {
            float a[16];
            const float4 *bv = reinterpret_cast<const float4 *>(sm + OFF_B1 + j0);
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const float4 b = bv[q];
                a[4 * q] = b.x; a[4 * q + 1] = b.y; a[4 * q + 2] = b.z; a[4 * q + 3] = b.w;
            }
#pragma unroll
            for (int i = 0; i < IN; i++) {
                const float4 *wv = reinterpret_cast<const float4 *>(sm + OFF_W1T + i * HID + j0);
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
            for (int q = 0; q < 16; q++) h1s[(j0 + q) * 32 + lane] = act(a[q]);
        }
        __syncwarp();
        // ---- layer 2, 8 envs x TN outputs per lane, packed over environment pairs ----
        const int r = lane >> 3, c = lane & 7;
        unsigned long long acc[4][TN];
#pragma unroll
        for (int m = 0; m < NM; m++) {
            const float2 b = *reinterpret_cast<const float2 *>(sm + OFF_B2 + 16 * m + 2 * c);
#pragma unroll
            for (int ip = 0; ip < 4; ip++) {
                acc[ip][2 * m] = pack2(b.x, b.x);
                acc[ip][2 * m + 1] = pack2(b.y, b.y);
            }
        }
        const float *ap = h1s + 8 * r;
        const float *wp = sm + OFF_W2D + 4 * c;
#pragma unroll 8
        for (int k = 0; k < HID; k++) {
            const ulonglong2 a01 = *reinterpret_cast<const ulonglong2 *>(ap + k * 32);
            const ulonglong2 a23 = *reinterpret_cast<const ulonglong2 *>(ap + k * 32 + 4);
            const unsigned long long a[4] = {a01.x, a01.y, a23.x, a23.y};
#pragma unroll
            for (int m = 0; m < NM; m++) {
                const ulonglong2 w = *reinterpret_cast<const ulonglong2 *>(wp + 2 * (k * HID + 16 * m));
#pragma unroll
                for (int ip = 0; ip < 4; ip++) {
                    acc[ip][2 * m] = fma2(a[ip], w.x, acc[ip][2 * m]);
                    acc[ip][2 * m + 1] = fma2(a[ip], w.y, acc[ip][2 * m + 1]);
                }
            }
        }
        // ---- epilogue: tanh, layer 3 partial sums over this lane's TN outputs ----
        unsigned long long part[OUT][4];
#pragma unroll
        for (int o = 0; o < OUT; o++)
#pragma unroll
            for (int ip = 0; ip < 4; ip++) part[o][ip] = 0ull;
#pragma unroll
        for (int m = 0; m < NM; m++) {
#pragma unroll
            for (int ip = 0; ip < 4; ip++) {
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    float lo, hi;
                    unpack2(acc[ip][2 * m + h], lo, hi);
                    acc[ip][2 * m + h] = pack2(act(lo), act(hi));
                }
            }
#pragma unroll
            for (int o = 0; o < OUT; o++) {
                const ulonglong2 w3 = *reinterpret_cast<const ulonglong2 *>(sm + OFF_W3D + 2 * (o * HID + 16 * m + 2 * c));
#pragma unroll
                for (int ip = 0; ip < 4; ip++) {
                    part[o][ip] = fma2(acc[ip][2 * m], w3.x, part[o][ip]);
                    part[o][ip] = fma2(acc[ip][2 * m + 1], w3.y, part[o][ip]);
                }
            }
        }
        // ---- transpose-reduce over the 8 lanes of the row group: lane (r, c) ends with env 8r + c ----
#pragma unroll
        for (int o = 0; o < OUT; o++) {
            float v[8];
#pragma unroll
            for (int ip = 0; ip < 4; ip++) unpack2(part[o][ip], v[2 * ip], v[2 * ip + 1]);
            float u[4];
            const bool b4 = c & 4;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const float send = b4 ? v[i] : v[4 + i], keep = b4 ? v[4 + i] : v[i];
                u[i] = keep + __shfl_xor_sync(0xffffffffu, send, 4);
            }
            float t[2];
            const bool b2 = c & 2;
#pragma unroll
            for (int i = 0; i < 2; i++) {
                const float send = b2 ? u[i] : u[2 + i], keep = b2 ? u[2 + i] : u[i];
                t[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
            }
            const bool b1 = c & 1;
            const float send = b1 ? t[0] : t[1], keep = b1 ? t[1] : t[0];
            out[o] = sm[OFF_B3 + o] + (keep + __shfl_xor_sync(0xffffffffu, send, 1));
        }
    }
};

// ---- generic FNN: any depth <= SRE_MAX_LAYERS, any width <= SRE_MAX_WIDTH; activations in local memory ----
SRE_DEV void stage_generic(const KPolicy &pol, const float *w, float *sm) {
    for (long long i = threadIdx.x; i < pol.n_weights; i += blockDim.x) sm[i] = w[i];
}

SRE_DEV void fnn_generic_forward(const KPolicy &pol, const float *sm, const float *obs, int n_obs, float *act,
                                 int n_act) {
    float bufa[SRE_MAX_WIDTH], bufb[SRE_MAX_WIDTH];
    float *cur = bufa, *nxt = bufb;
    int n_in = n_obs;
    if (pol.cartpole_feats) {  // fnn.py:214-218: [x, th, xd, thd] -> [x, sin th, cos th, xd, thd]
        cur[0] = obs[0]; cur[1] = sinf(obs[1]); cur[2] = cosf(obs[1]); cur[3] = obs[2]; cur[4] = obs[3];
        n_in = 5;
    } else {
        for (int i = 0; i < n_in; i++) cur[i] = obs[i];
    }
    const float *w = sm;
    for (int L = 0; L < pol.n_layers; L++) {
        const int ni = pol.sizes[L], no = pol.sizes[L + 1];
        const float *W = w, *b = w + ni * no;
        for (int o = 0; o < no; o++) {
            float s = b[o];
            for (int i = 0; i < ni; i++) s = fmaf(W[o * ni + i], cur[i], s);
            if (L < pol.n_layers - 1) s = nonlin_f(pol.hidden_nonlin, s);
            else if (pol.output_nonlin == 1) s = tanhf(s);
            nxt[o] = s;
        }
        w += ni * no + no;
        float *t = cur; cur = nxt; nxt = t;
    }
    for (int i = 0; i < n_act; i++) act[i] = cur[i];
}

// ---- Linear policy on a feature stack -------------------------------------------------------------------
SRE_DEV float feat_f(int kind, float x) {
    switch (kind) {
    case SRE_FEAT_IDENTITY: return x;
    case SRE_FEAT_SIGN: return (float)((x > 0.f) - (x < 0.f));
    case SRE_FEAT_ABS: return fabsf(x);
    case SRE_FEAT_SQUARED: return x * x;
    case SRE_FEAT_CUBIC: return x * x * x;
    case SRE_FEAT_SIN: return sinf(x);
    case SRE_FEAT_COS: return cosf(x);
    case SRE_FEAT_SINSIN: { const float s = sinf(x); return s * s; }
    default: return sinf(x) * cosf(x);
    }
}

// act = W feats(obs), W [n_act][n_lin_feats] row-major, the feature values formed on the fly in stack order
// (LinearPolicy.forward linear.py:95-110 over FeatureStack.__call__ features.py:66-75)
SRE_DEV void linear_accumulate(const KPolicy &pol, const float *w, const float *obs, int n_obs, float *act, int n_act) {
    const int nf = pol.n_lin_feats;
    for (int o = 0; o < n_act; o++) act[o] = 0.f;
    int col = 0;
    auto add = [&](float f) {
        for (int o = 0; o < n_act; o++) act[o] = fmaf(w[o * nf + col], f, act[o]);
        col++;
    };
    for (int k = 0; k < pol.n_feats; k++) {
        const int fk = pol.feats[k];
        const float *p = pol.fparams + pol.feat_arg[k];
        if (fk == SRE_FEAT_CONST) {
            add(1.0f);
        } else if (fk <= SRE_FEAT_SINCOS) {
            for (int i = 0; i < n_obs; i++) add(feat_f(fk, obs[i]));
        } else if (fk == SRE_FEAT_SIG) {
            for (int i = 0; i < n_obs; i++) add(1.0f / (1.0f + expf(-(p[0] * obs[i]))));
        } else if (fk == SRE_FEAT_BELL) {
            for (int i = 0; i < n_obs; i++) add(expf(-p[0] * (obs[i] * obs[i])));
        } else if (fk == SRE_FEAT_MULT) {
            const int n = (int)p[0];
            float v = obs[(int)p[1]];
            for (int j = 1; j < n; j++) v *= obs[(int)p[1 + j]];
            add(v);
        } else if (fk == SRE_FEAT_ATAN2) {
            add(atan2f(obs[(int)p[0]], obs[(int)p[1]]));
        } else if (fk == SRE_FEAT_RF) {
            const int D = (int)p[0];
            const float *freq = p + 2, *shift = p + 2 + D * n_obs;
            for (int d = 0; d < D; d++) {
                float s = 0.f;
                for (int i = 0; i < n_obs; i++) s = fmaf(obs[i], freq[d * n_obs + i], s);
                add(p[1] * cosf(s + shift[d]));
            }
        } else {  // SRE_FEAT_RBF
            const int F = (int)p[0];
            const bool state_wise = p[1] != 0.f;
            const float *scale = p + 2, *cen = p + 2 + n_obs;
            float tot = 0.f;
            if (!state_wise) {
                for (int i = 0; i < n_obs; i++)
                    for (int f = 0; f < F; f++) {
                        const float d = obs[i] - cen[f * n_obs + i];
                        tot += expf(-scale[i] * (d * d));
                    }
                if (!(tot > 1e-8f)) tot = 1.0f;
            }
            for (int i = 0; i < n_obs; i++) {
                float nrm = tot;
                if (state_wise) {
                    nrm = 0.f;
                    for (int f = 0; f < F; f++) {
                        const float d = obs[i] - cen[f * n_obs + i];
                        nrm += expf(-scale[i] * (d * d));
                    }
                    if (!(nrm > 1e-8f)) nrm = 1.0f;
                }
                for (int f = 0; f < F; f++) {
                    const float d = obs[i] - cen[f * n_obs + i];
                    add(expf(-scale[i] * (d * d)) / nrm);
                }
            }
        }
    }
}

template <int N_OBS, int N_ACT>
SRE_DEV void linear_forward(const KPolicy &pol, const float *sm, const float *obs, float *act) {
    linear_accumulate(pol, sm, obs, N_OBS, act, N_ACT);
}

SRE_DEV void linear_forward_rt(const KPolicy &pol, const float *w, const float *obs, int n_obs, float *act, int n_act) {
    linear_accumulate(pol, w, obs, n_obs, act, n_act);
}

}  // namespace sre
