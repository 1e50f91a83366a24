// sre_policy.cuh -- policies fused into the rollout: one thread evaluates the policy of its own environment,
// weights are staged once per CTA in shared memory and read with warp-broadcast loads (all lanes read the same
// address, so a 128-bit load feeds four FMAs per lane).
//   FNN     policies/feed_back/fnn.py:138-160  (Linear -> nonlin per hidden layer, Linear output)
//   Linear  policies/feed_back/linear.py:95-110 + policies/features.py:127-175
// float32 throughout, like torch's default dtype; summation order = bias first, inputs ascending.
#pragma once
#include "sre_device.cuh"

namespace sre {

struct KPolicy {
    int kind, n_layers;
    int sizes[SRE_MAX_LAYERS + 1];
    int n_feats;
    int feats[SRE_MAX_FEATS];
    int hidden_nonlin, output_nonlin, cartpole_feats;
    const float *weights;  // device, torch layout
    long long n_weights;
    int smem_floats;  // size of the staged (re-laid-out) copy
};

SRE_DEV float nonlin_f(int kind, float x) {
    if (kind == 0) return tanhf(x);
    if (kind == 1) return x > 0.f ? x : 0.f;
    return x;
}

// ---- shared-memory layout -----------------------------------------------------------------------------
// Generic path: the torch layout unchanged.
// Fast path FnnFast<IN,HID,OUT>: every layer transposed to [in][out] so that four consecutive OUTPUT neurons of
// one input are a single 128-bit broadcast load:  W1T[IN][HID] | b1[HID] | W2T[HID][HID] | b2[HID] | W3[OUT][HID] | b3[OUT]
template <int IN, int HID, int OUT> struct FnnFast {
    static constexpr int OFF_W1T = 0;
    static constexpr int OFF_B1 = OFF_W1T + IN * HID;
    static constexpr int OFF_W2T = OFF_B1 + HID;
    static constexpr int OFF_B2 = OFF_W2T + HID * HID;
    static constexpr int OFF_W3 = OFF_B2 + HID;
    static constexpr int OFF_B3 = OFF_W3 + OUT * HID;
    static constexpr int SMEM_FLOATS = ((OFF_B3 + OUT + 3) / 4) * 4;
    static constexpr int N_WEIGHTS = IN * HID + HID + HID * HID + HID + OUT * HID + OUT;
    static constexpr int JC = 8;  // output neurons of layer 2 accumulated together (independent FMA chains)

    // all threads of the CTA cooperate; caller syncs afterwards
    SRE_DEV static void stage(const float *__restrict__ w, float *sm) {
        const float *W1 = w, *b1 = W1 + IN * HID, *W2 = b1 + HID, *b2 = W2 + HID * HID, *W3 = b2 + HID,
                    *b3 = W3 + OUT * HID;
        for (int i = threadIdx.x; i < IN * HID; i += blockDim.x) {
            const int j = i / IN, k = i % IN;  // W1[j][k]
            sm[OFF_W1T + k * HID + j] = W1[i];
        }
        for (int i = threadIdx.x; i < HID * HID; i += blockDim.x) {
            const int j = i / HID, k = i % HID;  // W2[j][k]
            sm[OFF_W2T + k * HID + j] = W2[i];
        }
        for (int i = threadIdx.x; i < HID; i += blockDim.x) { sm[OFF_B1 + i] = b1[i]; sm[OFF_B2 + i] = b2[i]; }
        for (int i = threadIdx.x; i < OUT * HID; i += blockDim.x) sm[OFF_W3 + i] = W3[i];
        for (int i = threadIdx.x; i < OUT; i += blockDim.x) sm[OFF_B3 + i] = b3[i];
    }

    // tanh hidden layers, linear output
    SRE_DEV static void forward(const float *sm, const float *x, float *out) {
        float h1[HID];
        const float4 *b1v = reinterpret_cast<const float4 *>(sm + OFF_B1);
#pragma unroll
        for (int j = 0; j < HID / 4; j++) {
            const float4 b = b1v[j];
            h1[4 * j] = b.x; h1[4 * j + 1] = b.y; h1[4 * j + 2] = b.z; h1[4 * j + 3] = b.w;
        }
#pragma unroll
        for (int i = 0; i < IN; i++) {
            const float4 *wv = reinterpret_cast<const float4 *>(sm + OFF_W1T + i * HID);
#pragma unroll
            for (int j = 0; j < HID / 4; j++) {
                const float4 w = wv[j];
                h1[4 * j] = fmaf(w.x, x[i], h1[4 * j]);
                h1[4 * j + 1] = fmaf(w.y, x[i], h1[4 * j + 1]);
                h1[4 * j + 2] = fmaf(w.z, x[i], h1[4 * j + 2]);
                h1[4 * j + 3] = fmaf(w.w, x[i], h1[4 * j + 3]);
            }
        }
#pragma unroll
        for (int j = 0; j < HID; j++) h1[j] = tanhf(h1[j]);
#pragma unroll
        for (int o = 0; o < OUT; o++) out[o] = sm[OFF_B3 + o];
        // layers 2 and 3 fused: h2[j] is consumed by the output layer as soon as its dot product is complete
#pragma unroll 1
        for (int j0 = 0; j0 < HID; j0 += JC) {
            float acc[JC];
            {
                const float4 *bv = reinterpret_cast<const float4 *>(sm + OFF_B2 + j0);
#pragma unroll
                for (int q = 0; q < JC / 4; q++) {
                    const float4 b = bv[q];
                    acc[4 * q] = b.x; acc[4 * q + 1] = b.y; acc[4 * q + 2] = b.z; acc[4 * q + 3] = b.w;
                }
            }
            const float *w2 = sm + OFF_W2T + j0;
#pragma unroll
            for (int k = 0; k < HID; k++) {
                const float4 *wv = reinterpret_cast<const float4 *>(w2 + k * HID);
#pragma unroll
                for (int q = 0; q < JC / 4; q++) {
                    const float4 w = wv[q];
                    acc[4 * q] = fmaf(w.x, h1[k], acc[4 * q]);
                    acc[4 * q + 1] = fmaf(w.y, h1[k], acc[4 * q + 1]);
                    acc[4 * q + 2] = fmaf(w.z, h1[k], acc[4 * q + 2]);
                    acc[4 * q + 3] = fmaf(w.w, h1[k], acc[4 * q + 3]);
                }
            }
#pragma unroll
            for (int jj = 0; jj < JC; jj++) acc[jj] = tanhf(acc[jj]);
#pragma unroll
            for (int o = 0; o < OUT; o++) {
                const float4 *wv = reinterpret_cast<const float4 *>(sm + OFF_W3 + o * HID + j0);
#pragma unroll
                for (int q = 0; q < JC / 4; q++) {
                    const float4 w = wv[q];
                    out[o] = fmaf(w.x, acc[4 * q], out[o]);
                    out[o] = fmaf(w.y, acc[4 * q + 1], out[o]);
                    out[o] = fmaf(w.z, acc[4 * q + 2], out[o]);
                    out[o] = fmaf(w.w, acc[4 * q + 3], out[o]);
                }
            }
        }
    }
};

// ---- generic FNN: any depth <= SRE_MAX_LAYERS, any width <= SRE_MAX_WIDTH; activations in local memory ----
SRE_DEV void stage_generic(const KPolicy &pol, float *sm) {
    for (long long i = threadIdx.x; i < pol.n_weights; i += blockDim.x) sm[i] = pol.weights[i];
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

template <int N_OBS, int N_ACT>
SRE_DEV void linear_forward(const KPolicy &pol, const float *sm, const float *obs, float *act) {
    int nf = 0;
    for (int k = 0; k < pol.n_feats; k++) nf += (pol.feats[k] == SRE_FEAT_CONST) ? 1 : N_OBS;
#pragma unroll
    for (int o = 0; o < N_ACT; o++) act[o] = 0.f;
    int col = 0;
    for (int k = 0; k < pol.n_feats; k++) {
        const int fk = pol.feats[k];
        if (fk == SRE_FEAT_CONST) {
#pragma unroll
            for (int o = 0; o < N_ACT; o++) act[o] = fmaf(sm[o * nf + col], 1.0f, act[o]);
            col += 1;
        } else {
#pragma unroll
            for (int i = 0; i < N_OBS; i++) {
                const float f = feat_f(fk, obs[i]);
#pragma unroll
                for (int o = 0; o < N_ACT; o++) act[o] = fmaf(sm[o * nf + col + i], f, act[o]);
            }
            col += N_OBS;
        }
    }
}

SRE_DEV void linear_forward_rt(const KPolicy &pol, const float *w, const float *obs, int n_obs, float *act, int n_act) {
    int nf = 0;
    for (int k = 0; k < pol.n_feats; k++) nf += (pol.feats[k] == SRE_FEAT_CONST) ? 1 : n_obs;
    for (int o = 0; o < n_act; o++) act[o] = 0.f;
    int col = 0;
    for (int k = 0; k < pol.n_feats; k++) {
        const int fk = pol.feats[k];
        if (fk == SRE_FEAT_CONST) {
            for (int o = 0; o < n_act; o++) act[o] = fmaf(w[o * nf + col], 1.0f, act[o]);
            col += 1;
        } else {
            for (int i = 0; i < n_obs; i++) {
                const float f = feat_f(fk, obs[i]);
                for (int o = 0; o < n_act; o++) act[o] = fmaf(w[o * nf + col + i], f, act[o]);
            }
            col += n_obs;
        }
    }
}

}  // namespace sre
