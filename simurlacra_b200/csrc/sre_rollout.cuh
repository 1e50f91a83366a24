// sre_rollout.cuh -- the persistent-horizon rollout kernel: one thread owns one environment lane for the whole
// launch.  Replaces the hot loop of rollout() (sampling/rollout.py:185-312) run for N envs by
// ParallelRolloutSampler.sample (sampling/parallel_rollout_sampler.py:248-323):
//   per lane: [DR sample] -> reset -> loop { obs (+noise) -> policy (fp32) -> [ActNorm] -> SimPyEnv.step } -> closing row
// State, derived constants and policy activations stay in registers; policy weights sit in shared memory;
// trajectory rows are staged per warp in shared memory and written with 128-bit coalesced stores.
#pragma once
#include "sre_core.cuh"
#include "sre_policy.cuh"
#include "sre_policy_tc.cuh"
#include <cstdlib>
#include <type_traits>

namespace sre {

// policy variants compiled per (env, dtype)
#define SRE_PV_ACTIONS 0  // injected actions or idle
#define SRE_PV_GENERIC 1  // generic FNN
#define SRE_PV_LINEAR 2
#define SRE_PV_FNN16 16
#define SRE_PV_FNN32 32
#define SRE_PV_FNN64 64
#define SRE_PV_FNN16W 116  // SRE_PV_FNN16 (float build) with the weights as launch arguments (sre_policy.weights_host given)
#ifndef SRE_COOP_MAX_LANES
#define SRE_COOP_MAX_LANES 5  // up to this many restarting lanes of a warp draw their parameters cooperatively
#endif
#define SRE_PV_TC 1000    // [do,64,64,da] tanh, hidden layer on the tensor cores, 3 x TF32 (sre_policy_tc.cuh)
#define SRE_PV_TC1 1001   // the same, single TF32 pass
#define SRE_PV_TCW 1002   // SRE_PV_TC with layer 1 / biases / layer 3 as launch arguments (sre_policy.weights_host given)

struct KIO {
    long long n_envs, n_rows;
    unsigned long long seed, env_offset;
    unsigned episode;
    int stop_on_done, auto_reset;
    const void *params, *init, *actions;
    int n_init_cols;
    void *traj;
    long long traj_stride, traj_off;  // record / flag buffers are [T, traj_stride, ...], lane n writes column traj_off + n
    unsigned char *row_flag;
    void *ret;
    int *length;
    void *params_out;
    double *stats;
    int *status;
    int idle_policy;
    int prep_slots;  // host: shared memory holds one prepared episode start per lane (auto-reset sampling, see the kernel)
    int stage_bufs;  // host: 1, or 2 = the warp's record tile is double-buffered and leaves with cp.async.bulk (see the kernel)
};

template <int KIND, typename R> struct Rec {
    using E = Env<KIND, R>;
    static constexpr int OFF_OBS = 0;
    static constexpr int OFF_ACT = E::DO;
    static constexpr int OFF_ACT_APP = OFF_ACT + E::DA;
    static constexpr int OFF_REW = OFF_ACT_APP + E::DA;
    static constexpr int OFF_STATE = OFF_REW + 1;
    static constexpr int OFF_HIDDEN = OFF_STATE + E::DS;
    static constexpr int W = OFF_HIDDEN + E::NH;
};

// cooperative store of one warp's row tile: n_valid contiguous reals from shared to global memory
template <typename R> SRE_DEV void warp_store_row(const R *stage, R *g, int n_valid, bool vec_ok, int lane) {
    constexpr int V = 16 / sizeof(R);
    if (vec_ok) {
        const int nv = n_valid / V;
        const float4 *s4 = reinterpret_cast<const float4 *>(stage);
        float4 *g4 = reinterpret_cast<float4 *>(g);
        for (int i = lane; i < nv; i += 32) __stcs(g4 + i, s4[i]);  // streaming: written once, not re-read here
        for (int i = nv * V + lane; i < n_valid; i += 32) g[i] = stage[i];
    } else {
        for (int i = lane; i < n_valid; i += 32) g[i] = stage[i];
    }
}

// Bulk (TMA) store of a warp's record tile: ONE instruction issued by one lane moves the `bytes` contiguous bytes from
// shared to global memory asynchronously (UBLKCP in SASS), instead of a 6-trip LDS.128 / STG.128 loop on every lane.
// bytes % 16 == 0, both addresses 16-byte aligned.  The writers fence their generic-proxy stores first (fence_async_smem).
SRE_DEV void bulk_store_tile(void *gdst, const void *ssrc, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
// wait until at most `N` of this thread's bulk groups are still READING shared memory
template <int N> SRE_DEV void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }

SRE_DEV void atomic_min_double(double *addr, double v) {
    unsigned long long *a = reinterpret_cast<unsigned long long *>(addr);
    unsigned long long old = *a, assumed;
    do {
        assumed = old;
        if (__longlong_as_double((long long)assumed) <= v) break;
        old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v));
    } while (assumed != old);
}
SRE_DEV void atomic_max_double(double *addr, double v) {
    unsigned long long *a = reinterpret_cast<unsigned long long *>(addr);
    unsigned long long old = *a, assumed;
    do {
        assumed = old;
        if (__longlong_as_double((long long)assumed) >= v) break;
        old = atomicCAS(a, assumed, (unsigned long long)__double_as_longlong(v));
    } while (assumed != old);
}

// CTA shape per variant: the tiled FNN production build runs 256-thread CTAs, two per SM (16 warps, 128 registers)
#ifndef SRE_TC_BLOCKS
#define SRE_TC_BLOCKS 4  // resident CTAs (= UMMA tiles) per SM of the tensor-core variant: 128 TMEM columns each
#endif
#ifndef SRE_FNN_BLOCKS
#define SRE_FNN_BLOCKS 2  // resident 256-thread CTAs per SM of the fused-FNN (CUDA-core) variants of the float build
#endif
#ifndef SRE_FNN16_THREADS
#define SRE_FNN16_THREADS 256
#endif
#ifndef SRE_FNN16_BLOCKS
#define SRE_FNN16_BLOCKS 2
#endif
template <typename R, int PV> struct LaunchShape {
    static constexpr bool TC = PV >= SRE_PV_TC;  // one CTA = one 128-row UMMA tile
    static constexpr bool LANE16 = (PV == SRE_PV_FNN16 || PV == SRE_PV_FNN16W) && sizeof(R) == 4;  // register-resident FNN16: no warp cooperation
    static constexpr int THREADS = TC ? 128 : (LANE16 ? SRE_FNN16_THREADS : ((PV >= 16 && sizeof(R) == 4) ? 256 : 128));
    static constexpr int MIN_BLOCKS =
        TC ? SRE_TC_BLOCKS : (LANE16 ? SRE_FNN16_BLOCKS : ((PV >= 16 && sizeof(R) == 4) ? SRE_FNN_BLOCKS : 1));
};

template <int KIND, typename R> SRE_DEV float qbb_plate_ang(R arm_radius, R plate_length, R th_in) {
    if constexpr (KIND == SRE_QBB) return Env<SRE_QBB, R>::plate_ang(arm_radius, plate_length, th_in);
    else return 0.f;
}

// ---- stateful step-path wrappers inside the persistent kernel (sre_step_wrappers, include/sre.h) --------------------
// One wrapper state per environment lane, alive for the whole launch: the delay queue, the held action and step counter
// of the downsampling, its observation ring, the velocity-filter memory, the augmented parameters.  Only the generic
// body of the un-tiled policy variants carries it (WRAP below); the state lives in local memory, nothing of it touches
// the specialised hot bodies.
template <typename R, int DO, int DA> struct WrapState {
    R dq[SRE_MAX_DELAY * DA];
    R hold[DA];
    R ob[SRE_MAX_DOWN * DO];
    R vz[SRE_MAX_VF];
    R aug[SRE_MAX_AUG];
    int dq_head, cnt, ob_n, ob_head;
};
template <typename R, int DO, int DA> SRE_DEV void wrap_init(WrapState<R, DO, DA> &w) {
    for (int i = 0; i < SRE_MAX_DELAY * DA; i++) w.dq[i] = (R)0;
    for (int i = 0; i < DA; i++) w.hold[i] = (R)0;
    for (int i = 0; i < SRE_MAX_VF; i++) w.vz[i] = (R)0;
    for (int i = 0; i < SRE_MAX_AUG; i++) w.aug[i] = (R)0;
    w.dq_head = w.cnt = w.ob_n = w.ob_head = 0;
}
// ActDelayWrapper._process_act (action_delay.py:102-110): append the action, pop the oldest
template <typename R, int DO, int DA> SRE_DEV void wrap_delay(const KWrap<R> &k, WrapState<R, DO, DA> &w, R *act) {
    for (int i = 0; i < DA; i++) {
        const R old = w.dq[w.dq_head * DA + i];
        w.dq[w.dq_head * DA + i] = act[i];
        act[i] = old;
    }
    w.dq_head = (w.dq_head + 1 == k.delay) ? 0 : w.dq_head + 1;
}
// The observation path for ONE new observation of the lane (episode start or after a step): ObsVelFiltWrapper
// (observation_velfilter.py:119-145; init_filter :107-117 at a reset), DownsamplingWrapper._process_obs
// (downsampling.py:134-141; reset :118-131: the first observation is filtered with what the previous episode left in the
// buffer and the result seeds the new one), ObsPartialWrapper (observation_partial.py:63-75), StateAugmentationWrapper
// (state_augmentation.py:75-81).  Also resets the action-side state at an episode start.
template <typename R, int DO, int DA, int DS>
SRE_DEV int wrap_observe(const KWrap<R> &k, WrapState<R, DO, DA> &w, const R *state, const R *obs, bool reset, R *out) {
    R v[DO];
#pragma unroll
    for (int i = 0; i < DO; i++) v[i] = obs[i];
    if (reset) {
        for (int i = 0; i < SRE_MAX_DELAY * DA; i++) w.dq[i] = (R)0;
        w.dq_head = 0;
        w.cnt = 0;
    }
    for (int j = 0; j < k.vf_n; j++) {
        R x = (R)0;
#pragma unroll
        for (int i = 0; i < DS; i++)
            if (i == k.vf_pos[j]) x = state[i];
        if (reset) w.vz[j] = k.vf_zi * x;
        const R y = k.vf_b0 * x + w.vz[j];
        w.vz[j] = k.vf_b1 * x - k.vf_a1 * y;
#pragma unroll
        for (int i = 0; i < DO; i++)
            if (i == k.vf_vel[j]) v[i] = y;
    }
    if (k.down > 1) {
        if (w.ob_n < k.down) {
            const int at = (w.ob_head + w.ob_n) % k.down;
            for (int i = 0; i < DO; i++) w.ob[at * DO + i] = v[i];
            w.ob_n++;
        } else {
            for (int i = 0; i < DO; i++) w.ob[w.ob_head * DO + i] = v[i];
            w.ob_head = (w.ob_head + 1) % k.down;
        }
        for (int i = 0; i < DO; i++) {
            R sum = w.ob[w.ob_head * DO + i];  // oldest first, like np.mean over the deque
            for (int q = 1; q < w.ob_n; q++) sum += w.ob[((w.ob_head + q) % k.down) * DO + i];
            v[i] = sum / (R)w.ob_n;
        }
        if (reset) {
            for (int i = 0; i < DO; i++) w.ob[i] = v[i];
            w.ob_head = 0;
            w.ob_n = 1;
        }
    }
    int n = 0;
    for (int i = 0; i < DO; i++)
        if (k.keep_mask == 0 || ((k.keep_mask >> i) & 1)) out[n++] = v[i];
    for (int j = 0; j < k.n_aug; j++) out[n++] = k.aug_fixed ? k.aug_nominal[j] : w.aug[j];
    return n;
}

// MODE (compile time) -- what the step body has to be able to do:
//   SRE_MODE_GENERIC  everything decided at run time (records, observation noise / normalisation, action noises, dense
//                     Q / R, FinalRewTask): the kernel of round 1
//   SRE_MODE_PLAIN    no records, no row flags, none of the optional step-path wrappers, diagonal Q / R, no final reward:
//                     the body of the sampling hot path (returns + lengths + statistics) with none of those branches
//   SRE_MODE_PLAIN_REC  the same with trajectory records and row flags always on
// The host picks the mode from the launch arguments (launch_rollout_pv); results are identical by construction.
#define SRE_MODE_GENERIC 0
#define SRE_MODE_PLAIN 1
#define SRE_MODE_PLAIN_REC 2

template <int KIND, typename R, int PV, int MODE = SRE_MODE_GENERIC>
__global__ void __launch_bounds__((LaunchShape<R, PV>::THREADS), (LaunchShape<R, PV>::MIN_BLOCKS))
rollout_kernel(const __grid_constant__ KCfg<R> cfg, const __grid_constant__ KDistr distr,
               const __grid_constant__ KPolicy pol, const __grid_constant__ KIO io) {
    using E = Env<KIND, R>;
    using RC = Rec<KIND, R>;
    constexpr int DS = E::DS, DO = E::DO, DA = E::DA, DP = E::DP, NH = E::NH, W = RC::W;
    constexpr bool PLAIN = MODE != SRE_MODE_GENERIC;
    constexpr bool WRAP = MODE == SRE_MODE_GENERIC && PV < 16;  // the bodies that can run the stateful step wrappers
    constexpr int NIN = WRAP ? SRE_MAX_POL_IN : DO;
    // verification build (double) keeps the accurate tanhf; the production build uses the 2-MUFU form
    using Fast = std::conditional_t<PV == SRE_PV_FNN16W, FnnLane16<DO, DA, true>,
                                    FnnTiled<DO, ((PV >= 16 && PV <= 64) ? PV : 16), DA, sizeof(R) == 4>>;
    constexpr bool TC = PV >= SRE_PV_TC;
    using Tc = FnnTC2<DO, DA, PV == SRE_PV_TC1, PV == SRE_PV_TCW>;
    typename Tc::Ctx tc;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *sm_w = reinterpret_cast<float *>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool has_stage = !(PV >= SRE_PV_TC) || (MODE == SRE_MODE_GENERIC ? io.traj != nullptr : MODE == SRE_MODE_PLAIN_REC);
    const int n_bufs = io.stage_bufs == 2 ? 2 : 1;
    const size_t stage_bytes = has_stage ? (size_t)n_bufs * (blockDim.x / 32) * 32 * W * sizeof(R) : 0;
    R *const stage0 = reinterpret_cast<R *>(smem_raw + (size_t)pol.smem_floats * sizeof(float)) + (size_t)warp * n_bufs * 32 * W;
    constexpr int NSLOT = Lane<KIND, R>::NSLOT;
    // prepared episode starts: [warp][field][lane]
    R *slot = reinterpret_cast<R *>(smem_raw + (size_t)pol.smem_floats * sizeof(float) + stage_bytes) +
              (size_t)warp * 32 * NSLOT + (threadIdx.x & 31);
    const size_t slot_bytes = io.prep_slots ? (size_t)(blockDim.x / 32) * 32 * NSLOT * sizeof(R) : 0;
    float *h1s = sm_w + Fast::W_FLOATS + warp * Fast::H1S_FLOATS;  // only meaningful for the tiled FNN variants

    // parameter exploration with CTA-aligned groups: this CTA's parameter set is staged like a single set
    const float *w_cta = pol.weights;
    if (pol.cta_groups) w_cta += ((long long)blockIdx.x * blockDim.x / pol.group_size) * pol.n_weights;
    const bool per_lane_w = pol.group_size > 0 && !pol.cta_groups;
#ifdef SRE_TC_TRACE
    tc.trace = reinterpret_cast<uint32_t *>(io.status);
    tc.step = 0;
#endif
    if (TC) Tc::setup(tc, smem_raw + (size_t)pol.smem_floats * sizeof(float) + stage_bytes + slot_bytes, w_cta);  // the tensor-core block sits behind the staging area and the slots (pol.smem_floats = 0)
    else if (PV >= 16) Fast::stage(w_cta, sm_w);
    else if ((PV == SRE_PV_GENERIC || PV == SRE_PV_LINEAR) && !per_lane_w) stage_generic(pol, w_cta, sm_w);
    __syncthreads();

    const long long N = io.n_envs;
    const long long T = io.n_rows;
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long warp_env0 = n - lane;
    const bool valid = n < N;
    const uint64_t genv = io.env_offset + (uint64_t)n;
    const int n_valid_lanes = (int)min((long long)32, max((long long)0, N - warp_env0));
    const long long NS = io.traj_stride, NO = io.traj_off;
    const bool vec_ok = ((NS * W * (long long)sizeof(R)) % 16 == 0) && ((NO * W * (long long)sizeof(R)) % 16 == 0) &&
                        ((reinterpret_cast<size_t>(io.traj) % 16) == 0);
    const bool record = MODE == SRE_MODE_GENERIC ? io.traj != nullptr : MODE == SRE_MODE_PLAIN_REC;
    const bool want_flag = MODE == SRE_MODE_GENERIC ? io.row_flag != nullptr : MODE == SRE_MODE_PLAIN_REC;

    enum { RUN = 0, CLOSING = 1, IDLE = 2 };
    Lane<KIND, R> L;
    R obs[DO];
#pragma unroll
    for (int i = 0; i < DO; i++) obs[i] = (R)0;
    uint32_t episode = io.episode;
    int lane_state = valid ? RUN : IDLE;
    int t_ep = 0;
    R ret_ep = (R)0;
    bool first_episode = true;
    bool failed_ep = false;
    int status = 0;
    // stateful step wrappers (only in the bodies that carry them; `wrapped` is launch-uniform)
    const bool wrapped = WRAP && cfg.wrap.active;
    WrapState<R, DO, DA> ws;
    R wobs[NIN];  // the observation the policy sees (and the record keeps) when `wrapped`
    int n_in = DO;
    if (WRAP) {
        if (wrapped) wrap_init(ws);
#pragma unroll
        for (int i = 0; i < NIN; i++) wobs[i] = (R)0;
    }
    // statistics: accumulated per CTA in shared memory when an episode closes (rare), so that no statistics
    // registers stay live across the hot loop; one global atomic set per CTA at the end
    __shared__ double sm_stats[SRE_N_STATS];
    if (threadIdx.x < SRE_N_STATS)
        sm_stats[threadIdx.x] = threadIdx.x == SRE_STAT_MIN_RET ? 1e300 : (threadIdx.x == SRE_STAT_MAX_RET ? -1e300 : 0.0);
    __syncthreads();

    // Episode start for the lanes in `mask` (DomainRandWrapperLive.reset + SimPyEnv.reset).  WARP-UNIFORM: every lane
    // of the warp calls it, `mine` says whether this lane starts an episode.  When only a few lanes restart (the common
    // case after the synchronous first episode) the other lanes are put to work instead of idling through the
    // divergent path: lane i draws table entry i of the restarting lane (one Philox block + Box-Muller per warp
    // instead of distr.n per lane), and QBB's two inverse-kinematics solves of each restarting lane run on two lanes.
    // `D` is the lane object the start is written to (the live lane, or a temporary that is parked in a slot), `ep_l` this
    // lane's index of the episode being started, `live`: the start takes effect now (observation, counters, params_out).
    auto start_episode = [&](Lane<KIND, R> &D, uint32_t ep_l, bool mine, unsigned mask, bool live) {
        const unsigned FULL = 0xffffffffu;
        // the episode's domain parameters live only inside this call (the lane keeps the derived constants): nothing of
        // them stays in registers across the hot loop
        R dp[DP];
#pragma unroll
        for (int i = 0; i < DP; i++)
            dp[i] = (io.params && mine) ? static_cast<const R *>(io.params)[n * DP + i] : cfg.base_params[i];
        const int k = __popc(mask);
        // DomainRandWrapperLive.reset (environment_wrappers/domain_randomization.py:141-148)
        if (distr.n > 0 && k <= SRE_COOP_MAX_LANES) {
            for (unsigned m = mask; m; m &= m - 1) {
                const int src = __ffs(m) - 1;
                const uint32_t g_lo = __shfl_sync(FULL, (uint32_t)genv, src);
                const uint32_t g_hi = __shfl_sync(FULL, (uint32_t)(genv >> 32), src);
                const uint32_t ep = __shfl_sync(FULL, ep_l, src);
                float v = 0.f;
                if (lane < distr.n)
                    v = sample_entry(distr.e[lane], io.seed, ((uint64_t)g_hi << 32) | g_lo, ep, (uint32_t)lane);
#pragma unroll
                for (int j = 0; j < DP; j++) {
                    const int e = distr.entry_of_param[j];  // warp-uniform
                    if (e >= 0) {
                        const float vj = __shfl_sync(FULL, v, e);
                        if (lane == src) dp[j] = (R)vj;
                    }
                }
            }
        } else if (distr.n > 0 && mine) {
            for (int i = 0; i < distr.n; i++) {
                const float v = sample_entry(distr.e[i], io.seed, genv, ep_l, (uint32_t)i);
                const int pi = distr.e[i].param_index;
#pragma unroll
                for (int j = 0; j < DP; j++)
                    if (j == pi) dp[j] = (R)v;  // static register indexing
            }
        }
        if (mine) {
            D.set_params(cfg, dp);
            if (WRAP && wrapped && live) {  // StateAugmentationWrapper: this episode's values of the selected parameters
                for (int j = 0; j < cfg.wrap.n_aug; j++) {
                    R v = (R)0;
#pragma unroll
                    for (int i = 0; i < DP; i++)
                        if (i == cfg.wrap.aug_param[j]) v = dp[i];
                    ws.aug[j] = v;
                }
            }
            if (live && first_episode && io.params_out) {
#pragma unroll
                for (int i = 0; i < DP; i++) static_cast<R *>(io.params_out)[n * DP + i] = dp[i];
            }
            R init[DS];
            int n_cols = E::DINIT;
            if (io.init) {
                n_cols = io.n_init_cols;
#pragma unroll
                for (int i = 0; i < DS; i++)
                    if (i < n_cols) init[i] = static_cast<const R *>(io.init)[n * n_cols + i];
            } else {
                sample_init<KIND, R>(cfg, dp, io.seed, genv, ep_l, init);
            }
            D.reset_state(init, n_cols);
        }
        if (KIND == SRE_QBB && !(cfg.flags & SRE_F_SIMPLE_DYN)) {
            // quanser_ball_balancer.py:231-245: plate angles from the inverse kinematics, two 100-iteration solves per
            // episode start.  Worker lane w serves the (w / 2)-th restarting lane of the chunk, axis w & 1.
            const R in0 = mine ? dp[4 % DP] : (R)0, in1 = mine ? dp[3 % DP] : (R)0;
            const R in2 = mine ? D.s[0] + dp[18 % DP] : (R)0, in3 = mine ? D.s[1 % DS] + dp[19 % DP] : (R)0;
            const int my_rank = __popc(mask & ((1u << lane) - 1u));
            for (int base = 0; base < k; base += 16) {
                const int serve = base + (lane >> 1);
                const unsigned owner = serve < k ? __fns(mask, 0, serve + 1) : 0u;
                const R ar = shfl_r(in0, (int)owner), pl = shfl_r(in1, (int)owner);
                const R t0 = shfl_r(in2, (int)owner), t1 = shfl_r(in3, (int)owner);
                float ang = 0.f;
                if (serve < k) ang = qbb_plate_ang<KIND, R>(ar, pl, (lane & 1) ? t1 : t0);
                const int w0 = 2 * (my_rank - base);
                const float a0 = __shfl_sync(FULL, ang, w0 & 31), a1 = __shfl_sync(FULL, ang, (w0 + 1) & 31);
                if (mine && my_rank >= base && my_rank < base + 16) {
                    D.h[0] = (R)a0;
                    D.h[1 % (NH > 0 ? NH : 1)] = (R)a1;
                }
            }
        } else if (mine) {
            E::reset_hidden(dp, cfg.flags, D.s, D.h);
        }
        if (mine && live) {
            D.observe(obs);
            t_ep = 0;
            ret_ep = (R)0;
            failed_ep = false;
        }
    };
    // Prepared episode starts (auto-reset sampling).  What an episode start computes depends only on (seed, env, episode),
    // so it can be computed ahead of time and parked in shared memory; a lane that restarts then only loads its slot (a
    // few dozen instructions at one active lane instead of the whole DR draw + derive + init + kinematics).  The slots
    // are filled in CONVERGENT passes whose cost does not depend on how many lanes take part (the draws run lane-parallel,
    // QBB's kinematics solves of up to 16 lanes share one 100-iteration loop): one pass right after the first start, and
    // from then on ON DEMAND -- when a lane restarts and finds its slot empty, the pass that computes its start also
    // fills the empty slot of every other running lane of the warp with that lane's NEXT start.  Every expensive pass
    // therefore serves as many episode starts as the warp has open, whatever the episode lengths are (a pass every
    // PREP_PERIOD rows, the first design, left lanes that restart twice within a period on a divergent direct path:
    // 1.7 k warp-instructions on most rows once episodes are shorter than the period).  Same values, same order of
    // episodes.  `ready` is warp-uniform: bit l = slot of lane l holds the start of episode `episode + 1` of that lane.
    const bool prep = io.prep_slots && io.auto_reset;
    unsigned ready = 0u;
    // need: lanes whose slot is filled by this pass; now: the subset that restarts in this row (their `episode` is
    // already the one to start), the others get their next one
    auto prepare_slots = [&](unsigned need, unsigned now) {
        Lane<KIND, R> P;
        const bool mine = (need >> lane) & 1u;
        start_episode(P, episode + (((now >> lane) & 1u) ? 0u : 1u), mine, need, false);
        if (mine) P.store_slot(slot, 32);
        ready |= need;
        __syncwarp();
    };
    // Episode end.  Per lane: the first episode's return / length.  Statistics: WARP-UNIFORM -- the closing lanes' values
    // are gathered with shuffles, every lane forms the same partial sums and lane 0 issues ONE set of shared-memory
    // atomics per warp and row (fp64 shared atomics are CAS loops: per-lane updates serialise when lanes close together)
    auto finish_episodes = [&](bool mine, bool complete, unsigned mask) {
        if (mine && first_episode) {
            if (io.ret) static_cast<R *>(io.ret)[n] = ret_ep;
            if (io.length) io.length[n] = t_ep;
            first_episode = false;
        }
        if (!io.stats) return;
        const unsigned cmask = __ballot_sync(0xffffffffu, mine && complete);
        const unsigned fmask = __ballot_sync(0xffffffffu, mine && complete && failed_ep);
        const int steps = __reduce_add_sync(0xffffffffu, mine ? t_ep : 0);
        const int len = __reduce_add_sync(0xffffffffu, (mine && complete) ? t_ep : 0);
        double sum = 0.0, sum2 = 0.0, mn = 1e300, mx = -1e300;
        for (unsigned m = cmask; m; m &= m - 1) {
            const double r = (double)shfl_r(ret_ep, __ffs(m) - 1);
            sum += r;
            sum2 += r * r;
            mn = fmin(mn, r);
            mx = fmax(mx, r);
        }
        if (lane == 0) {
            atomicAdd(&sm_stats[SRE_STAT_STEPS], (double)steps);
            if (cmask) {
                atomicAdd(&sm_stats[SRE_STAT_EPISODES], (double)__popc(cmask));
                atomicAdd(&sm_stats[SRE_STAT_SUM_RET], sum);
                atomicAdd(&sm_stats[SRE_STAT_SUM_RET2], sum2);
                atomicAdd(&sm_stats[SRE_STAT_SUM_LEN], (double)len);
                if (fmask) atomicAdd(&sm_stats[SRE_STAT_FAILED], (double)__popc(fmask));
                atomic_min_double(&sm_stats[SRE_STAT_MIN_RET], mn);
                atomic_max_double(&sm_stats[SRE_STAT_MAX_RET], mx);
            }
        }
    };

    start_episode(L, episode, valid, __ballot_sync(0xffffffffu, valid), true);
    if (prep) prepare_slots(__ballot_sync(0xffffffffu, valid), 0u);

    for (long long row = 0; row < T; row++) {
        // the record of this row goes straight into the warp's staging tile (the pre-step fields before the step: nothing
        // of a record stays in registers across the dynamics)
        R *const stage = stage0 + ((n_bufs == 2 && (row & 1)) ? 32 * W : 0);
        R *my_rec = stage + lane * W;
        if (n_bufs == 2 && record) {  // the bulk store of two rows ago has to be done reading this buffer
            if (lane == 0) bulk_wait_read<1>();
            __syncwarp();
        }
        unsigned char flag = SRE_ROW_IDLE;

        bool restart = false, closed = false, closed_complete = false;
        const bool stepping = (lane_state == RUN && row < T - 1);
        const bool closing = !stepping && (lane_state == RUN || lane_state == CLOSING);
        if (record && !stepping) {  // idle rows are all zero, closing rows carry obs / state / hidden only
#pragma unroll
            for (int i = 0; i < W; i++) my_rec[i] = (R)0;
        }
        if (!PLAIN && (stepping || closing) && cfg.obs_noise) add_obs_noise<DO, R>(cfg, io.seed, genv, episode, (uint32_t)t_ep, obs);
        if (!PLAIN && (stepping || closing) && cfg.obs_norm) apply_obs_norm<DO, R>(cfg, obs);
        if (WRAP && wrapped && (stepping || closing))  // exactly one new observation per stepping / closing row
            n_in = wrap_observe<R, DO, DA, DS>(cfg.wrap, ws, L.s, obs, t_ep == 0, wobs);
        // ---- policy (float32).  The tiled FNN is warp-cooperative, so it runs outside the per-lane branches ----
        R act[DA];
        if (PV == SRE_PV_ACTIONS) {
#pragma unroll
            for (int i = 0; i < DA; i++)
                act[i] = (stepping && io.actions && !io.idle_policy)
                             ? static_cast<const R *>(io.actions)[(row * N + n) * DA + i]
                             : (R)0;
        } else {
            float of[NIN], af[DA];
#pragma unroll
            for (int i = 0; i < NIN; i++) of[i] = (float)((WRAP && wrapped) ? wobs[i] : (i < DO ? obs[i] : (R)0));
            if (TC) {
                // CTA-cooperative (tcgen05), no CTA barrier inside.  Every 16 rows the CTA agrees on whether any of
                // its lanes is still running (the exit has to be CTA-uniform: all four warps feed each MMA)
                if (!record && !want_flag && !io.auto_reset && (row & 15) == 0) {
                    if (!__syncthreads_or(lane_state != IDLE)) break;
                }
                Tc::forward(tc, of, af, pol.cw);
            } else if (PV >= 16) {
                if constexpr (Fast::TAKES_CW) Fast::forward(sm_w, h1s, of, af, lane, pol.cw);
                else Fast::forward(sm_w, h1s, of, af, lane);
            }
            else {
                // parameter exploration: this lane's own parameter set, straight from global memory (L2-resident)
                const float *wsrc = per_lane_w ? pol.weights + (valid ? n / pol.group_size : 0) * pol.n_weights : sm_w;
                if (PV == SRE_PV_GENERIC) fnn_generic_forward(pol, wsrc, of, (WRAP && wrapped) ? n_in : DO, af, DA);
                else linear_forward<DO, DA>(pol, wsrc, of, af);
            }
            if (!PLAIN && pol.act_noise && stepping)
                add_act_noise<DA>(pol.act_noise, pol.noise_scale, pol.noise_mean, io.seed, genv, episode, (uint32_t)t_ep, af);
#pragma unroll
            for (int i = 0; i < DA; i++) act[i] = (R)af[i];
        }
        if (stepping) {
            // ---- wrappers: ActNormWrapper._process_act (action_normalization.py:69-75) ----
            R act_env[DA];
#pragma unroll
            for (int i = 0; i < DA; i++) act_env[i] = act[i];
            if (WRAP && wrapped) {
                // DownsamplingWrapper._process_act (downsampling.py:143-151) and the outer ActDelayWrapper, in chain order
                const KWrap<R> &kw = cfg.wrap;
                if (kw.delay > 0 && !kw.delay_inner && kw.delay_first) wrap_delay(kw, ws, act_env);
                if (kw.down > 1) {
                    if (ws.cnt % kw.down == 0) {
#pragma unroll
                        for (int i = 0; i < DA; i++) ws.hold[i] = act_env[i];
                    } else {
#pragma unroll
                        for (int i = 0; i < DA; i++) act_env[i] = ws.hold[i];
                    }
                    ws.cnt++;
                }
                if (kw.delay > 0 && !kw.delay_inner && !kw.delay_first) wrap_delay(kw, ws, act_env);
            }
            // GaussianActNoiseWrapper outside / inside ActNormWrapper (the recorded action stays the policy's)
            if (!PLAIN && cfg.env_act_noise == SRE_ENV_ACT_NOISE_OUTER)
                add_env_act_noise<DA, R>(cfg, io.seed, genv, episode, (uint32_t)t_ep, act_env);
            R a_lo[DA], a_up[DA];
            L.act_bounds(cfg.flags, a_lo, a_up);
#pragma unroll
            for (int i = 0; i < DA; i++)
                if (cfg.act_norm) act_env[i] = a_lo[i] + (act_env[i] + (R)1) * (a_up[i] - a_lo[i]) / (R)2;
            if (!PLAIN && cfg.env_act_noise == SRE_ENV_ACT_NOISE_INNER)
                add_env_act_noise<DA, R>(cfg, io.seed, genv, episode, (uint32_t)t_ep, act_env);
            if (WRAP && wrapped && cfg.wrap.delay > 0 && cfg.wrap.delay_inner) wrap_delay(cfg.wrap, ws, act_env);
#pragma unroll
            for (int i = 0; i < DA; i++) {
                if (m_isnan(act[i])) status |= SRE_ST_NAN_ACT;
                // rollout.py:244 env.limit_act(act): through ActNormWrapper the NORMALISED action is clipped to [-1, 1]
                const R lo = cfg.act_norm ? (R)-1 : a_lo[i], up = cfg.act_norm ? (R)1 : a_up[i];
                if (record) {
                    my_rec[RC::OFF_ACT + i] = act[i];
                    my_rec[RC::OFF_ACT_APP + i] = m_min(m_max(act[i], lo), up);
                }
            }
            if (KIND == SRE_BOB && (cfg.flags & SRE_F_DISC_ACT) && !cfg.act_norm && record) {
                R app[DA];
                L.limit_act(cfg, act, app);
#pragma unroll
                for (int i = 0; i < DA; i++) my_rec[RC::OFF_ACT_APP + i] = app[i];
            }
#pragma unroll
            for (int i = 0; i < DO; i++) {
                if (record) my_rec[RC::OFF_OBS + i] = (WRAP && wrapped) ? (i < n_in ? wobs[i] : (R)0) : obs[i];
                if (m_isnan(obs[i])) status |= SRE_ST_NAN_OBS;
            }
            if (record) {
#pragma unroll
                for (int i = 0; i < DS; i++) my_rec[RC::OFF_STATE + i] = L.s[i];
#pragma unroll
                for (int i = 0; i < NH; i++) my_rec[RC::OFF_HIDDEN + i] = L.h[i];
            }
            // ---- SimPyEnv.step ----
            R rew;
            int inside;
            const bool done = L.template step<PLAIN>(cfg, act_env, rew, inside);
            if (inside < 0) status |= SRE_ST_NAN_STATE;
            L.observe(obs);
            if (record) my_rec[RC::OFF_REW] = rew;
            ret_ep += rew;
            t_ep += 1;
            flag = done ? SRE_ROW_DONE : SRE_ROW_STEP;
            if (done && io.stop_on_done) {
                lane_state = CLOSING;
                failed_ep = (inside != 1);
            }
        } else if (closing) {
            // closing row (rollout.py:309-311): final observation / state, nothing else
            if (record) {
#pragma unroll
                for (int i = 0; i < DO; i++) my_rec[RC::OFF_OBS + i] = (WRAP && wrapped) ? (i < n_in ? wobs[i] : (R)0) : obs[i];
#pragma unroll
                for (int i = 0; i < DS; i++) my_rec[RC::OFF_STATE + i] = L.s[i];
#pragma unroll
                for (int i = 0; i < NH; i++) my_rec[RC::OFF_HIDDEN + i] = L.h[i];
            }
            const bool complete = (lane_state == CLOSING);
            flag = complete ? SRE_ROW_TERMINAL : SRE_ROW_TRUNCATED;
            closed = true;
            closed_complete = complete;
        }
        {
            // warp-uniform tail of the closing rows: statistics, then the next episodes (the closing row's record is
            // already in `rec`)
            const unsigned cm = __ballot_sync(0xffffffffu, closed);
            if (cm) {
                finish_episodes(closed, closed_complete, cm);
                if (closed) {
                    if (io.auto_reset && row < T - 1) {
                        episode += 1;
                        restart = true;
                        lane_state = RUN;
                    } else {
                        lane_state = IDLE;
                    }
                }
                const unsigned rs = __ballot_sync(0xffffffffu, restart);
                if (prep && (rs & ~ready) && row < T - 2) {
                    // a restarting lane has no prepared start: one pass for it and for every other empty slot
                    const unsigned running = __ballot_sync(0xffffffffu, lane_state != IDLE);
                    prepare_slots(running & ~ready, rs & ~ready);
                }
                const unsigned fast = rs & ready;
                if (fast) {
                    if ((fast >> lane) & 1u) {
                        L.load_slot(slot, 32);
                        L.observe(obs);
                        t_ep = 0;
                        ret_ep = (R)0;
                        failed_ep = false;
                    }
                    ready &= ~fast;
                }
                if (rs & ~fast) start_episode(L, episode, restart && !((fast >> lane) & 1u), rs & ~fast, true);
            }
        }

        if (record) {
            R *const gdst = static_cast<R *>(io.traj) + (row * NS + NO + warp_env0) * W;
            const unsigned tile_bytes = (unsigned)(n_valid_lanes * W * sizeof(R));
            if (n_bufs == 2 && (tile_bytes & 15u) == 0u) {
                fence_async_smem();
                __syncwarp();
                if (lane == 0 && tile_bytes) bulk_store_tile(gdst, stage, tile_bytes);
            } else {
                __syncwarp();
                if (n_valid_lanes > 0) warp_store_row<R>(stage, gdst, n_valid_lanes * W, vec_ok, lane);
                __syncwarp();
            }
        }
        if (want_flag && valid) io.row_flag[row * NS + NO + n] = flag;
        if (!TC && !record && !want_flag) {
            // early exit accounting: the warp leaves when every lane is idle
            if (__ballot_sync(0xffffffffu, lane_state != IDLE) == 0u) break;
        }
    }
    if (n_bufs == 2 && lane == 0) bulk_wait_read<0>();  // shared memory must outlive the last bulk stores
#ifndef SRE_TC_TRACE
    if (valid && io.status) io.status[n] = status;
#endif
    if (TC) Tc::teardown(tc);

    // ---- statistics: one atomic set per CTA ----
    if (io.stats) {
        __syncthreads();
        if (threadIdx.x < 6) {
            if (sm_stats[threadIdx.x] != 0.0) atomicAdd(io.stats + threadIdx.x, sm_stats[threadIdx.x]);
        } else if (threadIdx.x == SRE_STAT_MIN_RET) {
            if (sm_stats[SRE_STAT_EPISODES] > 0) atomic_min_double(io.stats + SRE_STAT_MIN_RET, sm_stats[SRE_STAT_MIN_RET]);
        } else if (threadIdx.x == SRE_STAT_MAX_RET) {
            if (sm_stats[SRE_STAT_EPISODES] > 0) atomic_max_double(io.stats + SRE_STAT_MAX_RET, sm_stats[SRE_STAT_MAX_RET]);
        }
    }
}

// host-side: pick the policy variant, size shared memory, launch
template <int KIND, typename R>
int launch_rollout(const KCfg<R> &cfg, const KDistr &distr, KPolicy pol, const KIO &io, cudaStream_t stream);

template <int KIND, typename R, int PV, int MODE>
static int launch_rollout_mode(const KCfg<R> &cfg, const KDistr &distr, const KPolicy &pol, const KIO &io_in,
                               cudaStream_t stream) {
    constexpr int W = Rec<KIND, R>::W;
    const int threads = LaunchShape<R, PV>::THREADS;
    const long long blocks = (io_in.n_envs + threads - 1) / threads;
    KIO io2 = io_in;
    const size_t tile_bytes = (size_t)(threads / 32) * 32 * W * sizeof(R);
    // record tiles leave with bulk stores from a double-buffered stage when the rows are 16-byte aligned in `traj`
    // (SRE_NO_BULK=1: the LDS / STG loop, for A/B measurements).  Measured: C4 (QBB, 256-thread CTAs) 1.78e10 -> 2.01e10
    // env-steps/s; the tensor-core variants (128-thread CTAs, four per SM) lose 3 % to the fence + wait and keep the loop.
    const char *nobulk = getenv("SRE_NO_BULK");
    io2.stage_bufs = (PV < SRE_PV_TC && io_in.traj && ((size_t)io_in.traj_stride * W * sizeof(R)) % 16 == 0 &&
                      ((size_t)io_in.traj_off * W * sizeof(R)) % 16 == 0 && (32 * W * sizeof(R)) % 16 == 0 &&
                      reinterpret_cast<size_t>(io_in.traj) % 16 == 0 && !(nobulk && nobulk[0] == '1'))
                         ? 2 : 1;
    size_t smem = (size_t)pol.smem_floats * sizeof(float) + io2.stage_bufs * tile_bytes;
    if (PV >= SRE_PV_TC && !io_in.traj) smem = 0;  // the staging area is only needed when records are written
    if (PV >= SRE_PV_TC) smem += FnnTC2<Env<KIND, R>::DO, Env<KIND, R>::DA>::SMEM_BYTES;
    // prepared episode starts: only where they pay (auto-reset sampling with something to compute per episode) and
    // where the resident CTAs of an SM still fit its shared memory
    const size_t slot_bytes = (size_t)threads * Lane<KIND, R>::NSLOT * sizeof(R);
    const char *noprep = getenv("SRE_NO_PREP");
    io2.prep_slots = io_in.auto_reset && !(noprep && noprep[0] == '1') && !cfg.wrap.active &&
                     (smem + slot_bytes + 1024) * LaunchShape<R, PV>::MIN_BLOCKS <= 220 * 1024;
    if (io2.prep_slots) smem += slot_bytes;
    const KIO &io = io2;
    auto kern = rollout_kernel<KIND, R, PV, MODE>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    kern<<<(unsigned)blocks, threads, smem, stream>>>(cfg, distr, pol, io);
    return (int)cudaGetLastError();
}

// The plain bodies exist for the production (float) build of the fused-FNN variants -- the sampling hot path; every other
// combination runs the generic body.
template <int KIND, typename R, int PV>
static int launch_rollout_pv(const KCfg<R> &cfg, const KDistr &distr, const KPolicy &pol, const KIO &io,
                             cudaStream_t stream) {
    constexpr bool HAS_PLAIN = sizeof(R) == 4 && PV >= 16;
    const char *force = getenv("SRE_FORCE_GENERIC");  // tests: the plain bodies must reproduce the generic one bit for bit
    if (force && force[0] == '1') return launch_rollout_mode<KIND, R, PV, SRE_MODE_GENERIC>(cfg, distr, pol, io, stream);
    if constexpr (HAS_PLAIN) {
        const bool plain = !cfg.obs_noise && !cfg.obs_norm && !cfg.env_act_noise && !pol.act_noise && cfg.task.q_diag &&
                           !(cfg.task.final_flags & SRE_FR_ENABLED);
        if (plain && !io.traj && !io.row_flag)
            return launch_rollout_mode<KIND, R, PV, SRE_MODE_PLAIN>(cfg, distr, pol, io, stream);
        if (plain && io.traj && io.row_flag)
            return launch_rollout_mode<KIND, R, PV, SRE_MODE_PLAIN_REC>(cfg, distr, pol, io, stream);
    }
    return launch_rollout_mode<KIND, R, PV, SRE_MODE_GENERIC>(cfg, distr, pol, io, stream);
}

template <int KIND, typename R>
int launch_rollout(const KCfg<R> &cfg, const KDistr &distr, KPolicy pol, const KIO &io, cudaStream_t stream) {
    using E = Env<KIND, R>;
    if (pol.kind == SRE_POL_ACTIONS || pol.kind == SRE_POL_IDLE) {
        pol.smem_floats = 0;
        return launch_rollout_pv<KIND, R, SRE_PV_ACTIONS>(cfg, distr, pol, io, stream);
    }
    // parameter exploration: groups that are multiples of the variant's CTA size get one staged set per CTA and run
    // every variant below unchanged; other group sizes read their weights per lane on the un-staged variants
    auto cta_aligned = [&](int threads) { return pol.group_size > 0 && pol.group_size % threads == 0; };
    pol.cta_groups = 0;
    if (pol.kind == SRE_POL_LINEAR) {
        pol.cta_groups = cta_aligned(LaunchShape<R, SRE_PV_LINEAR>::THREADS);
        pol.smem_floats = (pol.group_size > 0 && !pol.cta_groups) ? 4 : (int)((pol.n_weights + 3) / 4 * 4);
        return launch_rollout_pv<KIND, R, SRE_PV_LINEAR>(cfg, distr, pol, io, stream);
    }
    // FNN: the register-resident fast path needs [do, H, H, da] with tanh hidden and linear output
    // the stateful step wrappers live in the generic body of the un-tiled variants
    const bool two_hidden = !cfg.wrap.active && pol.n_layers == 3 && pol.sizes[0] == E::DO && pol.sizes[3] == E::DA &&
                            pol.sizes[1] == pol.sizes[2] && pol.hidden_nonlin == 0 && pol.output_nonlin == 0 &&
                            !pol.cartpole_feats;
    const int fast_threads = two_hidden && pol.sizes[1] == 64 && pol.precision != SRE_PREC_FP32 && sizeof(R) == 4
                                 ? LaunchShape<R, SRE_PV_TC>::THREADS
                                 : (two_hidden && (pol.sizes[1] == 64 || pol.sizes[1] == 32 || pol.sizes[1] == 16)
                                        ? LaunchShape<R, SRE_PV_FNN16>::THREADS
                                        : LaunchShape<R, SRE_PV_GENERIC>::THREADS);
    pol.cta_groups = cta_aligned(fast_threads);
    if (pol.group_size > 0 && !pol.cta_groups) {  // per-lane parameter sets: only the un-staged forward pass applies
        pol.smem_floats = 4;
        return launch_rollout_pv<KIND, R, SRE_PV_GENERIC>(cfg, distr, pol, io, stream);
    }
    if (two_hidden && pol.sizes[1] == 64 && pol.precision != SRE_PREC_FP32 && sizeof(R) == 4) {
        pol.smem_floats = 0;
        if (pol.precision == SRE_PREC_TF32) return launch_rollout_pv<KIND, R, SRE_PV_TC1>(cfg, distr, pol, io, stream);
        const char *nocw = getenv("SRE_NO_CW");  // A/B measurements and tests: ignore the host copy of the weights
        if (pol.has_cw == 1 && !pol.cta_groups && !(nocw && nocw[0] == '1'))
            return launch_rollout_pv<KIND, R, SRE_PV_TCW>(cfg, distr, pol, io, stream);
        return launch_rollout_pv<KIND, R, SRE_PV_TC>(cfg, distr, pol, io, stream);
    }
    if (two_hidden && pol.sizes[1] == 64) {
        pol.smem_floats = FnnTiled<E::DO, 64, E::DA, sizeof(R) == 4>::SMEM_FLOATS;
        return launch_rollout_pv<KIND, R, SRE_PV_FNN64>(cfg, distr, pol, io, stream);
    }
    if (two_hidden && pol.sizes[1] == 32) {
        pol.smem_floats = FnnTiled<E::DO, 32, E::DA, sizeof(R) == 4>::SMEM_FLOATS;
        return launch_rollout_pv<KIND, R, SRE_PV_FNN32>(cfg, distr, pol, io, stream);
    }
    if (two_hidden && pol.sizes[1] == 16) {
        pol.smem_floats = FnnTiled<E::DO, 16, E::DA, sizeof(R) == 4>::SMEM_FLOATS;
        const char *nocw = getenv("SRE_NO_CW");
        if (sizeof(R) == 4 && pol.has_cw == 2 && !pol.cta_groups && !(nocw && nocw[0] == '1'))
            return launch_rollout_pv<KIND, R, SRE_PV_FNN16W>(cfg, distr, pol, io, stream);
        return launch_rollout_pv<KIND, R, SRE_PV_FNN16>(cfg, distr, pol, io, stream);
    }
    pol.smem_floats = (int)((pol.n_weights + 3) / 4 * 4);
    return launch_rollout_pv<KIND, R, SRE_PV_GENERIC>(cfg, distr, pol, io, stream);
}

}  // namespace sre
