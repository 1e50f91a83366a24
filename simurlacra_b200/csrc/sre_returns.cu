// sre_returns.cu -- post-processing of the time-major record buffer on the device: discounted returns-to-go and
// generalised advantage estimates for every episode of every lane, without copying trajectories to the host.
//
// Replaces discounted_reverse_cumsum / discounted_value(s) / gae_returns (pyrado/sampling/step_sequence.py:1049-1116)
// applied to each rollout of a sample() call.  One thread per env lane scans its rows from the last to the first; the
// row flags written by the rollout kernel delimit the episodes:
//   SRE_ROW_STEP       G = r + gamma G',          delta = r + gamma v' - v,   A = delta + gamma lambda A'
//   SRE_ROW_DONE       G = r,                     delta = r - v,              A = delta          (step.done: next value 0)
//   SRE_ROW_TERMINAL   closing row of a finished episode: G' = A' = 0 for the row before (which is a DONE row)
//   SRE_ROW_TRUNCATED  closing row of an episode cut by the row budget: the row before bootstraps from this row's value
//                      (v' = v[row], G' = v[row] if values are given, else 0) -- the reference never truncates
//   SRE_ROW_IDLE       nothing
// Rows are contiguous over lanes (time-major), so every load / store of a warp is one coalesced line; the loads of
// SRE_RET_UNROLL rows are issued before the dependent recurrence consumes them.  HBM-bound: 4 (r) + 1 (flag) + 4 (v) bytes read
// and 4 + 4 written per row and lane when the inputs are contiguous [T, N] arrays.
#include <cuda_runtime.h>
#include <stdint.h>

#include "sre_host.h"

#ifndef SRE_RET_UNROLL
#define SRE_RET_UNROLL 2
#endif
#ifndef SRE_RET_VEC
#define SRE_RET_VEC 1
#endif
#ifndef SRE_RET_THREADS
#define SRE_RET_THREADS 128
#endif

namespace sre {

// V consecutive lanes per thread: V = 4 (float, contiguous inputs, N % 4 == 0) turns every access into a 128-bit
// load / store (a 256-thread block then touches 4 KB contiguous per row and array instead of 1 KB, and carries four
// independent recurrences per thread); V = 1 is the general path (strided `rew` column of the record buffer, fp64).
template <typename R, int V> struct RowVec { R x[V]; };

template <typename R, int V>
__device__ __forceinline__ RowVec<R, V> load_row(const R *p, long long lane_stride) {
    RowVec<R, V> o;
    if constexpr (V == 4 && sizeof(R) == 4) {
        const float4 q = *reinterpret_cast<const float4 *>(p);
        o.x[0] = q.x; o.x[1] = q.y; o.x[2] = q.z; o.x[3] = q.w;
    } else {
#pragma unroll
        for (int i = 0; i < V; i++) o.x[i] = p[i * lane_stride];
    }
    return o;
}
template <typename R, int V> __device__ __forceinline__ void store_row(R *p, const RowVec<R, V> &o) {
    if constexpr (V == 4 && sizeof(R) == 4) {
        *reinterpret_cast<float4 *>(p) = make_float4(o.x[0], o.x[1], o.x[2], o.x[3]);
    } else {
#pragma unroll
        for (int i = 0; i < V; i++) p[i] = o.x[i];
    }
}

template <typename R, bool HAS_V, int V>
__global__ void __launch_bounds__(SRE_RET_THREADS)
returns_kernel(const R *__restrict__ rew, long long rew_row_stride, long long rew_lane_stride,
               const unsigned char *__restrict__ flag, const R *__restrict__ val, R gamma, R lamb,
               R *__restrict__ out_ret, R *__restrict__ out_adv, long long T, long long N) {
    const long long n = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * V;
    if (n >= N) return;
    R G[V], A[V], v_next[V];
#pragma unroll
    for (int i = 0; i < V; i++) G[i] = A[i] = v_next[i] = (R)0;
    const R gl = gamma * lamb;
    constexpr int U = SRE_RET_UNROLL;
    struct Row {
        RowVec<R, V> r, v;
        unsigned char f[V];
    };
    auto load = [&](long long t) {
        Row o;
        if constexpr (V == 4) {
            const uchar4 q = *reinterpret_cast<const uchar4 *>(flag + t * N + n);
            o.f[0] = q.x; o.f[1] = q.y; o.f[2] = q.z; o.f[3] = q.w;
        } else {
#pragma unroll
            for (int i = 0; i < V; i++) o.f[i] = flag[t * N + n + i];
        }
        o.r = load_row<R, V>(rew + t * rew_row_stride + n * rew_lane_stride, rew_lane_stride);
        if (HAS_V) o.v = load_row<R, V>(val + t * N + n, 1);
        return o;
    };
    auto body = [&](long long t, const Row &in) {
        RowVec<R, V> o_ret, o_adv;
#pragma unroll
        for (int i = 0; i < V; i++) {
            const unsigned char f = in.f[i];
            const R r = in.r.x[i], v = HAS_V ? in.v.x[i] : (R)0;
            o_ret.x[i] = (R)0;
            o_adv.x[i] = (R)0;
            if (f == SRE_ROW_STEP || f == SRE_ROW_DONE) {
                const bool done = (f == SRE_ROW_DONE);
                G[i] = done ? r : r + gamma * G[i];
                o_ret.x[i] = G[i];
                if (HAS_V) {
                    const R delta = r + (done ? (R)0 : gamma * v_next[i]) - v;
                    A[i] = done ? delta : delta + gl * A[i];
                    o_adv.x[i] = A[i];
                }
            } else if (f == SRE_ROW_TRUNCATED) {
                G[i] = HAS_V ? v : (R)0;
                A[i] = (R)0;
            } else {  // TERMINAL, IDLE
                G[i] = (R)0;
                A[i] = (R)0;
            }
            v_next[i] = v;
        }
        store_row<R, V>(out_ret + t * N + n, o_ret);
        if (HAS_V) store_row<R, V>(out_adv + t * N + n, o_adv);
    };
    long long row = T - 1;
    // rows in groups of U from the end (all loads of a group are issued before the recurrence consumes them); the
    // T % U rows left at the front go one at a time
    for (; row >= U - 1; row -= U) {
        Row in[U];
#pragma unroll
        for (int u = 0; u < U; u++) in[u] = load(row - u);
#pragma unroll
        for (int u = 0; u < U; u++) body(row - u, in[u]);
    }
    for (; row >= 0; row--) body(row, load(row));
}

}  // namespace sre

using namespace sre;

extern "C" int sre_returns(int dtype, const void *rew, int64_t rew_row_stride, int64_t rew_lane_stride,
                           const uint8_t *row_flag, const void *values, double gamma, double lamb, void *out_ret,
                           void *out_adv, int64_t n_rows, int64_t n_envs, void *stream) {
    SRE_CHECK_ARG(dtype == SRE_F32 || dtype == SRE_F64, SRE_E_ARG, "unknown dtype %d", dtype);
    SRE_CHECK_ARG(n_rows >= 0 && n_envs >= 0, SRE_E_ARG, "n_rows %lld / n_envs %lld invalid", (long long)n_rows,
                  (long long)n_envs);
    SRE_CHECK_ARG(gamma >= 0.0 && gamma <= 1.0 && lamb >= 0.0 && lamb <= 1.0, SRE_E_ARG, "gamma %g / lamb %g outside [0, 1]",
                  gamma, lamb);
    if (n_rows == 0 || n_envs == 0) return 0;  // empty buffers carry NULL pointers
    SRE_CHECK_ARG(rew && row_flag && out_ret, SRE_E_ARG, "rew, row_flag and out_ret must not be NULL");
    SRE_CHECK_ARG((values == nullptr) == (out_adv == nullptr), SRE_E_ARG,
                  "values and out_adv go together (advantages need the value estimates)");
    cudaStream_t s = (cudaStream_t)stream;
    // vector path: float, contiguous [T, N] inputs, four lanes per thread
    const bool vec = SRE_RET_VEC && dtype == SRE_F32 && n_envs % 4 == 0 && rew_lane_stride == 1 && rew_row_stride % 4 == 0 &&
                     ((uintptr_t)rew % 16 == 0) && ((uintptr_t)row_flag % 4 == 0) && ((uintptr_t)out_ret % 16 == 0) &&
                     (!values || ((uintptr_t)values % 16 == 0 && (uintptr_t)out_adv % 16 == 0));
    const long long threads_needed = vec ? n_envs / 4 : n_envs;
    const unsigned blocks = (unsigned)((threads_needed + SRE_RET_THREADS - 1) / SRE_RET_THREADS);
#define SRE_LAUNCH_RET(R, HV, VV)                                                                                      \
    returns_kernel<R, HV, VV><<<blocks, SRE_RET_THREADS, 0, s>>>((const R *)rew, rew_row_stride, rew_lane_stride,     \
                                                                 row_flag, (const R *)values, (R)gamma, (R)lamb,      \
                                                                 (R *)out_ret, (R *)out_adv, n_rows, n_envs)
    if (dtype == SRE_F64) {
        if (values) SRE_LAUNCH_RET(double, true, 1); else SRE_LAUNCH_RET(double, false, 1);
    } else if (vec) {
        if (values) SRE_LAUNCH_RET(float, true, 4); else SRE_LAUNCH_RET(float, false, 4);
    } else {
        if (values) SRE_LAUNCH_RET(float, true, 1); else SRE_LAUNCH_RET(float, false, 1);
    }
#undef SRE_LAUNCH_RET
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) set_error("sre_returns: %s", cudaGetErrorString(e));
    return (int)e;
}
