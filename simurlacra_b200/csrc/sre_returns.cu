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
// four rows are issued before the dependent recurrence consumes them.  HBM-bound: 4 (r) + 1 (flag) + 4 (v) bytes read
// and 4 + 4 written per row and lane when the inputs are contiguous [T, N] arrays.
#include <cuda_runtime.h>
#include <stdint.h>

#include "sre_host.h"

namespace sre {

template <typename R, bool HAS_V>
__global__ void __launch_bounds__(256)
returns_kernel(const R *__restrict__ rew, long long rew_row_stride, long long rew_lane_stride,
               const unsigned char *__restrict__ flag, const R *__restrict__ val, R gamma, R lamb,
               R *__restrict__ out_ret, R *__restrict__ out_adv, long long T, long long N) {
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    R G = (R)0, A = (R)0, v_next = (R)0;
    const R gl = gamma * lamb;
    constexpr int U = 4;
    long long row = T - 1;
    // rows in groups of U from the end; the remainder (T % U rows at the front) is handled by the same body with U = 1
    auto body = [&](long long t, R r, unsigned char f, R v) {
        R o_ret = (R)0, o_adv = (R)0;
        if (f == SRE_ROW_STEP || f == SRE_ROW_DONE) {
            const bool done = (f == SRE_ROW_DONE);
            G = done ? r : r + gamma * G;
            o_ret = G;
            if (HAS_V) {
                const R delta = r + (done ? (R)0 : gamma * v_next) - v;
                A = done ? delta : delta + gl * A;
                o_adv = A;
            }
        } else if (f == SRE_ROW_TRUNCATED) {
            G = HAS_V ? v : (R)0;
            A = (R)0;
        } else {  // TERMINAL, IDLE
            G = (R)0;
            A = (R)0;
        }
        v_next = v;
        out_ret[t * N + n] = o_ret;
        if (HAS_V) out_adv[t * N + n] = o_adv;
    };
    for (; row >= U - 1; row -= U) {
        R r[U], v[U];
        unsigned char f[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const long long t = row - u;
            f[u] = flag[t * N + n];
            r[u] = rew[t * rew_row_stride + n * rew_lane_stride];
            v[u] = HAS_V ? val[t * N + n] : (R)0;
        }
#pragma unroll
        for (int u = 0; u < U; u++) body(row - u, r[u], f[u], v[u]);
    }
    for (; row >= 0; row--)
        body(row, rew[row * rew_row_stride + n * rew_lane_stride], flag[row * N + n], HAS_V ? val[row * N + n] : (R)0);
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
    const unsigned blocks = (unsigned)((n_envs + 255) / 256);
#define SRE_LAUNCH_RET(R, HV)                                                                                           \
    returns_kernel<R, HV><<<blocks, 256, 0, s>>>((const R *)rew, rew_row_stride, rew_lane_stride, row_flag,            \
                                                 (const R *)values, (R)gamma, (R)lamb, (R *)out_ret, (R *)out_adv,    \
                                                 n_rows, n_envs)
    if (dtype == SRE_F64) {
        if (values) SRE_LAUNCH_RET(double, true); else SRE_LAUNCH_RET(double, false);
    } else {
        if (values) SRE_LAUNCH_RET(float, true); else SRE_LAUNCH_RET(float, false);
    }
#undef SRE_LAUNCH_RET
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) set_error("sre_returns: %s", cudaGetErrorString(e));
    return (int)e;
}
