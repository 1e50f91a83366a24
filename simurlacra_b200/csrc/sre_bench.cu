// sre_bench.cu -- measurement helpers: the roofline denominators are MEASURED on the box, because
// MEASURED_PEAKS.json has no FP32/FP64 FMA figure.  bench.py times these launches with CUDA events.
#include "sre_host.h"

namespace sre {

// 32 independent accumulator chains per thread: no dependency stalls, no memory traffic
template <typename R> __global__ void __launch_bounds__(256) fma_kernel(long long iters, R seed, R *sink) {
    R acc[32];
#pragma unroll
    for (int i = 0; i < 32; i++) acc[i] = seed + (R)(threadIdx.x + i);
    const R a = (R)1.0000001, b = (R)1e-7;
    for (long long it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 32; i++) acc[i] = acc[i] * a + b;
    }
    R s = (R)0;
#pragma unroll
    for (int i = 0; i < 32; i++) s += acc[i];
    if (s == (R)123.456) sink[0] = s;  // keeps the loop alive, never true in practice
}

// packed FP32 (Blackwell fma.rn.f32x2): 16 independent 64-bit accumulator pairs = the same 32 FMA chains per thread
__global__ void __launch_bounds__(256) fma2_kernel(long long iters, float seed, float *sink) {
    unsigned long long acc[16];
#pragma unroll
    for (int i = 0; i < 16; i++) {
        float lo = seed + (float)(threadIdx.x + 2 * i), hi = lo + 1.f;
        asm("mov.b64 %0, {%1, %2};" : "=l"(acc[i]) : "f"(lo), "f"(hi));
    }
    unsigned long long a, b;
    asm("mov.b64 %0, {%1, %1};" : "=l"(a) : "f"(1.0000001f));
    asm("mov.b64 %0, {%1, %1};" : "=l"(b) : "f"(1e-7f));
    for (long long it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[i]) : "l"(a), "l"(b));
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i]));
        s += lo + hi;
    }
    if (s == 123.456f) sink[0] = s;
}

// FMA with three DISTINCT, changing register operands per instruction (acc[i] = acc[i+1] * acc[i+2] + acc[i]): what
// simulation code issues -- no operand comes from an immediate, the constant bank or the operand-reuse cache, so the
// register file has to deliver three reads per lane and instruction.  PACKED: the same on fma.rn.f32x2.
template <bool PACKED> __global__ void __launch_bounds__(256) fma3reg_kernel(long long iters, float seed, float *sink) {
    if constexpr (!PACKED) {
        float acc[32];
#pragma unroll
        for (int i = 0; i < 32; i++) acc[i] = 1.0f + 1e-7f * (seed + (float)(threadIdx.x + i));
        for (long long it = 0; it < iters; it++) {
#pragma unroll
            for (int i = 0; i < 32; i++)
                asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc[i]) : "f"(acc[(i + 11) & 31]), "f"(acc[(i + 22) & 31]));
        }
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 32; i++) s += acc[i];
        if (s == 123.456f) sink[0] = s;
    } else {
        unsigned long long acc[16];
#pragma unroll
        for (int i = 0; i < 16; i++) {
            float lo = 1.0f + 1e-7f * (seed + (float)(threadIdx.x + 2 * i)), hi = lo + 1e-7f;
            asm("mov.b64 %0, {%1, %2};" : "=l"(acc[i]) : "f"(lo), "f"(hi));
        }
        for (long long it = 0; it < iters; it++) {
#pragma unroll
            for (int i = 0; i < 16; i++)
                asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc[i]) : "l"(acc[(i + 5) & 15]), "l"(acc[(i + 11) & 15]));
        }
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 16; i++) {
            float lo, hi;
            asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i]));
            s += lo + hi;
        }
        if (s == 123.456f) sink[0] = s;
    }
}

// Do the FMA pipe and the XU (MUFU) pipe overlap?  Per iteration 16 FFMA2 with three register operands (MODE & 1) and / or
// 8 MUFU.EX2 on independent chains (MODE & 2): if time(both) ~ max(time(fma), time(mufu)) the pipes run concurrently.
template <int MODE> __global__ void __launch_bounds__(256) pipe_mix_kernel(long long iters, float seed, float *sink) {
    unsigned long long acc[16];
    float m[8];
#pragma unroll
    for (int i = 0; i < 16; i++) {
        float lo = 1.0f + 1e-7f * (seed + (float)(threadIdx.x + 2 * i)), hi = lo + 1e-7f;
        asm("mov.b64 %0, {%1, %2};" : "=l"(acc[i]) : "f"(lo), "f"(hi));
    }
#pragma unroll
    for (int i = 0; i < 8; i++) m[i] = -1.0f - seed * (float)i;
    for (long long it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) {
            if (MODE & 1)
                asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc[i]) : "l"(acc[(i + 5) & 15]), "l"(acc[(i + 11) & 15]));
            if ((MODE & 2) && (i & 1)) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(m[i >> 1]));
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i]));
        s += lo + hi;
    }
#pragma unroll
    for (int i = 0; i < 8; i++) s += m[i];
    if (s == 123.456f) sink[0] = s;
}

// Cost of packed FP32 instructions by operand form (cycles per warp instruction on one scheduler):
//   FORM 0  fma2 d, a, b, d   a, b, d changing registers            (fma3reg_kernel<true>)
//   FORM 1  fma2 d, a, X, d   X a loop-invariant per-thread register pair
//   FORM 2  fma2 d, a, U, d   U launch-uniform (kernel parameter -> uniform register / constant bank)
//   FORM 3  mul2 d, a, b      two changing registers
//   FORM 4  add2 d, a, 1.0    register + immediate
struct MixParams {
    float u[8];
};
template <int FORM>
__global__ void __launch_bounds__(256) pipe_form_kernel(const __grid_constant__ MixParams prm, long long iters, float seed, float *sink) {
    unsigned long long acc[16];
#pragma unroll
    for (int i = 0; i < 16; i++) {
        float lo = 1.0f + 1e-7f * (seed + (float)(threadIdx.x + 2 * i)), hi = lo + 1e-7f;
        asm("mov.b64 %0, {%1, %2};" : "=l"(acc[i]) : "f"(lo), "f"(hi));
    }
    unsigned long long X, U[4], one;
    asm("mov.b64 %0, {%1, %2};" : "=l"(X) : "f"(1.0f + 1e-7f * (float)threadIdx.x), "f"(1.0f - 1e-7f * (float)threadIdx.x));
    asm("mov.b64 %0, {%1, %1};" : "=l"(one) : "f"(1.0f));
#pragma unroll
    for (int i = 0; i < 4; i++) asm("mov.b64 %0, {%1, %2};" : "=l"(U[i]) : "f"(prm.u[2 * i]), "f"(prm.u[2 * i + 1]));
    for (long long it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) {
            if (FORM == 1) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc[i]) : "l"(acc[(i + 5) & 15]), "l"(X));
            if (FORM == 2) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc[i]) : "l"(acc[(i + 5) & 15]), "l"(U[i & 3]));
            if (FORM == 3) asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(acc[i]) : "l"(acc[(i + 5) & 15]), "l"(acc[(i + 11) & 15]));
            if (FORM == 4) asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(acc[i]) : "l"(acc[(i + 5) & 15]), "l"(one));
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; i++) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i]));
        s += lo + hi;
    }
    if (s == 123.456f) sink[0] = s;
}

__global__ void __launch_bounds__(256) hbm_write_kernel(float4 *dst, long long n_vec) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride) __stcs(dst + i, v);
}

}  // namespace sre

using namespace sre;

extern "C" {

int sre_bench_fma(int dtype, int64_t iters, int blocks, int threads, float *sink, double *flops, void *stream) {
    SRE_CHECK_ARG(dtype == SRE_F32 || dtype == SRE_F64 || (dtype >= 2 && dtype <= 11), SRE_E_DTYPE, "unknown dtype %d", dtype);
    SRE_CHECK_ARG(iters > 0 && blocks > 0 && threads > 0 && threads <= 256 && sink, SRE_E_ARG, "bad launch shape");
    if (dtype == 2) fma2_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, 0.5f, sink);
    else if (dtype == 3) fma3reg_kernel<false><<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, 0.5f, sink);
    else if (dtype == 4) fma3reg_kernel<true><<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, 0.5f, sink);
    else if (dtype == 5) pipe_mix_kernel<1><<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, 0.5f, sink);
    else if (dtype == 6) pipe_mix_kernel<2><<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, 0.5f, sink);
    else if (dtype == 7) pipe_mix_kernel<3><<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, 0.5f, sink);
    else if (dtype >= 8) {
        MixParams prm;
        for (int i = 0; i < 8; i++) prm.u[i] = 1.0f + 1e-7f * (float)i;
        if (dtype == 8) pipe_form_kernel<1><<<blocks, threads, 0, (cudaStream_t)stream>>>(prm, iters, 0.5f, sink);
        else if (dtype == 9) pipe_form_kernel<2><<<blocks, threads, 0, (cudaStream_t)stream>>>(prm, iters, 0.5f, sink);
        else if (dtype == 10) pipe_form_kernel<3><<<blocks, threads, 0, (cudaStream_t)stream>>>(prm, iters, 0.5f, sink);
        else pipe_form_kernel<4><<<blocks, threads, 0, (cudaStream_t)stream>>>(prm, iters, 0.5f, sink);
    }
    else if (dtype == SRE_F32) fma_kernel<float><<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, 0.5f, sink);
    else fma_kernel<double><<<blocks, threads, 0, (cudaStream_t)stream>>>(iters, 0.5, reinterpret_cast<double *>(sink));
    if (flops) *flops = 2.0 * 32.0 * (double)iters * (double)blocks * (double)threads;
    SRE_CUDA(cudaGetLastError());
    return 0;
}

int sre_bench_hbm_write(void *dst, int64_t bytes, void *stream) {
    SRE_CHECK_ARG(dst && bytes >= 16 && (reinterpret_cast<size_t>(dst) % 16) == 0, SRE_E_ARG, "bad buffer");
    hbm_write_kernel<<<148 * 8, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<float4 *>(dst), bytes / 16);
    SRE_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
