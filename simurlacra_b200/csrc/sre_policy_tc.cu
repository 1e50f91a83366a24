// sre_policy_tc.cu -- stand-alone batch forward pass of the tensor-core FNN (Policy.forward for a batch with
// sre_policy.precision = SRE_PREC_TF32X3 / SRE_PREC_TF32); also the unit under test for the tcgen05 path.
#include "sre_host.h"
#include "sre_policy_tc.cuh"

namespace sre {

template <int IN, int OUT, bool SP>
__global__ void __launch_bounds__(128) policy_forward_tc_kernel(const __grid_constant__ KPolicy pol, const float *obs,
                                                                float *act, long long N) {
    using F = FnnTC2<IN, OUT, SP>;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    typename F::Ctx c;
    F::setup(c, smem_raw, pol.weights);
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    float x[IN], a[OUT];
#pragma unroll
    for (int i = 0; i < IN; i++) x[i] = n < N ? obs[n * IN + i] : 0.f;
    F::forward(c, x, a, pol.cw);
    if (n < N) {
#pragma unroll
        for (int i = 0; i < OUT; i++) act[n * OUT + i] = a[i];
    }
    F::teardown(c);
}

template <int IN, int OUT, bool SP>
static int launch_pf(const KPolicy &kp, const float *obs, float *act, long long n, cudaStream_t s) {
    using F = FnnTC2<IN, OUT, SP>;
    auto kern = policy_forward_tc_kernel<IN, OUT, SP>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, F::SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
    kern<<<(unsigned)((n + 127) / 128), 128, F::SMEM_BYTES, s>>>(kp, obs, act, n);
    return (int)cudaGetLastError();
}

bool policy_is_tc_shape(const KPolicy &kp, int n_obs, int n_act) {
    return kp.kind == SRE_POL_FNN && kp.precision != SRE_PREC_FP32 && kp.n_layers == 3 && kp.sizes[0] == n_obs &&
           kp.sizes[1] == 64 && kp.sizes[2] == 64 && kp.sizes[3] == n_act && kp.hidden_nonlin == 0 &&
           kp.output_nonlin == 0 && !kp.cartpole_feats;
}

// returns -1000 if the (n_obs, n_act) pair has no tensor-core instantiation (the caller falls back to the FMA kernel)
int policy_forward_tc_dispatch(const KPolicy &kp, const float *obs, int n_obs, float *act, int n_act, long long n,
                               cudaStream_t s) {
    const bool sp = kp.precision == SRE_PREC_TF32;
#define SRE_PF_CASE(I, O)                                                                   \
    if (n_obs == I && n_act == O)                                                           \
        return sp ? launch_pf<I, O, true>(kp, obs, act, n, s) : launch_pf<I, O, false>(kp, obs, act, n, s);
    SRE_PF_CASE(6, 1)
    SRE_PF_CASE(5, 1)
    SRE_PF_CASE(8, 2)
    SRE_PF_CASE(4, 1)
    SRE_PF_CASE(2, 1)
    SRE_PF_CASE(3, 1)
#undef SRE_PF_CASE
    return -1000;
}

}  // namespace sre
