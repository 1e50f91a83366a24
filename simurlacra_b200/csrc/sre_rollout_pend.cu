// sre_rollout_pend.cu -- instantiates the rollout kernel for env kind 5 (pend), float and double builds.
#include "sre_host.h"

namespace sre {
int rollout_f32_5(const KCfg<float> &cfg, const KDistr &d, KPolicy pol, const KIO &io, cudaStream_t s) {
    return launch_rollout<5, float>(cfg, d, pol, io, s);
}
int rollout_f64_5(const KCfg<double> &cfg, const KDistr &d, KPolicy pol, const KIO &io, cudaStream_t s) {
    return launch_rollout<5, double>(cfg, d, pol, io, s);
}
}  // namespace sre
