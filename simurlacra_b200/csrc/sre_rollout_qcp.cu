// sre_rollout_qcp.cu -- instantiates the rollout kernel for env kind 1 (qcp), float and double builds.
#include "sre_host.h"

namespace sre {
int rollout_f32_1(const KCfg<float> &cfg, const KDistr &d, KPolicy pol, const KIO &io, cudaStream_t s) {
    return launch_rollout<1, float>(cfg, d, pol, io, s);
}
int rollout_f64_1(const KCfg<double> &cfg, const KDistr &d, KPolicy pol, const KIO &io, cudaStream_t s) {
    return launch_rollout<1, double>(cfg, d, pol, io, s);
}
}  // namespace sre
