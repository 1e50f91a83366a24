// sre_rollout_qbb.cu -- instantiates the rollout kernel for env kind 2 (qbb), float and double builds.
#include "sre_host.h"

namespace sre {
int rollout_f32_2(const KCfg<float> &cfg, const KDistr &d, KPolicy pol, const KIO &io, cudaStream_t s) {
    return launch_rollout<2, float>(cfg, d, pol, io, s);
}
int rollout_f64_2(const KCfg<double> &cfg, const KDistr &d, KPolicy pol, const KIO &io, cudaStream_t s) {
    return launch_rollout<2, double>(cfg, d, pol, io, s);
}
}  // namespace sre
