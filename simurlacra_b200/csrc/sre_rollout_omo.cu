// sre_rollout_omo.cu -- instantiates the rollout kernel for env kind 4 (omo), float and double builds.
#include "sre_host.h"

namespace sre {
int rollout_f32_4(const KCfg<float> &cfg, const KDistr &d, KPolicy pol, const KIO &io, cudaStream_t s) {
    return launch_rollout<4, float>(cfg, d, pol, io, s);
}
int rollout_f64_4(const KCfg<double> &cfg, const KDistr &d, KPolicy pol, const KIO &io, cudaStream_t s) {
    return launch_rollout<4, double>(cfg, d, pol, io, s);
}
}  // namespace sre
