// sre_rollout_bob.cu -- instantiates the rollout kernel for env kind 3 (bob), float and double builds.
#include "sre_host.h"

namespace sre {
int rollout_f32_3(const KCfg<float> &cfg, const KDistr &d, KPolicy pol, const KIO &io, cudaStream_t s) {
    return launch_rollout<3, float>(cfg, d, pol, io, s);
}
int rollout_f64_3(const KCfg<double> &cfg, const KDistr &d, KPolicy pol, const KIO &io, cudaStream_t s) {
    return launch_rollout<3, double>(cfg, d, pol, io, s);
}
}  // namespace sre
