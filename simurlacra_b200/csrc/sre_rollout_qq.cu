// sre_rollout_qq.cu -- instantiates the rollout kernel for env kind 0 (qq), float and double builds.
#include "sre_host.h"

namespace sre {
int rollout_f32_0(const KCfg<float> &cfg, const KDistr &d, KPolicy pol, const KIO &io, cudaStream_t s) {
    return launch_rollout<0, float>(cfg, d, pol, io, s);
}
int rollout_f64_0(const KCfg<double> &cfg, const KDistr &d, KPolicy pol, const KIO &io, cudaStream_t s) {
    return launch_rollout<0, double>(cfg, d, pol, io, s);
}
}  // namespace sre
