"""Small end-to-end case that launches every kernel family once (debugging aid; compute-sanitizer is closed on this pool): the (FMA, tcgen05, records, auto-reset,
exploration noise, QBB cooperative restart), the returns kernel and the adjoint.  python tools/kernel_families_case.py"""
import os.path as osp
import sys

import numpy as np
import torch

sys.path.insert(0, osp.dirname(osp.dirname(osp.abspath(__file__))))
from bench import fnn_weights  # noqa: E402
from simurlacra_b200 import _lib as L  # noqa: E402
from simurlacra_b200 import engine as E  # noqa: E402
from simurlacra_b200.domain_randomization.default_randomizers import create_default_randomizer_qbb, create_default_randomizer_qq  # noqa: E402
from simurlacra_b200.environments.pysim.quanser_ball_balancer import QBallBalancerSim  # noqa: E402
from simurlacra_b200.environments.pysim.quanser_qube import QQubeSwingUpSim  # noqa: E402

dev = torch.device("cuda", 0)
N = 1000  # ragged: not a multiple of 32 / 128
for kind, dt, H, sizes, distr in ((L.QQ, 1 / 250.0, 60, [6, 64, 64, 1], create_default_randomizer_qq().table(QQubeSwingUpSim._param_names)),
                                  (L.QBB, 0.01, 60, [8, 64, 64, 2], create_default_randomizer_qbb().table(QBallBalancerSim._param_names)),
                                  (L.QBB, 0.01, 60, [8, 16, 16, 2], create_default_randomizer_qbb().table(QBallBalancerSim._param_names))):
    w = torch.from_numpy(fnn_weights(sizes, seed=1) * 3).to(dev)
    cfg = E.make_cfg(kind, dt, H, dtype=torch.float32, act_norm=True)
    for prec in (L.PREC_FP32, L.PREC_TF32X3):
        for record in (False, True):
            pol = E.make_policy(L.POL_FNN, w, sizes=sizes, precision=prec, act_noise=L.ACT_NOISE_NORMAL, act_noise_scale=[0.2, 0.2])
            res = E.rollout(cfg, pol, N, H + 1, device=dev, distr=distr, seed=1, auto_reset=True, record=record)
            torch.cuda.synchronize()
            print(kind, sizes, prec, record, float(res["stats"][L.STAT_STEPS]), float(res["stats"][0]))
    if record:
        ret, adv = E.returns(res["rew"], res["row_flag"], 0.99, values=torch.zeros_like(res["rew"]))
        ret2, _ = E.returns(res["rew"].contiguous()[:, :996], res["row_flag"][:, :996].contiguous(), 0.99)
        torch.cuda.synchronize()
cfg = E.make_cfg(L.QCP, 1 / 20.0, 50, dtype=torch.float32, long_pole=True, wild_init=1)
s0 = torch.zeros(N, 4, device=dev)
act = torch.ones(50, N, device=dev)
st, thdd, rew, done = E.qcp_diff_forward(cfg, None, s0, act)
E.qcp_diff_backward(cfg, None, st, thdd, act, g_rew=torch.ones(50, N, device=dev))
torch.cuda.synchronize()
print("all kernel families ran")
