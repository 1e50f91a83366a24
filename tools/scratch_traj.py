import sys, os.path as osp, json
sys.path.insert(0, osp.dirname(osp.dirname(osp.abspath(__file__))))
import torch
from bench import time_launches, fnn_weights
from simurlacra_b200 import _lib as L, engine as E
from simurlacra_b200.domain_randomization.default_randomizers import create_default_randomizer_qbb
from simurlacra_b200.environments.pysim.quanser_ball_balancer import QBallBalancerSim
dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
N = 1 << 20
for name, kind, dt, H, pol_kind in (("omo idle", L.OMO, 0.02, 300, "idle"), ("qq idle", L.QQ, 0.004, 600, "idle"), ("bob idle", L.BOB, 0.01, 500, "idle"),
                                    ("qbb idle", L.QBB, 0.01, 500, "idle"), ("qbb idle simple", L.QBB, 0.01, 500, "idle_simple"),
                                    ("qbb fnn16 DR", L.QBB, 0.01, 500, "fnn")):
    cfg = E.make_cfg(kind, dt, H, dtype=torch.float32, act_norm=True, simple_dynamics=pol_kind == "idle_simple")
    lay = E.record_layout(kind)
    distr = None
    if pol_kind == "fnn":
        sizes = [8, 16, 16, 2]
        pol = E.make_policy(L.POL_FNN, torch.from_numpy(fnn_weights(sizes, seed=1)).to(dev), sizes=sizes)
        distr = create_default_randomizer_qbb().table(QBallBalancerSim._param_names)
    else:
        pol = E.make_policy(L.POL_IDLE)
    traj = torch.empty(H + 1, N, lay["width"], dtype=torch.float32, device=dev)
    flag = torch.empty(H + 1, N, dtype=torch.uint8, device=dev)
    for rec in (True, False):
        f = lambda i: E.rollout(cfg, pol, N, H + 1, device=dev, distr=distr, seed=i, auto_reset=True, record=rec, traj=traj if rec else None, row_flag=flag if rec else None)["stats"]
        f(0); f(1)
        t, st = time_launches(torch, lambda i: f(5 + i), 3, flush)
        s = (sum(st) / 3).cpu().numpy()
        gbs = 3 * (H + 1) * N * (lay["width"] * 4 + 1) / t / 1e9 if rec else 0
        print(json.dumps(dict(name=name, record=rec, ms=1e3 * t / 3, steps_per_s=3 * s[L.STAT_STEPS] / t, episodes=float(s[0]), GBs=gbs)), flush=True)
    del traj, flag
