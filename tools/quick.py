"""tools/quick.py -- one device-timed line for a kernel experiment (not the contract bench).

  python tools/quick.py [--workload qq|qcp|qbb] [--prec 0|1|2] [--envs N] [--reps K] [--record] [--hidden 64] [--policy bench|sweep]

Prints env-steps/s of the rollout launch (CUDA events, L2 flushed between launches) plus the launch statistics, so that two
builds (SRE_B200_LIB=...) can be compared on identical work: same seeds => same episodes, returns and step counts.
"""
import argparse
import json
import os.path as osp
import sys

import numpy as np

ROOT = osp.dirname(osp.dirname(osp.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="qq")
    ap.add_argument("--prec", type=int, default=1)
    ap.add_argument("--envs", type=int, default=1 << 20)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--hidden", type=int, default=64)
    ap.add_argument("--record", action="store_true")
    ap.add_argument("--policy", default="bench")
    ap.add_argument("--tag", default="")
    ap.add_argument("--no-dr", action="store_true")
    ap.add_argument("--wseed", type=int, default=-1)
    ap.add_argument("--host-weights", action="store_true", help="also pass the parameters as a host copy (sre_policy.weights_host)")
    a = ap.parse_args()
    import torch

    from bench import fnn_weights, qq_randomizer, time_launches
    from simurlacra_b200 import _lib as L
    from simurlacra_b200 import engine as E
    from simurlacra_b200.domain_randomization import DomainRandomizer, UniformDomainParam
    from simurlacra_b200.domain_randomization.default_randomizers import create_default_randomizer_qbb
    from simurlacra_b200.environments.pysim.quanser_ball_balancer import QBallBalancerSim
    from simurlacra_b200.environments.pysim.quanser_cartpole import QCartPoleSwingUpSim

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    hid = a.hidden
    if a.workload == "qq":
        kind, dt, H, n_obs, n_act, kw = L.QQ, 1 / 250.0, 600, 6, 1, {}
        distr = E.make_distr_table(qq_randomizer(E.nominal_params(L.QQ)))
    elif a.workload == "qcp":
        kind, dt, H, n_obs, n_act, kw = L.QCP, 1 / 250.0, 600, 5, 1, dict(long_pole=True)
        nom = QCartPoleSwingUpSim.get_nominal_domain_param(long=True)
        rz = DomainRandomizer(*[UniformDomainParam(name=k, mean=nom[k], halfspan=nom[k] / 5) for k in
                                ("cart_mass", "pole_mass", "pole_length", "pole_damping", "combined_damping", "cart_friction_coeff")])
        distr = rz.table(QCartPoleSwingUpSim._param_names)
    else:
        kind, dt, H, n_obs, n_act, kw = L.QBB, 0.01, 500, 8, 2, {}
        distr = create_default_randomizer_qbb().table(QBallBalancerSim._param_names)
        if a.hidden == 64:
            hid = 16
    sizes = [n_obs, hid, hid, n_act]
    w = torch.from_numpy(fnn_weights(sizes, seed=a.wseed if a.wseed >= 0 else (0 if a.policy == "bench" else 3))).to(dev)
    if a.policy == "zero":  # zero output layer: the action is 0, nothing fails, episodes run to max_steps (restart-free rate)
        w[-(hid * n_act + n_act):] = 0.0
    if a.no_dr:
        distr = None
    cfg = E.make_cfg(kind, dt, H, dtype=torch.float32, act_norm=True, **kw)
    pol = E.make_policy(L.POL_FNN, w, sizes=sizes, precision=a.prec, weights_host=w.cpu() if a.host_weights else None)
    extra = {}
    if a.record:
        lay = E.record_layout(kind)
        extra = dict(traj=torch.empty(H + 1, a.envs, lay["width"], dtype=torch.float32, device=dev),
                     row_flag=torch.empty(H + 1, a.envs, dtype=torch.uint8, device=dev))

    def step(i):
        return E.rollout(cfg, pol, a.envs, H + 1, device=dev, distr=distr, seed=i, auto_reset=True, record=a.record,
                         **extra)["stats"]

    for i in range(3):
        step(i)
    t, st = time_launches(torch, lambda i: step(10 + i), a.reps, flush)
    steps = sum(float(s[L.STAT_STEPS]) for s in st)
    s0 = st[0].cpu().numpy()
    out = dict(tag=a.tag, workload=a.workload, prec=a.prec, envs=a.envs, record=a.record, hidden=hid,
               env_steps_per_s=steps / t, ms_per_launch=1e3 * t / a.reps, episodes=float(s0[L.STAT_EPISODES]),
               mean_ret=float(s0[L.STAT_SUM_RET] / max(s0[L.STAT_EPISODES], 1)), steps=float(s0[L.STAT_STEPS]),
               failed=float(s0[L.STAT_FAILED]))
    if a.record:
        out["record_gbs"] = steps / t * (E.record_layout(kind)["width"] * 4 + 1) / 1e9
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
