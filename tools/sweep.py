"""
tools/sweep.py -- the BASELINE.json configs C2..C5 and the env-count sweep (1K .. 16M envs) on ONE B200.

  python tools/sweep.py [--out gpurun_out/sweep.json] [--quick]

Every line is device-timed with CUDA events after 2 warm-up launches, L2 flushed between launches, through the
C-ABI (`simurlacra_b200.engine`).  This is a reporting tool; the contract bench is bench.py.
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
    ap.add_argument("--out", default=osp.join(ROOT, "gpurun_out", "sweep.json"))
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--only", default="", help="comma-separated prefixes of the configs to run (e.g. F4)")
    args = ap.parse_args()

    import torch

    from bench import fnn_weights, measure_fma_peak, time_launches
    from simurlacra_b200 import _lib as L
    from simurlacra_b200 import engine as E
    from simurlacra_b200.domain_randomization import DomainRandomizer, UniformDomainParam
    from simurlacra_b200.domain_randomization.default_randomizers import (create_default_randomizer_qbb,
                                                                           create_default_randomizer_qq)
    from simurlacra_b200.environments.pysim.quanser_ball_balancer import QBallBalancerSim
    from simurlacra_b200.environments.pysim.quanser_cartpole import QCartPoleSwingUpSim
    from simurlacra_b200.environments.pysim.quanser_qube import QQubeSwingUpSim

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = dict(gpu=torch.cuda.get_device_name(0), lines=[])
    fma_peak = measure_fma_peak(E, L, torch, dev, L.F32)
    out["fp32_fma_peak_tflops"] = fma_peak

    def want(name):
        return not args.only or any(name.startswith(p) for p in args.only.split(","))

    def run(name, cfg, sizes, n, h, distr, precision, reps=3, record=False, flop=None, bytes_per_step=0):
        if not want(name):
            return
        w = torch.from_numpy(fnn_weights(sizes, seed=3)).to(dev)
        pol = E.make_policy(L.POL_FNN, w, sizes=sizes, precision=precision)
        kw = {}
        if record:
            lay = E.record_layout(cfg.kind)
            kw = dict(traj=torch.empty(h + 1, n, lay["width"], dtype=torch.float32, device=dev),
                      row_flag=torch.empty(h + 1, n, dtype=torch.uint8, device=dev))

        def step(i):
            return E.rollout(cfg, pol, n, h + 1, device=dev, distr=distr, seed=i, auto_reset=True, record=record, **kw)["stats"]

        for i in range(2):
            step(i)
        t, st = time_launches(torch, lambda i: step(10 + i), reps, flush)
        steps = sum(float(s[L.STAT_STEPS]) for s in st)
        rate = steps / t
        line = dict(config=name, envs=n, horizon=h, precision=["fp32_fma", "tcgen05_tf32x3", "tcgen05_tf32"][precision],
                    env_steps_per_s=rate, ms_per_launch=1e3 * t / reps)
        if flop:
            line["algorithmic_tflops"] = rate * flop / 1e12
            line["frac_fp32_fma_peak"] = rate * flop / 1e12 / fma_peak
        if bytes_per_step:
            line["record_gbs"] = rate * bytes_per_step / 1e9
        out["lines"].append(line)
        print(json.dumps(line), flush=True)
        del kw

    # ---- C2: QQube + ActNorm + default DR + FNN[64,64]; env-count sweep ----
    qq_cfg = E.make_cfg(L.QQ, 1 / 250.0, 600, dtype=torch.float32, act_norm=True)
    qq_distr = create_default_randomizer_qq().table(QQubeSwingUpSim._param_names)
    exps = (10, 14, 18, 20, 22) if args.quick else (10, 12, 14, 16, 17, 18, 19, 20, 21, 22, 24)
    for e in exps:
        for prec in (1, 0):
            if e == 24 and prec == 0:
                continue
            run("C2 qq+DR+FNN[64,64]", qq_cfg, [6, 64, 64, 1], 1 << e, 600, qq_distr, prec, reps=5 if e <= 20 else 2,
                flop=120 + 9217)
    run("C2 qq+DR+FNN[64,64] single-pass TF32", qq_cfg, [6, 64, 64, 1], 1 << 20, 600, qq_distr, 2, flop=120 + 9217)

    # ---- C3: QCartPoleSwingUp long pole, RK4, 4M envs, uniform DR over pole / cart parameters ----
    nom = QCartPoleSwingUpSim.get_nominal_domain_param(long=True)
    dr = DomainRandomizer(*[UniformDomainParam(name=k, mean=nom[k], halfspan=nom[k] / 5, clip_lo=1e-4)
                            for k in ("cart_mass", "pole_mass", "pole_length", "pole_damping", "combined_damping",
                                      "cart_friction_coeff")])
    qcp_cfg = E.make_cfg(L.QCP, 1 / 250.0, 600, dtype=torch.float32, long_pole=True, act_norm=True, wild_init=0)
    qcp_distr = dr.table(QCartPoleSwingUpSim._param_names)
    n3 = 1 << (20 if args.quick else 22)
    for prec in (1, 0):
        run("C3 qcp long RK4 + uniform DR + FNN[64,64]", qcp_cfg, [5, 64, 64, 1], n3, 600, qcp_distr, prec, reps=2,
            flop=375 + 9089)

    # ---- C4: QBallBalancer + default DR (20 params), full StepSequence records in HBM, 2M envs per GPU ----
    qbb_cfg = E.make_cfg(L.QBB, 0.01, 500, dtype=torch.float32, act_norm=True)
    qbb_distr = create_default_randomizer_qbb().table(QBallBalancerSim._param_names)
    n4 = 1 << (19 if args.quick else 21)
    run("C4 qbb+DR+FNN[16,16] records to HBM", qbb_cfg, [8, 16, 16, 2], n4, 500, qbb_distr, 0, reps=2, record=True,
        flop=260 + 2 * (8 * 16 + 16 * 16 + 16 * 2) + 34, bytes_per_step=93)
    run("C4 qbb+DR+FNN[16,16] no records", qbb_cfg, [8, 16, 16, 2], n4, 500, qbb_distr, 0, reps=2,
        flop=260 + 2 * (8 * 16 + 16 * 16 + 16 * 2) + 34)
    run("C4' qbb+DR+FNN[64,64] records to HBM", qbb_cfg, [8, 64, 64, 2], n4 // 2, 500, qbb_distr, 1, reps=2, record=True,
        bytes_per_step=93)
    torch.cuda.empty_cache()

    # ---- F4: parameter-exploration sampling, a population of policy-parameter sets in one launch ----
    def run_pes(name, sizes, pop, per_set, precision, h=600, reps=3):
        if not want(name):
            return
        nw = sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(len(sizes) - 1))
        g = torch.Generator(device="cpu").manual_seed(1)
        w = (torch.randn(pop, nw, generator=g) * 0.1).to(dev)
        pol = E.make_policy(L.POL_FNN, w, sizes=sizes, precision=precision, group_size=per_set)
        n = pop * per_set

        def step(i):
            return E.rollout(qq_cfg, pol, n, h + 1, device=dev, distr=qq_distr, seed=i, auto_reset=True)["stats"]

        for i in range(2):
            step(i)
        t, st = time_launches(torch, lambda i: step(10 + i), reps, flush)
        steps = sum(float(s[L.STAT_STEPS]) for s in st)
        line = dict(config=name, envs=n, horizon=h, population=pop, rollouts_per_set=per_set,
                    precision=["fp32_fma", "tcgen05_tf32x3", "tcgen05_tf32"][precision], env_steps_per_s=steps / t,
                    ms_per_launch=1e3 * t / reps)
        out["lines"].append(line)
        print(json.dumps(line), flush=True)

    run_pes("F4 PES qq+DR FNN[64,64] 8192 sets x 128 (one set per CTA, tcgen05)", [6, 64, 64, 1], 8192, 128, 1)
    run_pes("F4 PES qq+DR FNN[64,64] 4096 sets x 256 (fp32 FMA)", [6, 64, 64, 1], 4096, 256, 0)
    run_pes("F4 PES qq+DR FNN[16,16] 4096 sets x 256 (fp32 FMA)", [6, 16, 16, 1], 4096, 256, 0)
    run_pes("F4 PES qq+DR FNN[16,16] 32768 sets x 32 (per-lane weights)", [6, 16, 16, 1], 32768, 32, 0)
    run_pes("F4 PES qq+DR FNN[64,64] 16384 sets x 32 (per-lane weights)", [6, 64, 64, 1], 16384, 32, 0, reps=2)
    run_pes("F4 PES qq+DR FNN[16,16] 131072 sets x 8 (per-lane weights)", [6, 16, 16, 1], 131072, 8, 0, reps=2)

    # ---- C5: differentiable cart-pole rollout, forward + adjoint through 300 steps, 256K envs ----
    n5, h5 = 1 << 18, 300
    dcfg = E.make_cfg(L.QCP, 1 / 20.0, h5, dtype=torch.float32, long_pole=True, wild_init=1)
    g = torch.Generator(device="cpu").manual_seed(0)
    s0 = (torch.rand(n5, 4, generator=g) * 0.2 - 0.1).to(dev)
    act = (torch.rand(h5, n5, generator=g) * 2 - 1).to(dev) * 5.0
    g_rew = torch.ones(h5, n5, device=dev)

    def fwd(i):
        return E.qcp_diff_forward(dcfg, None, s0, act)

    def fwd_bwd(i):
        states, thdd, rew, done = E.qcp_diff_forward(dcfg, None, s0, act)
        return E.qcp_diff_backward(dcfg, None, states, thdd, act, g_rew=g_rew)

    for name, fn in (("C5 qcp diff forward", fwd), ("C5 qcp diff forward+adjoint", fwd_bwd)):
        if not want(name):
            continue
        for i in range(2):
            fn(i)
        t, _ = time_launches(torch, lambda i: fn(i) and None, 5, flush)  # results dropped: no allocation in the timed region
        line = dict(config=name, envs=n5, horizon=h5, precision="fp32", env_steps_per_s=5 * n5 * h5 / t,
                    ms_per_launch=1e3 * t / 5)
        out["lines"].append(line)
        print(json.dumps(line), flush=True)

    with open(args.out, "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
