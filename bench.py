"""
bench.py -- env-steps/s of the pysim rollout hot path (BASELINE.json metric) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload qq|qcp|qbb] [--envs M]

A "step" is one pass of the hot path over one batch: ONE launch of the persistent-horizon kernel that takes
2^20 environments per GPU through 600 simulation steps (config 2 of BASELINE.json: QQubeSwingUpSim, dt = 1/250,
ActNormWrapper, DomainRandWrapperLive with the default qq randomizer, fused FNN [64,64] tanh policy,
auto-reset so that every lane steps for the whole horizon).  Prints ONE JSON line (rank 0).

 value      env-steps/s, whole job, inputs resident in HBM (device-timed with CUDA events, max over ranks)
 e2e        the same through the public sampler call with HOST buffers: policy parameters copied from pinned host
            memory every step, returns / lengths / statistics read back to the host every step
 roofline   of the rollout kernel: algorithmic FLOP per env-step (9 337) x env-steps/s over the TF32
            tensor peak (half of MEASURED_PEAKS.json's bf16 figure) when the hidden layer runs on tcgen05;
 roofline_fp32  the same FLOP over the FP32-FMA peak MEASURED live with an FMA-only kernel (MEASURED_PEAKS.json
            carries no FP32 figure) -- the CUDA-core view, and the primary roofline with --precision 0
 roofline_traj  HBM roofline of the trajectory-write workload (QBB records)
 variants   the other arithmetic (--precision 0: packed fp32 FMA on the CUDA cores)
 cpu_baseline  the CPU oracle port (oracle/sre_oracle.c, OpenMP) on a bounded sample of the same workload

--impl reference times the reference's CPU implementation of the path.  The reference is Python and does not
exist on the GPU box, so this arm runs the C restatement of it (oracle/, `kind: "port"`) on all host threads.
"""
import argparse
import json
import os
import os.path as osp
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = osp.dirname(osp.abspath(__file__))
sys.path.insert(0, ROOT)

QQ_NAMES = ["gravity_const", "motor_resistance", "motor_back_emf", "mass_rot_pole", "length_rot_pole",
            "damping_rot_pole", "mass_pend_pole", "length_pend_pole", "damping_pend_pole"]

# algorithmic work per env-step (FLOP convention stated there): env + fused FNN [64,64]
FLOP_PER_STEP = {"qq": 120 + 9217, "qcp": 375 + 9089, "qbb": 260 + 2 * (8 * 16 + 16 * 16 + 16 * 2) + 34}
RECORD_BYTES = {"qq": 53, "qcp": 53, "qbb": 93}
# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of rollout_kernel<QQ, float, tcgen05> at the default size
# (2^20 envs x 600 steps), from the ncu --set full capture summarised in profiles/r01_n_final_rollout_qq_fnn64_1M.txt
# (153 856 B read + 5 018 880 B written: returns, lengths, first-episode parameters; the state never leaves the registers)
NCU_TRAFFIC = {(1 << 20, 600, 64, 1): 153856 + 5018880}


def qq_randomizer(nom):
    """create_default_randomizer_qq (domain_randomization/default_randomizers.py:376-416) as a table."""
    div = dict(gravity_const=10, damping_rot_pole=4, damping_pend_pole=4)
    clip = dict(gravity_const=1e-3, motor_resistance=1e-3, damping_rot_pole=1e-9, damping_pend_pole=1e-9)
    return [dict(param_index=i, kind="normal", a=nom[i], b=nom[i] / div.get(k, 5), clip_lo=clip.get(k, 1e-4))
            for i, k in enumerate(QQ_NAMES)]


def fnn_weights(sizes, seed=0, scale=0.3):
    rng = np.random.RandomState(seed)
    return (rng.randn(sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(len(sizes) - 1))) * scale).astype(np.float32)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.p = gpu_index, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                       "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.p is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.25)
        self.p.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 8:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


def cpu_port_run(workload, n_envs, horizon, threads, seed=0):
    """The oracle port on the host cores: same workload (DR + FNN [64,64] + ActNorm), bounded sample."""
    sys.path.insert(0, osp.join(ROOT, "oracle"))
    import oracle as O

    assert workload == "qq"
    sizes = [6, 64, 64, 1]
    w = fnn_weights(sizes)
    pol = O.make_policy(O.POL_FNN, w, sizes=sizes)
    tab = O.make_distr_table(qq_randomizer(O.nominal_params(O.QQ)))
    t0 = time.perf_counter()
    res = O.rollout_batch(O.QQ, 1 / 250.0, horizon, n_envs, horizon, policy=pol, act_norm=True, distr=tab, seed=seed,
                          record=False, n_threads=threads)
    dt = time.perf_counter() - t0
    return int(res["steps"]), dt


def cpu_port_bounded(workload, horizon, threads, target_s=12.0):
    cpu_port_run(workload, 64 * threads, horizon, threads)  # thread-pool warm-up
    steps, dt = cpu_port_run(workload, 2048 * threads, horizon, threads)  # calibration (about a second)
    rate = steps / dt
    n = int(max(256 * threads, min(1 << 21, rate * target_s / horizon)))
    steps, dt = cpu_port_run(workload, n, horizon, threads)
    if dt < 0.7 * target_s:  # the calibration under-estimated the rate: one more, longer run (10-30 s of CPU work)
        n = int(min(1 << 22, n * target_s / max(dt, 1e-3)))
        steps, dt = cpu_port_run(workload, n, horizon, threads)
    return steps / dt, f"{n} envs x {horizon} steps (QQube + DR + FNN[64,64]) in {dt:.1f} s on {threads} threads", n


def run_reference_arm(args):
    """--impl reference: the CPU implementation of the path on the host cores (the oracle port; see module doc)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    horizon = 600
    per_step_s = []
    sample = ""
    rate0, _, n = cpu_port_bounded(args.workload, horizon, threads, target_s=6.0)
    for i in range(args.warmup + args.steps):
        steps, dt = cpu_port_run(args.workload, n, horizon, threads, seed=i)
        if i >= args.warmup:
            per_step_s.append((steps, dt))
    tot_steps = sum(s for s, _ in per_step_s)
    tot_t = sum(t for _, t in per_step_s)
    value = tot_steps / tot_t
    sample = f"{n} envs x {horizon} steps per bench step, {threads} OpenMP threads"
    line = dict(impl="reference", metric="env-steps/sec", value=value, unit="env-steps/s", n_gpus=args.gpus, steps=args.steps,
                warmup=args.warmup, ms_per_step=1e3 * tot_t / max(1, args.steps), higher_is_better=True, scaling="weak",
                vs_baseline=None, dtype="f64", data="synthetic",
                config=dict(workload="QQubeSwingUpSim dt=1/250 H=600 + ActNorm + DomainRandWrapperLive(default qq) + fused FNN[64,64]",
                            envs_per_gpu=args.envs, horizon=horizon, auto_reset=True, trajectory="returns+lengths+stats",
                            reference_sample_envs_per_step=n),
                cpu_baseline=dict(value=value, unit="env-steps/s", cores=threads, kind="port", sample=sample),
                e2e=dict(value=value, unit="env-steps/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                gpu_launches=0)
    print(json.dumps(line), flush=True)


def measure_fma_peak(E, L, torch, dev, dtype_code):
    """FMA-only kernel, all SMs, 32 independent chains per thread: the FP32 (or FP64) roofline denominator."""
    import ctypes as C

    sink = torch.zeros(4, dtype=torch.float32, device=dev)
    flops = C.c_double()
    blocks, threads, iters = 148 * 8, 256, 4096
    best = 0.0
    s = torch.cuda.current_stream().cuda_stream
    for it in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        L.check(L.lib().sre_bench_fma(dtype_code, iters, blocks, threads, C.c_void_p(sink.data_ptr()), C.byref(flops),
                                      C.c_void_p(s)), "sre_bench_fma")
        e1.record()
        torch.cuda.synchronize()
        if it >= 2:
            best = max(best, flops.value / (e0.elapsed_time(e1) * 1e-3))
    return best / 1e12


def measured_peaks():
    """MEASURED_PEAKS.json (driver-written) or the fallback B200_PROFILING.md states."""
    path = osp.join(ROOT, "MEASURED_PEAKS.json")
    if osp.exists(path):
        d = json.load(open(path))
        return dict(hbm_gbs=float(d["hbm_gbs"]), bf16_tflops=float(d["bf16_tflops_sustained"]), source="measured")
    return dict(hbm_gbs=6650.0, bf16_tflops=1400.0, source="fallback")


def time_launches(torch, fn, k, flush):
    """CUDA events around each of k launches on torch's current stream (the one the C-ABI launches on)."""
    tot = 0.0
    outs = []
    for i in range(k):
        flush.fill_(i & 0xFF)  # L2 flush between timed iterations (not timed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        outs.append(fn(i))
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1) * 1e-3
    return tot, outs


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="qq")
    ap.add_argument("--envs", type=int, default=1 << 20, help="environments per GPU")
    ap.add_argument("--horizon", type=int, default=600)
    ap.add_argument("--hidden", type=int, default=64)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the variant / trajectory-write measurements")
    ap.add_argument("--precision", type=int, default=1,
                    help="hidden layer of the fused policy: 0 fp32 FMA (CUDA cores), 1 tcgen05 3xTF32 (fp32-level "
                         "accuracy, default), 2 tcgen05 single TF32")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch
    import torch.distributed as dist

    import simurlacra_b200 as pyrado
    import simurlacra_b200.engine as E
    from simurlacra_b200 import _lib as L
    from simurlacra_b200.domain_randomization.default_randomizers import (create_default_randomizer_qbb,
                                                                           create_default_randomizer_qq)
    from simurlacra_b200.environment_wrappers import ActNormWrapper, DomainRandWrapperLive
    from simurlacra_b200.environments.pysim.quanser_ball_balancer import QBallBalancerSim
    from simurlacra_b200.environments.pysim.quanser_qube import QQubeSwingUpSim
    from simurlacra_b200.policies import FNNPolicy
    from simurlacra_b200.sampling import ParallelRolloutSampler

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert args.warmup >= 3, "timing rules: at least 3 warm-up steps"

    N, H, hid = args.envs, args.horizon, args.hidden
    T = H + 1
    sizes = [6, hid, hid, 1]
    prec = args.precision if hid == 64 else 0
    # ---- the workload through the reference-facing API (config C2 of BASELINE.json) ----
    sim = QQubeSwingUpSim(dt=1 / 250.0, max_steps=H, device=dev, dtype=torch.float32)
    env = DomainRandWrapperLive(ActNormWrapper(sim), create_default_randomizer_qq())
    policy = FNNPolicy(env.spec, hidden_sizes=[hid, hid], hidden_nonlin=torch.tanh)  # parameters live on the host
    policy.param_values = torch.from_numpy(fnn_weights(sizes))
    policy.engine_precision = prec
    sampler = ParallelRolloutSampler(env, policy, num_workers=1, min_rollouts=1, seed=0)
    # device-resident descriptors for the `value` measurement
    w_dev = torch.from_numpy(fnn_weights(sizes)).to(dev)
    cfg = E.make_cfg(L.QQ, 1 / 250.0, H, dtype=torch.float32, act_norm=True)
    distr = E.make_distr_table(qq_randomizer(E.nominal_params(L.QQ)))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    launches = [0]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def resident_step(i, precision):
        pol = E.make_policy(L.POL_FNN, w_dev, sizes=sizes, precision=precision)
        res = E.rollout(cfg, pol, N, T, device=dev, distr=distr, seed=i, env_offset=rank * N, auto_reset=True,
                        record=False)
        launches[0] += L.lib().sre_last_launch_count()
        if world > 1:  # the statistics the learner logs: one small NCCL all-reduce
            dist.all_reduce(res["stats"][:6])
        return res["stats"]

    def e2e_step(i):
        """The call a user makes: policy parameters on the host (pinned staging -> H2D inside), results to the host."""
        res = sampler.sample_batch(N, auto_reset=True)
        return res["ret"].cpu(), res["length"].cpu(), sampler.last_stats.cpu()

    def timed(fn, k):
        barrier()
        t, outs = time_launches(torch, fn, k, flush)
        barrier()
        if world > 1:
            tt = torch.tensor([t], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t = float(tt)
        return t, outs

    # ---- device-resident timing (value) ----
    for i in range(args.warmup):
        resident_step(i, prec)
    clocks = ClockSampler(local)
    launches[0] = 0
    if rank == 0:
        clocks.start()
    t_dev, stats = timed(lambda i: resident_step(args.warmup + i, prec), args.steps)
    clk = clocks.stop() if rank == 0 else None
    n_launches = launches[0]
    total_env_steps = sum(float(s[L.STAT_STEPS]) for s in stats) * (1 if world > 1 else 1)  # all-reduced: whole job
    value = total_env_steps / t_dev

    # ---- end-to-end through the public sampler call with host buffers ----
    for i in range(2):
        e2e_step(i)
    barrier()
    t0 = time.perf_counter()
    e2e_steps = 0.0
    for i in range(args.steps):
        ret_h, len_h, st_h = e2e_step(i)
        e2e_steps += float(st_h[L.STAT_STEPS])
    barrier()
    t_e2e = time.perf_counter() - t0
    if world > 1:
        tt = torch.tensor([t_e2e], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_e2e = float(tt)
    e2e_value = e2e_steps / t_e2e

    extras = {}
    if rank == 0:
        peaks = measured_peaks()
        fma_peak = measure_fma_peak(E, L, torch, dev, L.F32)
        per_rank_rate = value / world
        flop = FLOP_PER_STEP["qq"] if hid == 64 else 120 + 2 * (6 * hid + hid * hid + hid) + 2 * hid + 1
        achieved = per_rank_rate * flop / 1e12
        prec_name = {0: "fp32 FMA (CUDA cores)", 1: "tcgen05 3xTF32 split (fp32-level accuracy)", 2: "tcgen05 TF32"}[prec]
        fp32_roof = dict(bound="fp32_fma", achieved=achieved, peak=fma_peak, unit="TFLOP/s", frac=achieved / fma_peak,
                         traffic=None, flop_per_env_step=flop,
                         peak_source="measured live: FMA-only kernel (sre_bench_fma), best of 4; MEASURED_PEAKS.json has "
                                     "no FP32 figure",
                         note="algorithmic FLOP (SURVEY 8d) of env + FNN per env-step over the FP32 CUDA-core FMA peak; "
                              "with the hidden layer on the tensor cores this fraction can exceed 1")
        if prec > 0:
            # the dominant kernel runs its 64x64 layer on tcgen05: report it against the tensor roofline (TF32 dense =
            # half the measured bf16 rate); `achieved` is ALGORITHMIC flop (one fp32 product per weight), `issued`
            # counts the 3 TF32 passes of the split actually executed.
            passes = 3 if prec == 1 else 1
            t_peak = peaks["bf16_tflops"] / 2
            t_iss = per_rank_rate * passes * 2 * hid * hid / 1e12
            roofline = dict(bound="tensor", achieved=achieved, peak=t_peak, unit="TFLOP/s", frac=achieved / t_peak,
                            traffic=NCU_TRAFFIC.get((N, H, hid, prec)),
                            traffic_source="profiles/r01_n_final_rollout_qq_fnn64_1M.txt (bytes per launch)", flop_per_env_step=flop, issued=t_iss, frac_issued=t_iss / t_peak,
                            peak_source=f"MEASURED_PEAKS.json bf16_tflops_sustained ({peaks['source']}) / 2: kind::tf32 "
                                        f"runs at half the bf16 rate",
                            note=f"{passes} tcgen05.mma kind::tf32 pass(es) per 64x64 layer (3xTF32 split keeps fp32-level "
                                 f"accuracy); per env-step only 8192 of {flop} FLOP are GEMM-shaped, the rest (layer 1, "
                                 f"128 tanh, layer 3, dynamics, sincos, reward) runs on the CUDA cores: ncu shows issue "
                                 f"slots 56% and the MUFU pipe 51% busy, tensor pipe 24% (profiles/r01_n_*), so the "
                                 f"kernel is CUDA-core/MUFU bound -- see roofline_fp32")
            extras["roofline_fp32"] = fp32_roof
        else:
            roofline = fp32_roof
        line = dict(metric="env-steps/sec", value=value, unit="env-steps/s", n_gpus=world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=1e3 * t_dev / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None,
                    dtype="f32", data="synthetic",
                    config=dict(workload=f"QQubeSwingUpSim dt=1/250 H={H} + ActNorm + DomainRandWrapperLive(default qq) + fused FNN[{hid},{hid}]",
                                envs_per_gpu=N, horizon=H, auto_reset=True, trajectory="returns+lengths+stats",
                                policy_hidden_layer=prec_name,
                                l2="flushed between timed iterations (256 MB fill)", parallelism=f"env-shard x{world}"),
                    e2e=dict(value=e2e_value, unit="env-steps/s", h2d_bytes_per_step=int(policy.num_param * 4),
                             d2h_bytes_per_step=int(N * 8 + 64)),
                    gpu_launches=n_launches, clocks=clk, roofline=roofline)

    # ---- extras on one GPU: the other arithmetic variant, and the trajectory-write (HBM) workload ----
    if world == 1 and not args.no_extras:
        other = 0 if prec else 1
        if hid == 64:
            for i in range(3):
                resident_step(i, other)
            t_o, st_o = timed(lambda i: resident_step(50 + i, other), args.steps)
            rate_o = sum(float(s[L.STAT_STEPS]) for s in st_o) / t_o
            extras["variants"] = {("fp32_fma" if other == 0 else "tcgen05_tf32x3"): dict(
                value=rate_o, unit="env-steps/s", fma_roofline_frac=rate_o * flop / 1e12 / fma_peak)}

            # the same launch with step-based exploration in the loop (NormalActNoiseExplStrat, std 0.1): one Philox
            # block + one Box-Muller pair per env-step on top of the policy
            def explore_step(i):
                pol = E.make_policy(L.POL_FNN, w_dev, sizes=sizes, precision=prec, act_noise=L.ACT_NOISE_NORMAL,
                                    act_noise_scale=[0.1])
                return E.rollout(cfg, pol, N, T, device=dev, distr=distr, seed=i, auto_reset=True, record=False)["stats"]

            for i in range(3):
                explore_step(i)
            t_x, st_x = timed(lambda i: explore_step(70 + i), args.steps)
            extras["variants"]["normal_act_noise"] = dict(value=sum(float(s[L.STAT_STEPS]) for s in st_x) / t_x,
                                                          unit="env-steps/s", note="headline arithmetic + exploration noise")
        # config C4: QBallBalancerSim + DR (20 params) + cheap fused policy, FULL StepSequence records written to HBM
        qn, qh = min(N, 1 << 20), 500
        qcfg = E.make_cfg(L.QBB, 0.01, qh, dtype=torch.float32, act_norm=True)
        qdistr = create_default_randomizer_qbb().table(QBallBalancerSim._param_names)
        qsizes = [8, 16, 16, 2]
        qpol = E.make_policy(L.POL_FNN, torch.from_numpy(fnn_weights(qsizes, seed=1)).to(dev), sizes=qsizes)
        lay = E.record_layout(L.QBB)
        traj = torch.empty(qh + 1, qn, lay["width"], dtype=torch.float32, device=dev)
        flag = torch.empty(qh + 1, qn, dtype=torch.uint8, device=dev)

        def qstep(i):
            return E.rollout(qcfg, qpol, qn, qh + 1, device=dev, distr=qdistr, seed=i, auto_reset=True, record=True,
                             traj=traj, row_flag=flag)["stats"]

        for i in range(2):
            qstep(i)
        t_q, st_q = timed(lambda i: qstep(10 + i), 3)
        bytes_per_launch = (qh + 1) * qn * (lay["width"] * 4 + 1)
        gbs = 3 * bytes_per_launch / t_q / 1e9
        extras["roofline_traj"] = dict(bound="hbm", achieved=gbs, peak=peaks["hbm_gbs"], unit="GB/s", frac=gbs / peaks["hbm_gbs"],
                                       traffic=None, peak_source=f"MEASURED_PEAKS.json hbm_gbs ({peaks['source']})",
                                       workload=f"QBallBalancerSim dt=0.01 H={qh} + ActNorm + DR(default qbb, 20 params) + fused "
                                                f"FNN[16,16], {qn} envs, records [T,N,{lay['width']}] f32 + flags written",
                                       bytes_per_env_step=lay["width"] * 4 + 1,
                                       env_steps_per_s=sum(float(s[L.STAT_STEPS]) for s in st_q) / t_q)
        del traj, flag
        # post-processing of a full record set: returns-to-go + GAE over [601, N] contiguous
        # rewards / values / flags -- a purely HBM-bound pass: 9 B read + 8 B written per row and lane
        rT = 601
        g = torch.Generator(device=dev).manual_seed(0)
        r_rew = torch.randn(rT, qn, device=dev, generator=g)
        r_val = torch.randn(rT, qn, device=dev, generator=g)
        r_flag = torch.zeros(rT, qn, dtype=torch.uint8, device=dev)
        r_flag[200::201] = L.ROW_TERMINAL
        r_flag[199::201] = L.ROW_DONE
        o_ret, o_adv = torch.empty_like(r_rew), torch.empty_like(r_rew)

        def ret_step(i):
            E.returns(r_rew, r_flag, 0.99, values=r_val, lamb=0.95, out_ret=o_ret, out_adv=o_adv)

        for i in range(2):
            ret_step(i)
        t_r, _ = timed(ret_step, 3)
        rb = 3 * rT * qn * 17 / t_r / 1e9
        extras["roofline_returns"] = dict(bound="hbm", achieved=rb, peak=peaks["hbm_gbs"], unit="GB/s", frac=rb / peaks["hbm_gbs"],
                                          traffic=(5671760000 + 5001247000) if (rT, qn) == (601, 1 << 20) else None,
                                          traffic_source="profiles/r01_h_returns_gae_601x1M.txt (dram bytes read + written per launch)", peak_source=f"MEASURED_PEAKS.json hbm_gbs ({peaks['source']})",
                                          workload=f"sre_returns (discounted returns + GAE), [{rT}, {qn}] f32 rewards / values / "
                                                   f"flags resident, 17 B per row and lane", rows_per_s=3 * rT * qn / t_r)
        del r_rew, r_val, r_flag, o_ret, o_adv

    if rank == 0:
        line.update(extras)
        if not args.no_cpu_baseline and world == 1:
            threads = os.cpu_count() or 1
            v, sample, _ = cpu_port_bounded("qq", H, threads)
            line["cpu_baseline"] = dict(value=v, unit="env-steps/s", cores=threads, kind="port", sample=sample)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
