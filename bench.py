"""
bench.py -- env-steps/s of the pysim rollout hot path (BASELINE.json metric) on N B200s of one node.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload qq|qcp|qbb|diff] [--envs M]

A "step" is one pass of the hot path over one batch: ONE launch of the persistent-horizon kernel.  The headline slot is
config C2 of BASELINE.json (`--workload qq`: QQubeSwingUpSim, dt = 1/250, 600 steps, ActNormWrapper, DomainRandWrapperLive
with the default qq randomizer, fused FNN [64,64] tanh policy, 2^20 envs per GPU, auto-reset so that every lane steps for
the whole horizon); `--workload qcp|qbb|diff` puts C3 / C4 / C5 into the headline slot (same JSON line, e.g. for the
1..8 GPU scaling run of the trajectory-write workload).  Prints ONE JSON line (rank 0).

 value       env-steps/s, whole job, inputs resident in HBM (device-timed with CUDA events, max over ranks)
 e2e         the same through the public sampler call with HOST buffers: policy parameters copied from pinned host memory
             every step, returns / lengths / statistics read back to the host every step
 e2e_sample  the reference-signature call, `ParallelRolloutSampler.sample()` -> List[StepSequence]: the reference
             script's setting min_steps = 30 * 600 (numpy StepSequences) and 2^16 rollouts (torch views of the records)
 roofline    of the headline kernel: algorithmic FLOP (or record bytes) per env-step x env-steps/s over the
             peak -- TF32 tensor peak (half of MEASURED_PEAKS.json's bf16 figure) when the hidden layer runs on tcgen05,
             FP32-FMA peak measured live otherwise, HBM copy bandwidth for the record-writing workload
 roofline_fp32 / roofline_c3 / roofline_traj / roofline_c5 / roofline_returns / c1_latency
             the other BASELINE.json configs (C3 cart-pole RK4 at 2^22 envs, C4 records at 2^21 envs, C5 forward and
             forward + adjoint at 2^18 x 300, C1 one env x 600 steps) and the FP32-FMA view of the headline
 variants    the other arithmetic of the hidden layer, a failure-heavy policy (more resets), exploration noise in the loop
 cpu_baseline  the reference's own ParallelRolloutSampler on the host cores (`kind: "reference"`, baseline/_ref built by
             baseline/make_ref.sh) when it is present, else the C port; the port is always reported as cpu_baseline_port

--impl reference times the reference's CPU implementation of the path: the UNMODIFIED Python reference from baseline/_ref
(B1 sampler on all cores, B2 one worker, B3 env-only hand loop) when that tree exists, else the C restatement (oracle/).
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

# algorithmic work per env-step (FLOP convention stated there): env + fused FNN
ENV_FLOP = {"qq": 120, "qcp": 375, "qbb": 260}
RECORD_BYTES = {"qq": 53, "qcp": 53, "qbb": 93}  # fp32 record + flag byte per env-step
C5_FLOP_FWD = 375         # one RK4 cart-pole step + reward
C5_FLOP_BWD = 2.5 * 375   # adjoint of the same step: reverse sweep through the four stages
C5_TAPE_BYTES = 24        # state (4) + carry + action, written by the forward pass and read back by the adjoint


def fnn_flop(sizes):
    return sum(2 * sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(len(sizes) - 1))


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


# ---------------------------------------------------------------------------------------------------------------------
# CPU arms: the unmodified reference (baseline/_ref) and the C port (oracle/)
# ---------------------------------------------------------------------------------------------------------------------
def cpu_port_run(workload, n_envs, horizon, threads, seed=0):
    """The oracle port on the host cores: same workload (DR + FNN [64,64] + ActNorm), bounded sample."""
    sys.path.insert(0, osp.join(ROOT, "oracle"))
    import oracle as O

    sizes = [6, 64, 64, 1]
    w = fnn_weights(sizes)
    pol = O.make_policy(O.POL_FNN, w, sizes=sizes)
    tab = O.make_distr_table(qq_randomizer(O.nominal_params(O.QQ)))
    t0 = time.perf_counter()
    res = O.rollout_batch(O.QQ, 1 / 250.0, horizon, n_envs, horizon, policy=pol, act_norm=True, distr=tab, seed=seed,
                          record=False, n_threads=threads)
    dt = time.perf_counter() - t0
    return int(res["steps"]), dt


def cpu_port_bounded(workload, horizon, threads, target_s=10.0):
    cpu_port_run(workload, 64 * threads, horizon, threads)  # thread-pool warm-up
    steps, dt = cpu_port_run(workload, 2048 * threads, horizon, threads)  # calibration (about a second)
    rate = steps / dt
    n = int(max(256 * threads, min(1 << 21, rate * target_s / horizon)))
    steps, dt = cpu_port_run(workload, n, horizon, threads)
    if dt < 0.7 * target_s:  # the calibration under-estimated the rate: one more, longer run
        n = int(min(1 << 22, n * target_s / max(dt, 1e-3)))
        steps, dt = cpu_port_run(workload, n, horizon, threads)
    return steps / dt, f"{n} envs x {horizon} steps (QQube + DR + FNN[64,64]) in {dt:.1f} s on {threads} threads", n


REF_DIR = osp.join(ROOT, "baseline", "_ref")


def reference_available():
    return osp.exists(osp.join(REF_DIR, "pyrado", "__init__.py")) and osp.exists(osp.join(REF_DIR, "ref_shim.py"))


def run_ref_runner(extra, timeout=900):
    """One run of baseline/ref_runner.py (the unmodified reference) in a subprocess; returns its JSON or None."""
    env = dict(os.environ, PYTHONPATH=REF_DIR + os.pathsep + os.environ.get("PYTHONPATH", ""), SRE_REF_SHIM="1",
               OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    try:
        p = subprocess.run([sys.executable, osp.join(ROOT, "baseline", "ref_runner.py")] + extra, env=env, cwd="/tmp",
                           capture_output=True, text=True, timeout=timeout)
    except subprocess.TimeoutExpired:
        return None
    for line in p.stdout.splitlines():
        if line.startswith("REF_RESULT "):
            return json.loads(line[len("REF_RESULT "):])
    sys.stderr.write("ref_runner failed:\n" + p.stderr[-2000:] + "\n")
    return None


def reference_sampler_baseline(cores, repeat=3, horizon=600):
    """B1 of SURVEY.md 8(d): ParallelRolloutSampler(num_workers = cores, min_rollouts = 8 * cores), one warm-up sample(),
    `repeat` timed ones."""
    r = run_ref_runner(["--mode", "sampler", "--workers", str(cores), "--repeat", str(repeat), "--horizon", str(horizon)])
    if r is None:
        return None
    return dict(value=r["steps_per_s"], unit="env-steps/s", cores=cores, kind="reference",
                sample=f"unmodified reference (baseline/_ref): ParallelRolloutSampler(num_workers={cores}, min_rollouts="
                       f"{r['min_rollouts']}), QQube + ActNorm + DomainRandWrapperLive + FNN[64,64] on the CPU, {r['repeat']} "
                       f"sample() calls = {r['steps']} env-steps in {r['seconds']:.1f} s (pool start {r['pool_start_s']:.0f} s "
                       f"not counted)")


def run_reference_arm(args):
    """--impl reference: the CPU implementation of the path on the host cores (see module doc)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    threads = os.cpu_count() or 1
    horizon = 600
    cfg = dict(workload="QQubeSwingUpSim dt=1/250 H=600 + ActNorm + DomainRandWrapperLive(default qq) + fused FNN[64,64]",
               envs_per_gpu=args.envs, horizon=horizon, auto_reset=True, trajectory="returns+lengths+stats")
    extra = {}
    if reference_available():
        b1 = run_ref_runner(["--mode", "sampler", "--workers", str(threads), "--repeat", str(args.steps), "--horizon", str(horizon)])
        if b1 is not None:
            b2 = run_ref_runner(["--mode", "sampler", "--workers", "1", "--min-rollouts", "8", "--repeat", "2"])
            b3 = run_ref_runner(["--mode", "hand", "--episodes", "10"])
            value = b1["steps_per_s"]
            sample = (f"unmodified reference: ParallelRolloutSampler(num_workers={threads}, min_rollouts={b1['min_rollouts']}), "
                      f"{b1['repeat']} sample() calls after one warm-up call = {b1['steps']} env-steps in {b1['seconds']:.1f} s")
            cfg["reference_sample_rollouts_per_step"] = b1["min_rollouts"]
            extra = dict(reference_1_worker=b2 and b2["steps_per_s"], reference_env_only_hand_loop=b3 and b3["steps_per_s"])
            pv, psample, _ = cpu_port_bounded("qq", horizon, threads, target_s=5.0)
            extra["cpu_baseline_port"] = dict(value=pv, unit="env-steps/s", cores=threads, kind="port", sample=psample)
            line = dict(impl="reference", metric="env-steps/sec", value=value, unit="env-steps/s", n_gpus=args.gpus,
                        steps=args.steps, warmup=1, ms_per_step=1e3 * b1["seconds"] / max(1, b1["repeat"]),
                        higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f64", data="synthetic", config=cfg,
                        cpu_baseline=dict(value=value, unit="env-steps/s", cores=threads, kind="reference", sample=sample),
                        e2e=dict(value=value, unit="env-steps/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                        gpu_launches=0, **extra)
            print(json.dumps(line), flush=True)
            return
    # fallback: the C restatement of the reference (oracle/), all host threads
    _, _, n = cpu_port_bounded(args.workload, horizon, threads, target_s=6.0)
    per = []
    for i in range(args.warmup + args.steps):
        steps, dt = cpu_port_run(args.workload, n, horizon, threads, seed=i)
        if i >= args.warmup:
            per.append((steps, dt))
    tot_steps, tot_t = sum(s for s, _ in per), sum(t for _, t in per)
    value = tot_steps / tot_t
    cfg["reference_sample_envs_per_step"] = n
    sample = f"{n} envs x {horizon} steps per bench step, {threads} OpenMP threads (baseline/_ref absent: C port)"
    line = dict(impl="reference", metric="env-steps/sec", value=value, unit="env-steps/s", n_gpus=args.gpus, steps=args.steps,
                warmup=args.warmup, ms_per_step=1e3 * tot_t / max(1, args.steps), higher_is_better=True, scaling="weak",
                vs_baseline=None, dtype="f64", data="synthetic", config=cfg,
                cpu_baseline=dict(value=value, unit="env-steps/s", cores=threads, kind="port", sample=sample),
                e2e=dict(value=value, unit="env-steps/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------------
# measurement helpers
# ---------------------------------------------------------------------------------------------------------------------
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


def ncu_traffic(key, units_per_launch=None):
    """dram bytes read + written per launch of a kernel from the committed ncu captures (profiles/r02_traffic.json, written
    by tools/ncu_traffic.py from `ncu --set full` reports), or None: nothing is pasted into this file.  A capture taken at
    another launch size is scaled by the work units (env-steps / rows) of the launch timed here; the source says so."""
    path = osp.join(ROOT, "profiles", "r02_traffic.json")
    if not osp.exists(path):
        return None, None
    d = json.load(open(path)).get(key)
    if not d:
        return None, None
    if units_per_launch is None or abs(units_per_launch - d["units"]) < 1e-6 * d["units"]:
        return d["dram_bytes"], d["source"]
    return (d["bytes_per_unit"] * units_per_launch,
            d["source"] + f"; {d['bytes_per_unit']:.4g} B per unit scaled to the {units_per_launch:.4g} units of this launch")


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


# ---------------------------------------------------------------------------------------------------------------------
# workloads (BASELINE.json configs) on the C-ABI
# ---------------------------------------------------------------------------------------------------------------------
def make_workload(name, E, L, torch, dev, hidden, prec):
    """Static part of a rollout workload: config, distribution table, network sizes, horizon, algorithmic work."""
    from simurlacra_b200.domain_randomization import DomainRandomizer, UniformDomainParam
    from simurlacra_b200.domain_randomization.default_randomizers import create_default_randomizer_qbb
    from simurlacra_b200.environments.pysim.quanser_ball_balancer import QBallBalancerSim
    from simurlacra_b200.environments.pysim.quanser_cartpole import QCartPoleSwingUpSim

    if name == "qq":  # C2
        H, sizes = 600, [6, hidden, hidden, 1]
        cfg = E.make_cfg(L.QQ, 1 / 250.0, H, dtype=torch.float32, act_norm=True)
        distr = E.make_distr_table(qq_randomizer(E.nominal_params(L.QQ)))
        label = f"QQubeSwingUpSim dt=1/250 H={H} + ActNorm + DomainRandWrapperLive(default qq) + fused FNN[{hidden},{hidden}]"
        record, wseed = False, 0
    elif name == "qcp":  # C3
        H, sizes = 600, [5, hidden, hidden, 1]
        cfg = E.make_cfg(L.QCP, 1 / 250.0, H, dtype=torch.float32, long_pole=True, act_norm=True, wild_init=0)
        nom = QCartPoleSwingUpSim.get_nominal_domain_param(long=True)
        dr = DomainRandomizer(*[UniformDomainParam(name=k, mean=nom[k], halfspan=nom[k] / 5, clip_lo=1e-4) for k in
                                ("cart_mass", "pole_mass", "pole_length", "pole_damping", "combined_damping",
                                 "cart_friction_coeff")])
        distr = dr.table(QCartPoleSwingUpSim._param_names)
        label = (f"QCartPoleSwingUpSim long pole dt=1/250 H={H} RK4 + ActNorm + uniform DR (6 params, +-20%) + fused "
                 f"FNN[{hidden},{hidden}]")
        record, wseed = False, 0
    elif name == "qbb":  # C4
        hidden = 16 if hidden == 64 else hidden
        H, sizes = 500, [8, hidden, hidden, 2]
        cfg = E.make_cfg(L.QBB, 0.01, H, dtype=torch.float32, act_norm=True)
        distr = create_default_randomizer_qbb().table(QBallBalancerSim._param_names)
        label = (f"QBallBalancerSim dt=0.01 H={H} + ActNorm + DomainRandWrapperLive(default qbb, 20 params) + fused "
                 f"FNN[{hidden},{hidden}], full StepSequence records [T,N,23] f32 + row flags written to HBM")
        record, wseed = True, 1
    else:
        raise ValueError(name)
    flop = ENV_FLOP[name] + fnn_flop(sizes)
    tc = hidden == 64 and prec > 0
    return dict(name=name, cfg=cfg, distr=distr, sizes=sizes, H=H, label=label, record=record, flop=flop, wseed=wseed,
                prec=prec if hidden == 64 else 0, tensor=tc, kind=cfg.kind)


class RolloutRunner:
    """Device-resident launches of one workload (weights, table and record buffers stay in HBM)."""

    def __init__(self, wl, E, L, torch, dev, n_envs, rank=0, wseed=None, prec=None, act_noise=False):
        self.wl, self.E, self.L, self.torch, self.dev, self.n, self.rank = wl, E, L, torch, dev, n_envs, rank
        w = fnn_weights(wl["sizes"], seed=wl["wseed"] if wseed is None else wseed)
        self.w = torch.from_numpy(w).to(dev)
        kw = dict(act_noise=L.ACT_NOISE_NORMAL, act_noise_scale=[0.1] * wl["sizes"][-1]) if act_noise else {}
        self.pol = E.make_policy(L.POL_FNN, self.w, sizes=wl["sizes"], precision=wl["prec"] if prec is None else prec,
                                 weights_host=torch.from_numpy(w), **kw)  # the host copy: small layers as launch arguments
        self.extra = {}
        if wl["record"]:
            lay = E.record_layout(wl["kind"])
            self.extra = dict(traj=torch.empty(wl["H"] + 1, n_envs, lay["width"], dtype=torch.float32, device=dev),
                              row_flag=torch.empty(wl["H"] + 1, n_envs, dtype=torch.uint8, device=dev))
        self.launches = 0

    def step(self, i):
        wl = self.wl
        res = self.E.rollout(wl["cfg"], self.pol, self.n, wl["H"] + 1, device=self.dev, distr=wl["distr"], seed=i,
                             env_offset=self.rank * self.n, auto_reset=True, record=wl["record"], **self.extra)
        self.launches += self.L.lib().sre_last_launch_count()
        return res["stats"]


def rate_of(stats_list, t, L):
    steps = sum(float(s[L.STAT_STEPS]) for s in stats_list)
    eps = sum(float(s[L.STAT_EPISODES]) for s in stats_list)
    return steps / t, eps / max(1, len(stats_list))


def roofline_for(wl, rate_per_gpu, peaks, fma_peak, n_envs, steps_per_launch=None):
    """The roofline object of a rollout workload at `rate_per_gpu` env-steps/s (steps_per_launch: env-steps one launch
    processed, for the ncu traffic figure)."""
    if wl["record"]:
        b = RECORD_BYTES[wl["name"]]
        gbs = rate_per_gpu * b / 1e9
        tr, src = ncu_traffic(f"{wl['name']}_rec", steps_per_launch)
        return dict(bound="hbm", achieved=gbs, peak=peaks["hbm_gbs"], unit="GB/s", frac=gbs / peaks["hbm_gbs"],
                    traffic=tr, traffic_source=src, bytes_per_env_step=b,
                    peak_source=f"MEASURED_PEAKS.json hbm_gbs ({peaks['source']})",
                    note="algorithmic bytes: one fp32 record (obs, act, act_applied, rew, state, hidden) + one flag byte per "
                         "env-step; the kernel is instruction-issue bound, not HBM bound (DESIGN.md section 5)")
    achieved = rate_per_gpu * wl["flop"] / 1e12
    fp32 = dict(bound="fp32_fma", achieved=achieved, peak=fma_peak, unit="TFLOP/s", frac=achieved / fma_peak, traffic=None,
                flop_per_env_step=wl["flop"],
                peak_source="measured live: FMA-only kernel (sre_bench_fma), best of 4; MEASURED_PEAKS.json has no FP32 figure",
                note="algorithmic FLOP (SURVEY 8d) of env + FNN per env-step over the FP32 CUDA-core FMA peak; with the "
                     "hidden layer on the tensor cores this fraction can exceed 1")
    if not wl["tensor"]:
        return fp32
    hid = wl["sizes"][1]
    passes = 3 if wl["prec"] == 1 else 1
    t_peak = peaks["bf16_tflops"] / 2
    t_iss = rate_per_gpu * passes * 2 * hid * hid / 1e12
    tr, src = ncu_traffic(f"{wl['name']}_tc", steps_per_launch)
    return dict(bound="tensor", achieved=achieved, peak=t_peak, unit="TFLOP/s", frac=achieved / t_peak, traffic=tr,
                traffic_source=src, flop_per_env_step=wl["flop"], issued=t_iss, frac_issued=t_iss / t_peak,
                peak_source=f"MEASURED_PEAKS.json bf16_tflops_sustained ({peaks['source']}) / 2: kind::tf32 runs at half "
                            f"the bf16 rate",
                note=f"{passes} tcgen05.mma kind::tf32 pass(es) per 64x64 layer; only 8192 of {wl['flop']} FLOP per env-step "
                     f"are GEMM-shaped, the rest (layer 1, 128 tanh, layer 3, dynamics) runs on the CUDA cores and the MUFU "
                     f"pipe, which bound the kernel (ncu summaries under profiles/): see roofline_fp32",
                fp32_view=fp32)


def build_sampler(name, torch, dev, hidden, prec, horizon):
    """The workload through the reference-facing classes (env chain + policy on the host + ParallelRolloutSampler)."""
    from simurlacra_b200.domain_randomization import DomainRandomizer, UniformDomainParam
    from simurlacra_b200.domain_randomization.default_randomizers import (create_default_randomizer_qbb,
                                                                           create_default_randomizer_qq)
    from simurlacra_b200.environment_wrappers import ActNormWrapper, DomainRandWrapperLive
    from simurlacra_b200.environments.pysim.quanser_ball_balancer import QBallBalancerSim
    from simurlacra_b200.environments.pysim.quanser_cartpole import QCartPoleSwingUpSim
    from simurlacra_b200.environments.pysim.quanser_qube import QQubeSwingUpSim
    from simurlacra_b200.policies import FNNPolicy
    from simurlacra_b200.sampling import ParallelRolloutSampler

    if name == "qq":
        sim, rz, n_obs, n_act, wseed = QQubeSwingUpSim(dt=1 / 250.0, max_steps=horizon, device=dev, dtype=torch.float32), \
            create_default_randomizer_qq(), 6, 1, 0
    elif name == "qcp":
        sim = QCartPoleSwingUpSim(dt=1 / 250.0, max_steps=horizon, long=True, device=dev, dtype=torch.float32)
        nom = QCartPoleSwingUpSim.get_nominal_domain_param(long=True)
        rz = DomainRandomizer(*[UniformDomainParam(name=k, mean=nom[k], halfspan=nom[k] / 5, clip_lo=1e-4) for k in
                                ("cart_mass", "pole_mass", "pole_length", "pole_damping", "combined_damping",
                                 "cart_friction_coeff")])
        n_obs, n_act, wseed = 5, 1, 0
    else:
        sim, rz, n_obs, n_act, wseed = QBallBalancerSim(dt=0.01, max_steps=horizon, device=dev, dtype=torch.float32), \
            create_default_randomizer_qbb(), 8, 2, 1
        hidden = 16 if hidden == 64 else hidden
    env = DomainRandWrapperLive(ActNormWrapper(sim), rz)
    policy = FNNPolicy(env.spec, hidden_sizes=[hidden, hidden], hidden_nonlin=torch.tanh)  # parameters live on the host
    policy.param_values = torch.from_numpy(fnn_weights([n_obs, hidden, hidden, n_act], seed=wseed))
    policy.engine_precision = prec if hidden == 64 else 0
    return env, policy, ParallelRolloutSampler(env, policy, num_workers=1, min_rollouts=1, seed=0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="qq", choices=["qq", "qcp", "qbb", "diff"])
    ap.add_argument("--envs", type=int, default=0, help="environments per GPU (default: the BASELINE.json size of the workload)")
    ap.add_argument("--hidden", type=int, default=64)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="only the headline slot (value, e2e, roofline)")
    ap.add_argument("--precision", type=int, default=1,
                    help="hidden layer of the fused policy: 0 fp32 FMA (CUDA cores), 1 tcgen05 3xTF32 (fp32-level "
                         "accuracy, default), 2 tcgen05 single TF32")
    args = ap.parse_args()
    default_envs = dict(qq=1 << 20, qcp=1 << 22, qbb=1 << 21, diff=1 << 18)
    if args.envs <= 0:
        args.envs = default_envs[args.workload]
    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch
    import torch.distributed as dist

    import simurlacra_b200.engine as E
    from simurlacra_b200 import _lib as L

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert args.warmup >= 3, "timing rules: at least 3 warm-up steps"
    N, K, hid, prec = args.envs, args.steps, args.hidden, args.precision
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        t, outs = time_launches(torch, fn, k, flush)
        barrier()
        if world > 1:
            tt = torch.tensor([t], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t = float(tt)
        return t, outs

    def wall_timed(fn, k):
        barrier()
        t0 = time.perf_counter()
        outs = [fn(i) for i in range(k)]
        barrier()
        t = time.perf_counter() - t0
        if world > 1:
            tt = torch.tensor([t], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t = float(tt)
        return t, outs

    peaks = measured_peaks()
    fma_peak = measure_fma_peak(E, L, torch, dev, L.F32)
    FMA_REG.update(scalar=measure_fma_peak(E, L, torch, dev, 3), packed=measure_fma_peak(E, L, torch, dev, 4))
    extras = {}
    clocks = ClockSampler(local)

    # =================================================================================================== headline slot
    if args.workload == "diff":
        head = run_diff(E, L, torch, dev, N, 300, K, args.warmup, timed, wall_timed, peaks, fma_peak, clocks, rank, world, dist)
    else:
        wl = make_workload(args.workload, E, L, torch, dev, hid, prec)
        run = RolloutRunner(wl, E, L, torch, dev, N, rank=rank)

        def resident(i):
            st = run.step(i)
            if world > 1:  # the statistics the learner logs: ONE small NCCL all-reduce of the sums
                dist.all_reduce(st[:6])
            return st

        for i in range(args.warmup):
            resident(i)
        run.launches = 0
        if rank == 0:
            clocks.start()
        t_dev, stats = timed(lambda i: resident(args.warmup + i), K)
        clk = clocks.stop() if rank == 0 else None
        value, episodes = rate_of(stats, t_dev, L)  # statistics are all-reduced: whole job
        # ---- end to end through the public sampler call with host buffers
        env, policy, sampler = build_sampler(args.workload, torch, dev, hid, prec, wl["H"])

        def e2e_step(i):
            res = sampler.sample_batch(N, auto_reset=True, record=wl["record"], **run.extra)
            ret, ln, st = sampler.read_results(res)  # pinned staging buffers of the sampler, one synchronisation
            return float(ret[0]), int(ln[0]), torch.from_numpy(st.copy())

        for i in range(2):
            e2e_step(i)
        t_e2e, outs = wall_timed(e2e_step, K)
        e2e_value = sum(float(o[2][L.STAT_STEPS]) for o in outs) / t_e2e
        head = dict(value=value, t_dev=t_dev, clk=clk, launches=run.launches, episodes_per_launch=episodes,
                    e2e=dict(value=e2e_value, unit="env-steps/s", h2d_bytes_per_step=int(policy.num_param * 4),
                             d2h_bytes_per_step=int(N * 8 + 64),
                             call="ParallelRolloutSampler.sample_batch(N) + read_results(): policy parameters from pinned host memory, "
                                  "returns + lengths + statistics read back into pinned host buffers"),
                    roofline=roofline_for(wl, value / world, peaks, fma_peak, N, steps_per_launch=value / world * t_dev / K),
                    config=dict(workload=wl["label"], envs_per_gpu=N, horizon=wl["H"], auto_reset=True,
                                trajectory="full records in HBM" if wl["record"] else "returns+lengths+stats",
                                policy_hidden_layer={0: "fp32 FMA (CUDA cores)", 1: "tcgen05 3xTF32 split (fp32-level accuracy)",
                                                     2: "tcgen05 TF32"}[wl["prec"]],
                                policy_small_layers="launch arguments: the host copy of the parameters (sre_policy.weights_host) "
                                                    "puts layer 1 / biases / layer 3 into the constant bank",
                                l2="flushed between timed iterations (256 MB fill)", parallelism=f"env-shard x{world}"))
        if wl["tensor"]:
            extras["roofline_fp32"] = head["roofline"].pop("fp32_view")
        del run
        torch.cuda.empty_cache()

    line = None
    if rank == 0:
        line = dict(metric="env-steps/sec", value=head["value"], unit="env-steps/s", n_gpus=world, steps=K, warmup=args.warmup,
                    ms_per_step=1e3 * head["t_dev"] / K, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
                    data="synthetic", config=head["config"], e2e=head["e2e"], gpu_launches=head["launches"],
                    clocks=head["clk"], roofline=head["roofline"], episodes_per_launch=head.get("episodes_per_launch"))

    # ================================================================ the other BASELINE.json configs (one GPU, rank 0)
    if world == 1 and not args.no_extras:
        extras.update(run_extras(args, E, L, torch, dev, timed, wall_timed, peaks, fma_peak, flush))
    if rank == 0:
        line.update(extras)
        if not args.no_cpu_baseline and world == 1:
            threads = os.cpu_count() or 1
            pv, psample, _ = cpu_port_bounded("qq", 600, threads)
            port = dict(value=pv, unit="env-steps/s", cores=threads, kind="port", sample=psample)
            ref = reference_sampler_baseline(threads) if reference_available() else None
            line["cpu_baseline"] = ref if ref is not None else port
            line["cpu_baseline_port"] = port
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def diff_setup(E, L, torch, dev, n, h):
    dcfg = E.make_cfg(L.QCP, 1 / 20.0, h, dtype=torch.float32, long_pole=True, wild_init=1)
    g = torch.Generator(device="cpu").manual_seed(0)
    s0_h = (torch.rand(n, 4, generator=g) * 0.2 - 0.1).pin_memory()
    act_h = ((torch.rand(h, n, generator=g) * 2 - 1) * 5.0).pin_memory()
    return dcfg, s0_h, act_h


FMA_REG = {}  # measured live in main(): FMA rate with three distinct register operands (scalar / packed f32x2), TFLOP/s


def reg_operand_fields(achieved):
    """The FMA-only peak above is reached with operands from immediates / the reuse cache; simulation arithmetic reads three
    changing registers per FMA, which the register file serves at a lower rate (measured by sre_bench_fma dtype 3 / 4)."""
    if not FMA_REG:
        return {}
    best = max(FMA_REG.values())
    return dict(peak_register_operands=best, frac_register_operands=achieved / best,
                peak_register_operands_source="measured live: FFMA / FFMA2 with three distinct changing register operands "
                                              f"(sre_bench_fma dtype 3: {FMA_REG['scalar']:.1f}, dtype 4: {FMA_REG['packed']:.1f} TFLOP/s)")


def c5_rooflines(rate_fwd, rate_fb, peaks, fma_peak, n, h):
    """C5: forward = FP32 FMA (375 FLOP per step); forward + adjoint also moves the 24 B/step tape twice."""
    a_f = rate_fwd * C5_FLOP_FWD / 1e12
    a_fb = rate_fb * (C5_FLOP_FWD + C5_FLOP_BWD) / 1e12
    tape = rate_fb * 2 * C5_TAPE_BYTES / 1e9
    return dict(workload=f"differentiable QCartPole (sac_wgrad path), RK4 dt=1/20, {n} envs x {h} steps, injected actions",
                forward=dict(env_steps_per_s=rate_fwd, bound="fp32_fma", achieved=a_f, peak=fma_peak, unit="TFLOP/s",
                             frac=a_f / fma_peak, flop_per_env_step=C5_FLOP_FWD, **reg_operand_fields(a_f)),
                forward_adjoint=dict(env_steps_per_s=rate_fb, bound="fp32_fma", achieved=a_fb, peak=fma_peak, unit="TFLOP/s",
                                     frac=a_fb / fma_peak, flop_per_env_step=C5_FLOP_FWD + C5_FLOP_BWD,
                                     flop_note="forward 375 (SURVEY 8a) + adjoint 2.5 x 375: the reverse sweep re-evaluates "
                                               "the four RK4 stages from the tape and propagates 5 adjoints through each",
                                     tape_gbs=tape, tape_frac_hbm=tape / peaks["hbm_gbs"], **reg_operand_fields(a_fb)))


def run_diff(E, L, torch, dev, n, h, K, warmup, timed, wall_timed, peaks, fma_peak, clocks, rank, world, dist):
    """--workload diff: C5 in the headline slot; one step = forward + adjoint over the horizon."""
    dcfg, s0_h, act_h = diff_setup(E, L, torch, dev, n, h)
    s0, act = s0_h.to(dev), act_h.to(dev)
    g_rew = torch.ones(h, n, device=dev)

    def fwd(i):
        return E.qcp_diff_forward(dcfg, None, s0, act)

    def fwd_bwd(i):
        states, thdd, rew, done = E.qcp_diff_forward(dcfg, None, s0, act)
        return E.qcp_diff_backward(dcfg, None, states, thdd, act, g_rew=g_rew)

    for i in range(warmup):
        fwd_bwd(i)
    if rank == 0:
        clocks.start()
    t_fb, _ = timed(lambda i: fwd_bwd(i) and None, K)
    clk = clocks.stop() if rank == 0 else None
    t_f, _ = timed(lambda i: fwd(i) and None, K)
    rate_fb, rate_f = K * n * h * world / t_fb, K * n * h * world / t_f

    out_h = (torch.empty(n, 4).pin_memory(), torch.empty(h, n).pin_memory(), torch.empty(n).pin_memory())

    def e2e_step(i):  # pinned host buffers in, gradients and returns out into pinned host buffers
        s0d, actd = s0_h.to(dev, non_blocking=True), act_h.to(dev, non_blocking=True)
        states, thdd, rew, done = E.qcp_diff_forward(dcfg, None, s0d, actd)
        g_s0, g_act, _ = E.qcp_diff_backward(dcfg, None, states, thdd, actd, g_rew=g_rew)
        out_h[0].copy_(g_s0, non_blocking=True)
        out_h[1].copy_(g_act, non_blocking=True)
        out_h[2].copy_(rew.sum(0), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(out_h[2][0])

    for i in range(2):
        e2e_step(i)
    t_e, _ = wall_timed(e2e_step, K)
    roof = c5_rooflines(rate_f / world, rate_fb / world, peaks, fma_peak, n, h)
    r = dict(roof["forward_adjoint"], traffic=None, forward=roof["forward"])
    return dict(value=rate_fb, t_dev=t_fb, clk=clk, launches=2 * K,
                e2e=dict(value=K * n * h * world / t_e, unit="env-steps/s", h2d_bytes_per_step=int(n * 4 * 4 + h * n * 4),
                         d2h_bytes_per_step=int(n * 4 * 4 + h * n * 4 + n * 4),
                         call="qcp_diff_forward + qcp_diff_backward: initial states and actions from pinned host memory, "
                              "gradients and returns read back to the host"),
                roofline=r,
                config=dict(workload=roof["workload"] + "; one step = forward + adjoint", envs_per_gpu=n, horizon=h,
                            l2="flushed between timed iterations (256 MB fill)", parallelism=f"env-shard x{world}"))


def run_extras(args, E, L, torch, dev, timed, wall_timed, peaks, fma_peak, flush):
    """Everything BASELINE.json names besides the headline slot, on one GPU."""
    out = {}
    K = max(2, min(args.steps, 3))

    def measure(wl_name, n, hidden=64, prec=1, wseed=None, act_noise=False, k=K):
        wl = make_workload(wl_name, E, L, torch, dev, hidden, prec)
        run = RolloutRunner(wl, E, L, torch, dev, n, wseed=wseed, act_noise=act_noise)
        for i in range(3):
            run.step(i)
        t, st = timed(lambda i: run.step(50 + i), k)
        rate, eps = rate_of(st, t, L)
        del run
        torch.cuda.empty_cache()
        return wl, rate, eps, 1e3 * t / k

    # ---- variants of the headline configuration (C2)
    if args.workload == "qq" and args.hidden == 64:
        n = args.envs
        var = {}
        other = 0 if args.precision else 1
        wl, rate, eps, ms = measure("qq", n, prec=other)
        var["fp32_fma" if other == 0 else "tcgen05_tf32x3"] = dict(value=rate, unit="env-steps/s", episodes_per_launch=eps,
                                                                   fma_roofline_frac=rate * wl["flop"] / 1e12 / fma_peak)
        wl, rate, eps, ms = measure("qq", n, prec=args.precision, wseed=3)
        var["failure_heavy_policy"] = dict(value=rate, unit="env-steps/s", episodes_per_launch=eps,
                                           note="policy weights of tools/sweep.py (seed 3): ~1.5x the episode starts of the "
                                                "bench policy in the same number of env-steps")
        wl, rate, eps, ms = measure("qq", n, prec=args.precision, act_noise=True)
        var["normal_act_noise"] = dict(value=rate, unit="env-steps/s", note="headline arithmetic + NormalActNoiseExplStrat "
                                                                               "(std 0.1) fused into the loop")
        out["variants"] = var

    # ---- C1: one env, 600 steps, nominal parameters, FNN [64,64]: latency per step
    wl1 = make_workload("qq", E, L, torch, dev, 64, args.precision)
    w1 = fnn_weights(wl1["sizes"])
    w1[-65:] = 0.0  # zero output layer: the episode runs its 600 steps
    pol1 = E.make_policy(L.POL_FNN, torch.from_numpy(w1).to(dev), sizes=wl1["sizes"], precision=args.precision)

    def c1(i):
        return E.rollout(wl1["cfg"], pol1, 1, 601, device=dev, seed=i, record=True)["length"]

    for i in range(3):
        c1(i)
    t1, ln = timed(c1, 5)
    out["c1_latency"] = dict(us_per_env_step=1e6 * t1 / sum(int(x[0]) for x in ln), ms_per_rollout=1e3 * t1 / 5,
                             workload="C1: QQubeSwingUpSim, 1 env x 600 steps, nominal parameters, fused FNN[64,64], full "
                                      "records; one kernel launch, device-timed", steps_per_rollout=int(ln[0][0]))

    # ---- C3 / C4 at the sizes BASELINE.json names
    if args.workload != "qcp":
        c3 = {}
        for prec, nm in ((1, "tcgen05_tf32x3"), (0, "fp32_fma")):
            wl, rate, eps, ms = measure("qcp", 1 << 22, prec=prec, k=2)
            c3[nm] = dict(roofline_for(wl, rate, peaks, fma_peak, 1 << 22, steps_per_launch=rate * ms * 1e-3), env_steps_per_s=rate, ms_per_launch=ms,
                          episodes_per_launch=eps)
            c3[nm].pop("fp32_view", None)
        c3["workload"] = wl["label"] + ", 4194304 envs"
        out["roofline_c3"] = c3
    if args.workload != "qbb":
        wl, rate, eps, ms = measure("qbb", 1 << 21, k=2)
        out["roofline_traj"] = dict(roofline_for(wl, rate, peaks, fma_peak, 1 << 21, steps_per_launch=rate * ms * 1e-3), env_steps_per_s=rate, ms_per_launch=ms,
                                    episodes_per_launch=eps, workload=wl["label"] + ", 2097152 envs")

    # ---- C5: differentiable cart-pole
    if args.workload != "diff":
        n5, h5 = 1 << 18, 300
        dcfg, s0_h, act_h = diff_setup(E, L, torch, dev, n5, h5)
        s0, act = s0_h.to(dev), act_h.to(dev)
        g_rew = torch.ones(h5, n5, device=dev)

        def fwd(i):
            return E.qcp_diff_forward(dcfg, None, s0, act) and None

        def fwd_bwd(i):
            states, thdd, rew, done = E.qcp_diff_forward(dcfg, None, s0, act)
            E.qcp_diff_backward(dcfg, None, states, thdd, act, g_rew=g_rew)

        for i in range(3):
            fwd_bwd(i)
        t_f, _ = timed(fwd, 5)
        t_fb, _ = timed(fwd_bwd, 5)
        out["roofline_c5"] = c5_rooflines(5 * n5 * h5 / t_f, 5 * n5 * h5 / t_fb, peaks, fma_peak, n5, h5)
        del s0, act, g_rew

    # ---- post-processing of a full record set: returns-to-go + GAE, purely HBM-bound
    rT, qn = 601, 1 << 20
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
    tr, src = ncu_traffic("returns_601x1M")
    out["roofline_returns"] = dict(bound="hbm", achieved=rb, peak=peaks["hbm_gbs"], unit="GB/s", frac=rb / peaks["hbm_gbs"],
                                   traffic=tr, traffic_source=src,
                                   peak_source=f"MEASURED_PEAKS.json hbm_gbs ({peaks['source']})",
                                   workload=f"sre_returns (discounted returns + GAE), [{rT}, {qn}] f32 rewards / values / flags "
                                            f"resident, 17 B per row and lane", rows_per_s=3 * rT * qn / t_r)
    del r_rew, r_val, r_flag, o_ret, o_adv
    torch.cuda.empty_cache()

    # ---- the reference-signature call: sample() -> List[StepSequence]
    if args.workload == "qq":
        from simurlacra_b200.sampling import ParallelRolloutSampler

        env, policy, _ = build_sampler("qq", torch, dev, args.hidden, args.precision, 600)
        es = {}
        for tag, kw, fmt in (("min_steps_18000_numpy", dict(min_steps=30 * 600), "numpy"),
                             ("rollouts_65536_torch", dict(min_rollouts=1 << 16), "torch")):
            smp = ParallelRolloutSampler(env, policy, num_workers=1, seed=0, data_format=fmt, **kw)
            smp.sample()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            n_steps = n_ro = 0
            reps = 3 if tag.startswith("min") else 2
            for i in range(reps):
                ros = smp.sample()
                n_ro += len(ros)
                n_steps += sum(ro.length for ro in ros)
                del ros
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            es[tag] = dict(value=n_steps / dt, unit="env-steps/s", rollouts_per_call=n_ro // reps, s_per_call=dt / reps)
        es["call"] = ("ParallelRolloutSampler.sample() -> List[StepSequence] (reference signature, "
                      "parallel_rollout_sampler.py:248): min_steps = 30 x 600 as in scripts/training/qq-su_udr-ppo.py with numpy "
                      "StepSequences on the host; 65 536 rollouts as torch StepSequences that are views of the record buffer")
        out["e2e_sample"] = es
    return out


if __name__ == "__main__":
    main()
