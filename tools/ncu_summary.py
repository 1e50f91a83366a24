"""Summarise an ncu launch list (csv) and/or a full .ncu-rep capture into a small text file under profiles/.
usage: python tools/ncu_summary.py --launches gpurun_out/launches.csv --rep gpurun_out/prof.ncu-rep --out profiles/x.txt"""
import argparse
import csv
import io
import subprocess
from collections import defaultdict

KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__waves_per_multiprocessor", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "sm__inst_issued.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "derived__smsp__sass_thread_inst_executed_op_ffma_pred_on_x2", "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum",
        "sass__thread_inst_executed_true_per_opcode", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__average_warp_latency_per_inst_issued.ratio",
        "sm__cycles_elapsed.avg", "smsp__cycles_active.avg",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "launch__shared_mem_per_block_dynamic",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"]


def launches(path, out):
    rows = [r for r in csv.reader(l for l in open(path) if not l.startswith("=="))]
    h = rows[0]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = defaultdict(lambda: [0, 0.0])
    unit = "ns"
    for r in rows[1:]:
        if len(r) <= vi:
            continue
        agg[r[ki][:100]][0] += 1
        agg[r[ki][:100]][1] += float(r[vi].replace(",", ""))
        unit = r[ui]
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(unit, 1e-6)
    tot = sum(v[1] for v in agg.values())
    out.write(f"== launch list {path} (gpu__time_duration.sum; cold-cache, serialised: compare shares) ==\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:15]:
        out.write(f"{v[1] * scale:12.3f} ms {v[0]:6d} launches {100 * v[1] / tot:6.2f}%  {k}\n")


def rep(path, out, extra):
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    h, units = rows[0], rows[1]
    out.write(f"== full capture {path} ==\n")
    for v in rows[2:]:
        name = v[h.index("Kernel Name")] if "Kernel Name" in h else "?"
        out.write(f"-- kernel: {name[:120]}\n")
        for i, n in enumerate(h):
            if n in KEYS or any(e in n for e in extra):
                out.write(f"{n:85s} {units[i]:>16s} {v[i]}\n")
        for i, n in enumerate(h):
            if "issue_stalled" in n and n.endswith("_per_warp_active.pct"):
                try:
                    if float(v[i]) >= 1.0:
                        out.write(f"{n:85s} {units[i]:>16s} {v[i]}\n")
                except ValueError:
                    pass


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--launches")
    ap.add_argument("--rep")
    ap.add_argument("--out", required=True)
    ap.add_argument("--extra", nargs="*", default=[])
    ap.add_argument("--note", default="")
    a = ap.parse_args()
    with open(a.out, "w") as f:
        if a.note:
            f.write(a.note + "\n")
        if a.launches:
            launches(a.launches, f)
        if a.rep:
            rep(a.rep, f, a.extra)
    print(open(a.out).read())
