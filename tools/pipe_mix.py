"""tools/pipe_mix.py -- do the FMA pipe (packed FFMA2, three register operands) and the XU pipe (MUFU.EX2) run
concurrently?  Times sre_bench_fma modes 5 (16 FFMA2 / iteration), 6 (8 MUFU / iteration), 7 (both) at the occupancy of
the headline kernel (512 threads per SM) and at full occupancy."""
import ctypes as C
import json
import os.path as osp
import sys

import torch

sys.path.insert(0, osp.dirname(osp.dirname(osp.abspath(__file__))))
from simurlacra_b200 import _lib as L  # noqa: E402

dev = torch.device("cuda", 0)
sink = torch.zeros(4, device=dev)
out = {}
for name, blocks, threads in (("4_warps_per_scheduler", 148 * 4, 128), ("16_warps_per_scheduler", 148 * 8, 256)):
    row = {}
    for mode, tag in ((5, "ffma2"), (6, "mufu"), (7, "both"), (8, "ffma2_invariant_reg"), (9, "ffma2_uniform"), (10, "fmul2"), (11, "fadd2_imm"), (4, "ffma2_3reg_16chains")):
        best = 1e9
        for it in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            fl = C.c_double()
            e0.record()
            L.check(L.lib().sre_bench_fma(mode, 20000, blocks, threads, C.c_void_p(sink.data_ptr()), C.byref(fl),
                                          C.c_void_p(torch.cuda.current_stream().cuda_stream)), "bench")
            e1.record()
            torch.cuda.synchronize()
            if it >= 1:
                best = min(best, e0.elapsed_time(e1))
        row[tag + "_ms"] = best
    # cycles per warp and iteration on one scheduler
    warps_per_smsp = blocks * threads / 32 / (148 * 4)
    for tag in [k[:-3] for k in list(row) if k.endswith("_ms")]:
        row[tag + "_cycles_per_warp_iter"] = row[tag + "_ms"] * 1e-3 * 1.965e9 / 20000 / warps_per_smsp
    row["sum_over_both"] = (row["ffma2_ms"] + row["mufu_ms"]) / row["both_ms"]
    row["max_over_both"] = max(row["ffma2_ms"], row["mufu_ms"]) / row["both_ms"]
    out[name] = row
print(json.dumps(out, indent=1))
