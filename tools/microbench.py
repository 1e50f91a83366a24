"""Measured denominators on the box: FFMA, packed FFMA2, DFMA rates and streaming HBM write bandwidth."""
import ctypes as C
import json
import sys, os.path as osp
sys.path.insert(0, osp.dirname(osp.dirname(osp.abspath(__file__))))
import torch
from simurlacra_b200 import _lib as L

dev = torch.device("cuda:0")
sink = torch.zeros(8, dtype=torch.float32, device=dev)
out = {}
for name, code in (("ffma", 0), ("ffma2", 2), ("dfma", 1)):
    best = 0
    for blocks in (148 * 4, 148 * 8):
        for it in range(5):
            fl = C.c_double()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            L.check(L.lib().sre_bench_fma(code, 4096, blocks, 256, C.c_void_p(sink.data_ptr()), C.byref(fl),
                                          C.c_void_p(torch.cuda.current_stream().cuda_stream)))
            e1.record(); torch.cuda.synchronize()
            if it >= 2: best = max(best, fl.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    out[name + "_tflops"] = best
buf = torch.empty(4 << 30, dtype=torch.uint8, device=dev)
best = 0
for it in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    L.check(L.lib().sre_bench_hbm_write(C.c_void_p(buf.data_ptr()), buf.numel(), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    e1.record(); torch.cuda.synchronize()
    if it >= 2: best = max(best, buf.numel() / (e0.elapsed_time(e1) * 1e-3) / 1e9)
out["hbm_write_gbs"] = best
print(json.dumps(out))
