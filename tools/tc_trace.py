"""tools/tc_trace.py -- debug: per-warp clock stamps of the tensor-core policy (library built with -DSRE_TC_TRACE, run with
SRE_B200_LIB=simurlacra_b200/lib/libsre_trace.so).  Prints where a policy step's wall time goes: compute before each
publish, skew between the warps of a tile, MMA wait after the last arrival."""
import os.path as osp
import sys

import numpy as np

ROOT = osp.dirname(osp.dirname(osp.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from bench import fnn_weights, qq_randomizer
from simurlacra_b200 import _lib as L
from simurlacra_b200 import engine as E

dev = torch.device("cuda", 0)
N, H = 1 << 20, 600
sizes = [6, 64, 64, 1]
w = torch.from_numpy(fnn_weights(sizes)).to(dev)
cfg = E.make_cfg(L.QQ, 1 / 250.0, H, dtype=torch.float32, act_norm=True)
pol = E.make_policy(L.POL_FNN, w, sizes=sizes, precision=int(sys.argv[1]) if len(sys.argv) > 1 else 1)
distr = E.make_distr_table(qq_randomizer(E.nominal_params(L.QQ)))
for i in range(2):
    res = E.rollout(cfg, pol, N, H + 1, device=dev, distr=distr, seed=i, auto_reset=True, record=False)
torch.cuda.synchronize()
tr = res["status"].cpu().numpy().astype(np.uint32)[: 592 * 8 * 4 * 8].reshape(592, 8, 4, 8).astype(np.int64)
smid = tr[:, 0, 0, 7]
d = lambda a, b: ((a - b) & 0xFFFFFFFF)
t0, a0, w0b, w0e, a1, w1b, w1e = (tr[..., k] for k in range(7))
step_len = d(t0[:, 1:], t0[:, :-1])  # [blk, 7, warp]
print("policy-entry to policy-entry (one env step of a warp), cycles: mean %.0f  p10 %.0f  p90 %.0f" % (
    step_len.mean(), np.percentile(step_len, 10), np.percentile(step_len, 90)))
print("entry -> arrive h0 (L1 + tanh + st of half 0):      mean %.0f" % d(a0, t0).mean())
print("arrive h0 -> begin wait h0 (L1 of half 1):           mean %.0f" % d(w0b, a0).mean())
print("wait h0 (mbarrier):                                  mean %.0f  p50 %.0f p90 %.0f" % (d(w0e, w0b).mean(), np.median(d(w0e, w0b)), np.percentile(d(w0e, w0b), 90)))
print("end wait h0 -> arrive h1 (tanh + st of half 1):      mean %.0f" % d(a1, w0e).mean())
print("wait h1 (mbarrier):                                  mean %.0f  p50 %.0f p90 %.0f" % (d(w1e, w1b).mean(), np.median(d(w1e, w1b)), np.percentile(d(w1e, w1b), 90)))
nxt = d(t0[:, 1:], w1e[:, :-1])
print("end wait h1 -> next policy entry (epilogue + env):   mean %.0f" % nxt.mean())
# skew: spread of the arrival stamps of the 4 warps of a tile; MMA latency: last arrival -> earliest wake-up
for nm, arr, wend in (("h0", a0, w0e), ("h1", a1, w1e)):
    last = arr.max(-1)
    first = arr.min(-1)
    print(f"{nm}: arrival spread over the 4 warps mean {d(last, first).mean():.0f} p90 {np.percentile(d(last, first), 90):.0f}; "
          f"last arrival -> first wake-up mean {d(wend.min(-1), last).mean():.0f} p50 {np.median(d(wend.min(-1), last)):.0f} p90 {np.percentile(d(wend.min(-1), last), 90):.0f}")
# phase relation of the 4 tiles of one SM: entry stamps of step 0 relative to the earliest tile of the SM
import collections
by_sm = collections.defaultdict(list)
for b in range(592):
    by_sm[int(smid[b])].append(b)
offs = []
for sm, bl in by_sm.items():
    if len(bl) == 4:
        e = np.array([a0[b, 3, 0] for b in bl])
        offs.append(np.sort(d(e, e.min())))
offs = np.array(offs)
print("h0 arrival offsets of the 4 tiles of an SM (sorted, cycles): mean", offs.mean(0).round(0), " (SMs with 4 traced tiles:", len(offs), ")")
np.save(osp.join(ROOT, "gpurun_out", "tc_trace.npy"), tr)
