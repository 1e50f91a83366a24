"""Dynamic opcode histogram from an ncu capture (SASS source page): thread-instructions per env-step by opcode.
usage: python tools/ncu_opcodes.py rep env_steps"""
import csv, io, subprocess, sys
from collections import defaultdict
rep, env_steps = sys.argv[1], float(sys.argv[2])
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
h = rows[1]
si, ti, wi = h.index("Source"), h.index("Thread Instructions Executed"), h.index("Warp Stall Sampling (All Samples)")
agg, stall = defaultdict(float), defaultdict(float)
tot = 0
for r in rows[2:]:
    if len(r) <= ti: continue
    toks = r[si].split()
    if not toks: continue
    op = toks[1] if toks[0].startswith("@") else toks[0]
    op = op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("MUFU", "LDS", "STS", "LDG", "STG")) and "." in op else "")
    agg[op] += float(r[ti]); stall[op] += float(r[wi]); tot += float(r[ti])
st = sum(stall.values())
print(f"total thread-inst per env-step: {tot / env_steps:.1f}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:40]:
    print(f"{k:14s} {v / env_steps:9.1f}  stall-samples {100 * stall[k] / st:5.1f}%")
