"""Sequential SASS listing of an ncu capture with per-instruction dynamic counts and stall samples, to attribute the
warp time to code regions.  usage: python tools/ncu_regions.py rep env_steps [--min-exec X]"""
import csv, io, subprocess, sys
rep, env_steps = sys.argv[1], float(sys.argv[2])
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
h = rows[1]
si, ti, wi = h.index("Source"), h.index("Thread Instructions Executed"), h.index("Warp Stall Sampling (All Samples)")
cols = {k: h.index(k) for k in ("stall_long_sb", "stall_short_sb", "stall_wait", "stall_math", "stall_mio", "stall_no_inst", "stall_not_selected", "stall_branch_resolving", "stall_barrier", "stall_selected", "stall_dispatch", "stall_lg")}
tot_st = sum(float(r[wi]) for r in rows[2:] if len(r) > wi)
cum = 0.0
for i, r in enumerate(rows[2:]):
    if len(r) <= wi: continue
    ex = float(r[ti]) / env_steps
    st = 100 * float(r[wi]) / tot_st
    cum += st
    top = sorted(((float(r[c]), k) for k, c in cols.items()), reverse=True)[:2]
    why = " ".join(f"{k[6:]}:{int(v)}" for v, k in top if v > 0)
    print(f"{i:5d} {ex:7.3f} {st:6.2f} {cum:6.1f}  {r[si].strip()[:90]:90s} {why}")
