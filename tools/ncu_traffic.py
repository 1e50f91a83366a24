"""Extract dram__bytes_read.sum + dram__bytes_write.sum of one captured kernel launch from an `ncu --set full` report and
record it, with the number of work units (env-steps / rows) that launch processed, in profiles/r02_traffic.json.
bench.py scales the figure to the launch it times (traffic per unit x units per launch) and names the capture.

usage: python tools/ncu_traffic.py <key> <rep.ncu-rep> <units in the captured launch> "<what was captured>"
keys bench.py reads: qq_tc, qcp_tc, qbb_rec, returns_601x1M"""
import csv
import io
import json
import os.path as osp
import subprocess
import sys

ROOT = osp.dirname(osp.dirname(osp.abspath(__file__)))


def dram_bytes(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    h, units, vals = rows[0], rows[1], rows[2]
    tot = 0.0
    for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        i = h.index(name)
        mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[units[i]]
        tot += float(vals[i].replace(",", "")) * mult
    return tot


def main():
    key, rep, units, what = sys.argv[1], sys.argv[2], float(sys.argv[3]), sys.argv[4]
    import os

    path = os.environ.get("SRE_TRAFFIC_JSON") or osp.join(ROOT, "profiles", "r02_traffic.json")
    d = json.load(open(path)) if osp.exists(path) else {}
    b = dram_bytes(rep)
    d[key] = dict(dram_bytes=b, units=units, bytes_per_unit=b / units, source=f"ncu --set full, {osp.basename(rep)}: {what}")
    json.dump(d, open(path, "w"), indent=1, sort_keys=True)
    print(key, d[key])


if __name__ == "__main__":
    main()
