#!/bin/bash
# tools/prof_final.sh -- the round's closing captures (each after the same command exited 0 without the profiler):
#   headline kernel (C2), C4 records, C3, C5 forward / adjoint, returns kernel; launch list of the default bench.
# The reports are summarised ON the GPU box (gpurun brings back at most 64 MiB): text summaries, opcode histograms and
# the dram-traffic json land in gpurun_out/prof/, the .ncu-rep files are deleted (SRE_KEEP_REP=<tag> keeps one).
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out/prof
cp profiles/r02_traffic.json gpurun_out/prof/r02_traffic.json 2>/dev/null
export SRE_TRAFFIC_JSON=gpurun_out/prof/r02_traffic.json
NCU="ncu --set full --clock-control none --import-source on -f"
steps_of() { python -c "import json,sys; print(json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])['steps'])" "$1"; }
summarise() {  # summarise <tag> <units> <traffic key or -> <what>
    local rep=gpurun_out/$1.ncu-rep
    [ -f "$rep" ] || { echo "$1: no report"; return; }
    python tools/ncu_summary.py --rep "$rep" --out "gpurun_out/prof/$1.txt" --note "$4" > /dev/null 2>&1
    python tools/ncu_opcodes.py "$rep" "$2" > "gpurun_out/prof/$1_opcodes.txt" 2>&1
    [ "$3" != "-" ] && python tools/ncu_traffic.py "$3" "$rep" "$2" "$4" > /dev/null
    [ "${SRE_KEEP_REP:-}" == "$1" ] || rm -f "$rep"
}
quick_cap() {  # quick_cap <tag> <traffic key> <what> <quick.py args...>
    local tag=$1 key=$2 what=$3; shift 3
    bash tools/prof.sh "$tag" "$@" > "gpurun_out/prof/${tag}_run.log" 2>&1
    local line; line=$(tail -1 "gpurun_out/${tag}_plain.log" | cut -c1-330)
    summarise "$tag" "$(steps_of gpurun_out/${tag}_plain.log)" "$key" "$what; without the profiler: $line"
    echo "$tag: $line"
}
quick_cap r02_h_headline qq_tc "rollout_kernel<QQ,float,TCW,PLAIN>, 2^20 envs x 600 (C2, headline; policy weights as launch arguments)" --workload qq --prec 1 --host-weights
quick_cap r02_h_c4 qbb_rec "rollout_kernel<QBB,float,FNN16W,PLAIN_REC>, 2^20 envs x 500, full records, bench policy (C4; policy weights as launch arguments)" --workload qbb --record --envs 1048576 --wseed 1 --host-weights
quick_cap r02_h_c3 qcp_tc "rollout_kernel<QCP,float,TCW,PLAIN>, 2^20 envs x 600 (C3; policy weights as launch arguments)" --workload qcp --prec 1 --host-weights
python bench.py --workload diff --no-cpu-baseline --no-extras --steps 2 > gpurun_out/prof/r02_h_c5_plain.log 2>&1 || { echo "diff bench failed"; exit 1; }
$NCU -k regex:qcp_diff_forward2_kernel --launch-skip 3 -c 1 -o gpurun_out/r02_h_c5_forward python bench.py --workload diff --no-cpu-baseline --no-extras --steps 2 > /dev/null 2>&1; echo "c5 fwd ncu $?"
summarise r02_h_c5_forward 78643200 c5_forward "qcp_diff_forward2_kernel (two environments per thread), 2^18 envs x 300 (C5 forward)"
$NCU -k regex:qcp_diff_backward_kernel --launch-skip 3 -c 1 -o gpurun_out/r02_h_c5_backward python bench.py --workload diff --no-cpu-baseline --no-extras --steps 2 > /dev/null 2>&1; echo "c5 bwd ncu $?"
summarise r02_h_c5_backward 78643200 c5_backward "qcp_diff_backward_kernel<float>, 2^18 envs x 300 (C5 adjoint)"
$NCU -k regex:returns_kernel --launch-skip 2 -c 1 -o gpurun_out/r02_h_returns python bench.py --no-cpu-baseline --steps 2 > /dev/null 2>&1; echo "returns ncu $?"
summarise r02_h_returns 630194176 returns_601x1M "returns_kernel<float,true,4>, [601, 2^20] rows (returns-to-go + GAE)"
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/prof/r02_h_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1; echo "launch list $?"
python tools/ncu_summary.py --launches gpurun_out/prof/r02_h_launches_bench.csv --out gpurun_out/prof/r02_h_launches_bench.txt > /dev/null 2>&1
du -sh gpurun_out
