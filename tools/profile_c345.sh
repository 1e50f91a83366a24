#!/bin/bash
# tools/profile_c345.sh -- ncu captures of the dominant kernel of BASELINE configs C3 (cart-pole RK4 + uniform DR),
# C4 (ball balancer, StepSequence records to HBM) and C5 (differentiable cart-pole forward + adjoint).
# Each capture runs only after the same command exited 0 without the profiler (B200_PROFILING.md).
# usage (on the GPU box): bash tools/profile_c345.sh   -> gpurun_out/r01_l_*.ncu-rep + sweep json
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on -f"
cap() {  # cap <tag> <sweep prefix> <kernel regex> <launches to skip>
    python tools/sweep.py --quick --only "$2" --out "gpurun_out/sweep_$1.json" > "gpurun_out/sweep_$1.log" 2>&1 || { echo "$1: plain run failed"; return 1; }
    $NCU -k "regex:$3" --launch-skip "$4" -c 1 -o "gpurun_out/r01_l_$1" \
        python tools/sweep.py --quick --only "$2" --out /dev/null > "gpurun_out/ncu_$1.log" 2>&1
    echo "$1: ncu exit $?"
    python tools/ncu_summary.py --rep "gpurun_out/r01_l_$1.ncu-rep" --out "gpurun_out/r01_l_$1.txt" \
        --note "ncu --set full --clock-control none, one launch of $3 ($2), tools/sweep.py --quick" > /dev/null 2>&1
}
cap c3_qcp_rk4_tc "C3" "rollout_kernel" 2
cap c3_qcp_rk4_fma "C3" "rollout_kernel" 6
cap c4_qbb_records "C4 qbb+DR+FNN[16,16] records" "rollout_kernel" 2
cap c5_diff_forward "C5 qcp diff forward" "qcp_diff_forward_kernel" 2
cap c5_diff_backward "C5 qcp diff forward+adjoint" "qcp_diff_backward_kernel" 2
ls -la gpurun_out/*.ncu-rep
