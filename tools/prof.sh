#!/bin/bash
# tools/prof.sh <tag> [quick.py args...] -- one `ncu --set full` capture of the rollout kernel of a tools/quick.py run
# (after the same command exited 0 without the profiler), report to gpurun_out/<tag>.ncu-rep (B200_PROFILING.md recipe).
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
tag=$1; shift
python tools/quick.py "$@" --reps 2 --tag "$tag" > "gpurun_out/${tag}_plain.log" 2>&1 || { echo "plain run failed"; tail -5 "gpurun_out/${tag}_plain.log"; exit 1; }
tail -1 "gpurun_out/${tag}_plain.log"
ncu --set full --clock-control none --import-source on -f -k "regex:rollout_kernel" --launch-skip 3 -c 1 -o "gpurun_out/${tag}" \
    python tools/quick.py "$@" --reps 1 > "gpurun_out/${tag}_ncu.log" 2>&1
echo "ncu exit $?"
ls -la "gpurun_out/${tag}.ncu-rep"
