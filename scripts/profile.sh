#!/bin/bash
# The profiling recipe used for profiles/ (B200_PROFILING.md), to be run under gpurun on ONE GPU:
#   gpurun --timeout 900 -- 'bash scripts/profile.sh launches'           # per-kernel device time of a step
#   gpurun --timeout 900 -- 'bash scripts/profile.sh full "conv2|conv1"'  # ncu --set full of matching kernels
# then, on the build box:  python scripts/launches.py gpurun_out/launches.csv conv1_fwd
#                          ncu -i gpurun_out/prof.ncu-rep --page raw --csv > raw.csv && python scripts/ncu_summary.py raw.csv
set -euo pipefail
mode=${1:-launches}
cmd="python bench.py --steps 3 --warmup 3 --no-e2e --no-graph"
mkdir -p gpurun_out
timeout 200 $cmd > gpurun_out/plain.log 2>&1          # ncu only after the same command exited 0 without it
if [ "$mode" = launches ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 400 --csv \
    --log-file gpurun_out/launches.csv $cmd > gpurun_out/ncu.log 2>&1
else
  timeout 700 ncu --set full --clock-control none --import-source on -k "regex:${2:-conv}" -s 12 -c 12 \
    -o gpurun_out/prof $cmd > gpurun_out/ncu_full.log 2>&1
fi
for f in gpurun_out/ncu*.log; do tail -n 2 "$f"; done
