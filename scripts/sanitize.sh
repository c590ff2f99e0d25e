#!/bin/bash
# Race / memory checking of the kernels (SURVEY section 5 "race detection / sanitizers").
# ONE tool per gpurun call, on the smallest case that exercises the kernel (B200_PROFILING.md):
#   gpurun --timeout 900 -- 'bash scripts/sanitize.sh memcheck  "conv2 and 3"'
#   gpurun --timeout 900 -- 'bash scripts/sanitize.sh racecheck "fc2"'
#   gpurun --timeout 900 -- 'bash scripts/sanitize.sh synccheck "maxpool"'
# The host-side C++ (csrc/host) has its own target: bash scripts/asan_host.sh (ASan + UBSan under the CPU test suite).
set -euo pipefail
tool=${1:-memcheck}
sel=${2:-"maxpool or im2col"}
mkdir -p gpurun_out
timeout 800 compute-sanitizer --tool "$tool" --error-exitcode 99 --log-file "gpurun_out/sanitizer_${tool}.log" \
  python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -p no:cacheprovider -k "$sel" 2>&1 | tail -5
tail -5 "gpurun_out/sanitizer_${tool}.log"
