#!/bin/bash
# AddressSanitizer + UBSan run of the host-side C++ (csrc/host: shared-memory process group, IDX loader, crc32c)
# under the whole CPU test suite (SURVEY section 5 "race detection / sanitizers").  No GPU needed.
#   bash scripts/asan_host.sh            # last result: 48 passed, no sanitizer report
# /opt/gcc's wrapper cannot find libasan on this image, hence the system compiler.
set -euo pipefail
cd "$(dirname "$0")/.."
export CXX=/usr/bin/g++
CXXFLAGS="-fsanitize=address,undefined -fno-omit-frame-pointer -g" python -m mpi_tensorflow_b200.ops.build --host --force
trap 'unset CXXFLAGS; python -m mpi_tensorflow_b200.ops.build --host --force >/dev/null' EXIT     # put the normal build back
ASAN=$(/usr/bin/gcc -print-file-name=libasan.so)
UBSAN=$(/usr/bin/gcc -print-file-name=libubsan.so)
LD_PRELOAD="$ASAN $UBSAN" ASAN_OPTIONS=detect_leaks=0:halt_on_error=1 UBSAN_OPTIONS=halt_on_error=1:print_stacktrace=1 \
  python -m pytest tests -q -x -p no:cacheprovider -m "not gpu"
