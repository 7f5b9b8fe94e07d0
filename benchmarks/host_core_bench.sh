#!/bin/bash
# Build and run benchmarks/host_core_bench.cpp (needs baseline/_ref: bash baseline/build_ref.sh,
# or only the copy step of it when no CUDA toolkit is around).
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
cd "$ROOT"
REF="${REF:-/root/reference}"
if [ ! -f baseline/_ref/src/parameter_server.cpp ]; then
  mkdir -p baseline/_ref/src baseline/_ref/include
  cp "$REF/src/parameter_server.cpp" baseline/_ref/src/
  cp "$REF/include/parameter_server.h" baseline/_ref/include/
fi
mkdir -p build/bin
${CXX:-g++} -std=c++17 -O2 -pthread -I baseline/_ref/include benchmarks/host_core_bench.cpp \
  baseline/_ref/src/parameter_server.cpp csrc/ps/ps_core.cpp csrc/ps/ps_device.cpp -ldl \
  -o build/bin/host_core_bench
exec build/bin/host_core_bench "$@"
