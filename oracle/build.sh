#!/usr/bin/env bash
# Builds the C++ CPU restatement (test infrastructure / CPU baseline): oracle/liboracle_cpu.so.
# x86-64-v3 (AVX2 + FMA), not -march=native: the .so is built in the CPU container and travels to the GPU box.
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
g++ -O3 -march=x86-64-v3 -fopenmp -shared -fPIC -std=c++17 -Wall -fvisibility=hidden \
    -o "$here/liboracle_cpu.so" "$here/oracle_cpu.cpp"
echo "built $here/liboracle_cpu.so"
