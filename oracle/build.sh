#!/usr/bin/env bash
# Builds the CPU oracle (test infrastructure) into oracle/_build/liboracle_odom.so.
# -ffp-contract=off is part of the arithmetic contract (see cpu_odom.cpp header).
set -euo pipefail
cd "$(dirname "$0")"
mkdir -p _build
g++ -std=c++17 -O2 -fPIC -shared -fopenmp -ffp-contract=off -fno-fast-math -mfma -mavx2 \
    -Wall -Wextra -o _build/liboracle_odom.so cpu_odom.cpp
echo "built oracle/_build/liboracle_odom.so"
