#!/usr/bin/env bash
# Builds the C++ drop-in-surface test against the in-tree library (rpath: relative, the tree travels to the GPU box).
set -euo pipefail
cd "$(dirname "$0")"
mkdir -p _build
g++ -std=c++14 -O1 -I../../include -I/usr/local/cuda/include shim_test.cpp -o _build/shim_test \
    -L../../densemonoslam_b200/lib -ldmsodom -L/usr/local/cuda/lib64 -lcudart \
    -Wl,-rpath,'$ORIGIN/../../../densemonoslam_b200/lib' -Wl,-rpath,/usr/local/cuda/lib64
g++ -std=c++14 -O1 -I../../include -I/usr/local/cuda/include frontend_test.cpp -o _build/frontend_test \
    -L../../densemonoslam_b200/lib -ldmsodom -L/usr/local/cuda/lib64 -lcudart \
    -Wl,-rpath,'$ORIGIN/../../../densemonoslam_b200/lib' -Wl,-rpath,/usr/local/cuda/lib64
echo "built tests/cpp/_build/shim_test tests/cpp/_build/frontend_test"
