#!/usr/bin/env bash
# Builds the product library densemonoslam_b200/lib/libdmsodom.so for sm_100a.
# -fmad=false: every fused multiply-add in the kernels is an explicit intrinsic (see common.cuh);
# -ffp-contract=off: same rule for the few host-side float expressions (Rprev^-1).
set -euo pipefail
cd "$(dirname "$0")"
mkdir -p ../lib
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
$NVCC -std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -fmad=false \
      -Xcompiler -fPIC,-ffp-contract=off -Xptxas -v \
      -I../../include -I. -shared -o ../lib/libdmsodom.so pyramid.cu track.cu odom_c.cu frontend.cu nid.cu splat.cu sequence.cu ring.cu "$@"
echo "built densemonoslam_b200/lib/libdmsodom.so"
