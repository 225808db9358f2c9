#!/bin/bash
# A/B of several builds of the library on the same box: every densemonoslam_b200/lib/libdmsodom*.so in turn.
# usage: tools/ab_perf.sh [workload] [steps] cfg...
W=${1:-kitti}; S=${2:-160}; shift 2
export CUDA_DEVICE_MAX_CONNECTIONS=32
for lib in densemonoslam_b200/lib/libdmsodom*.so; do
  echo "== $lib"
  DMSODOM_LIB=$PWD/$lib timeout 600 python tools/quick_perf.py $W $S "$@" 2>&1 | grep frames
done
