#!/usr/bin/env bash
# Builds the reference's own legacy tracking kernels, from the sources where they lie under /root/reference,
# into oracle/_ref/libref_legacy.so (git-ignored; it travels to the GPU box with the snapshot).
# No reference source is copied or modified: oracle/ref/ref_shim.h is force-included to supply the two
# things CUDA 12.9 / sm_100a no longer provide (__shfl_down without mask, legacy texture references).
# Flags are the reference's own (Core/src/CMakeLists.txt:48-49) with the arch changed to sm_100a.
set -euo pipefail
cd "$(dirname "$0")"
REF=${REF:-/root/reference/elasticfusion/Core/src/Cuda}
[ -d "$REF" ] || { echo "reference tree not found at $REF (the GPU box uses the prebuilt oracle/_ref)"; exit 0; }
mkdir -p _ref
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-O3 -std=c++14 -gencode arch=compute_100a,code=sm_100a --ftz=true --prec-div=false --prec-sqrt=false -Xcompiler -fPIC -w"
$NVCC $FLAGS -include "$PWD/ref/ref_shim.h" -I"$REF" -c "$REF/reduce.cu" -o _ref/reduce.o
$NVCC $FLAGS -include "$PWD/ref/ref_shim.h" -I"$REF" -c "$REF/cudafuncs.cu" -o _ref/cudafuncs.o
$NVCC $FLAGS -I"$REF" -x cu -c "$REF/containers/device_memory.cpp" -o _ref/device_memory.o
$NVCC $FLAGS -I"$REF" -c ref/ref_driver.cu -o _ref/ref_driver.o
$NVCC -shared -o _ref/libref_legacy.so _ref/reduce.o _ref/cudafuncs.o _ref/device_memory.o _ref/ref_driver.o
rm -f _ref/*.o
echo "built oracle/_ref/libref_legacy.so"
