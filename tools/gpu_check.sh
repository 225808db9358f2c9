#!/bin/bash
# One GPU call while iterating on k_track: parity suite, then the tuning helper in the regimes that matter.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
export CUDA_DEVICE_MAX_CONNECTIONS=32
timeout 600 python tools/quick_perf.py kitti 160 1x768 12x512x24 20x512x14 2>&1 | grep -v "^$" | tee gpurun_out/quick_perf.log
timeout 300 python tools/quick_perf.py vga 160 1x768 20x512x14 2>&1 | tee -a gpurun_out/quick_perf.log
