#!/bin/bash
# One GPU call: driver-shaped and long bench lines, configurations 1, 3, 4, timelines, the ncu launch list of the bench command
# and ncu --set full captures of k_track (a tracker that owns the GPU; a tracker's share when twenty share it; VGA) and of the
# pyramid kernels.  Everything lands in gpurun_out/.
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_kitti_20.json 2> gpurun_out/r02_bench_kitti_20.err
python bench.py --steps 300 --warmup 5 --no-cpu-baseline > gpurun_out/r02_bench_kitti_300.json 2> gpurun_out/r02_bench_kitti_300.err
bash tools/run_configs.sh
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches_kitti.csv python bench.py --steps 4 --warmup 3 --quick > gpurun_out/ncu_launch.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_track -s 2 -c 1 -o gpurun_out/r02_k_track_single -f python tools/one_track.py kitti 4 > gpurun_out/ncu_s.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:k_track -s 2 -c 1 -o gpurun_out/r02_k_track_shared -f python tools/one_track.py kitti 4 512 14 > gpurun_out/ncu_h.log 2>&1
ncu --set full --clock-control none -k regex:k_track -s 2 -c 1 -o gpurun_out/r02_k_track_vga -f python tools/one_track.py vga 4 > gpurun_out/ncu_v.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:"k_icp_depth|k_populate|k_maps" -s 4 -c 4 -o gpurun_out/r02_pyramid -f python tools/one_track.py kitti 4 > gpurun_out/ncu_p.log 2>&1
ls -la gpurun_out/*.ncu-rep
