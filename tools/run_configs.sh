#!/bin/bash
# SURVEY 8(d) configurations 1-4 on one GPU; results under gpurun_out/cfg_*.  (Config 5 = bench.py --gpus N under torchrun.)
mkdir -p gpurun_out
S=${STEPS:-300}
python bench.py --workload vga   --steps $S --warmup 5 --no-cpu-baseline > gpurun_out/cfg1_vga.json   2> gpurun_out/cfg1_vga.err
python tools/track_timing.py vga > gpurun_out/cfg1_vga_timeline.txt 2>&1
python tools/track_timing.py kitti > gpurun_out/cfg2_kitti_timeline.txt 2>&1
python bench.py --workload fast  --steps $S --warmup 5 --no-cpu-baseline > gpurun_out/cfg3_fast.json  2> gpurun_out/cfg3_fast.err
python bench.py --workload 1080p --steps 100  --warmup 5 --no-cpu-baseline > gpurun_out/cfg4_1080p.json 2> gpurun_out/cfg4_1080p.err
python tools/track_timing.py 1080p > gpurun_out/cfg4_1080p_timeline.txt 2>&1 && \
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:'k_track|k_populate|k_icp_depth|k_maps_pyramid' -s 10 -c 10 --csv --log-file gpurun_out/cfg4_1080p_dram.csv python tools/track_timing.py 1080p > gpurun_out/cfg4_ncu.log 2>&1
for f in cfg1_vga cfg3_fast cfg4_1080p; do python - <<PY
import json
try:
    j = json.load(open("gpurun_out/$f.json"))
    s = j.get("latency", {})
    print("$f", "value", round(j["value"], 1), "e2e", round(j["e2e"]["value"], 1), "frac", round(j["roofline"]["frac"], 3), "step-anchored", round(j["roofline"]["frac_step_anchored"], 3),
          "one tracker", round(s.get("value", 0), 1), "lat", s.get("host_latency_ms"), "us/GN it", round(s.get("us_per_gn_iteration", 0), 2), "track_us", s.get("track_us"))
except Exception as e:
    print("$f failed", e)
PY
done
