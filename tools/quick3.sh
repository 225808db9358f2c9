#!/bin/bash
# three tuning points: KITTI pipelined, KITTI single tracker, VGA pipelined
p() { python -c "
import sys,json
j=json.loads(sys.stdin.read()); print('$1', round(j['value'],1), 'frac', round(j['roofline']['frac'],3), 'kernel_ms', round(j['roofline']['kernel_ms'],3))"; }
python bench.py --steps 400 --warmup 5 --quick 2>/dev/null | p kitti_pipe8
python bench.py --steps 300 --warmup 5 --quick --pipeline 1 2>/dev/null | p kitti_single
python bench.py --steps 300 --warmup 5 --quick --workload vga 2>/dev/null | p vga_pipe8
