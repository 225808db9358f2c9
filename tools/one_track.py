"""A few frames of the open-loop sequence on one tracker (target of ncu captures): python tools/one_track.py [workload] [frames] [threads] [grid]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from densemonoslam_b200 import synth
from densemonoslam_b200.sequence import SequenceTracker
name = sys.argv[1] if len(sys.argv) > 1 else "kitti"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4
threads = int(sys.argv[3]) if len(sys.argv) > 3 else 0
grid = int(sys.argv[4]) if len(sys.argv) > 4 else 0
cam = {"vga": synth.CAM_VGA, "kitti": synth.CAM_KITTI, "1080p": synth.CAM_1080P}[name]
frames = synth.kitti_sequence(cam, 4, seed=2021)
dev = [{"v": torch.from_numpy(f["model"]["vmap4"]).cuda(), "n": torch.from_numpy(f["model"]["nmap4"]).cuda(), "mi": torch.from_numpy(f["model"]["rgba"]).cuda(),
        "d": torch.from_numpy(f["live"]["depth"]).cuda(), "li": torch.from_numpy(f["live"]["rgba"]).cuda(), "prior": f["prior"]} for f in frames]
seq = SequenceTracker(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, n_trackers=1, block_threads=threads, grid=grid, wait_mode=0, timing=int(os.environ.get("OT_TIMING", "3")))  # 3 = the production build of k_track (1 = instrumented)
res = seq.run(seq.make_frames([dev[i % 4] for i in range(n)]), dev[-1]["li"], False)
torch.cuda.synchronize()
print("track_ms", [round(r.track_ms, 3) for r in res])
