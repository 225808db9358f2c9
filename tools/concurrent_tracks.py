"""Experiment: aggregate frames/s when n independent trackers (handles) run concurrently on one GPU,
each on its own stream with 1/n of the SMs (DMS_TRACK_GRID)."""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
grid = int(sys.argv[2]) if len(sys.argv) > 2 else 148 // n
frames_per = int(sys.argv[3]) if len(sys.argv) > 3 else 200
os.environ["DMS_TRACK_GRID"] = str(grid)
from densemonoslam_b200 import synth
from densemonoslam_b200.odometry import RGBDOdometry

cam = synth.CAM_KITTI
fr = synth.kitti_sequence(cam, 4, seed=2021)
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
D = [{"v": dev(f["model"]["vmap4"]), "n": dev(f["model"]["nmap4"]), "mi": dev(f["model"]["rgba"]), "d": dev(f["live"]["depth"]), "li": dev(f["live"]["rgba"])} for f in fr]
handles = [RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy) for _ in range(n)]
streams = [torch.cuda.Stream() for _ in range(n)]
for h, s in zip(handles, streams):
    h.initFirstRGB(D[0]["li"], stream=s)
torch.cuda.synchronize()

def work(h, s, count):
    for i in range(count):
        f, d = fr[i % 4], D[i % 4]; p = f["prior"]
        h.initICPModel(d["v"], d["n"], 25.0, p, stream=s); h.initRGBModel(d["mi"], stream=s); h.initICP(d["d"], 25.0, stream=s); h.initRGB(d["li"], stream=s)
        h.getIncrementalTransformation(p[:3, 3], p[:3, :3], False, 10.0, True, False, True, stream=s)

for h, s in zip(handles, streams):
    work(h, s, 5)
torch.cuda.synchronize()
t0 = time.perf_counter()
ths = [threading.Thread(target=work, args=(h, s, frames_per)) for h, s in zip(handles, streams)]
[t.start() for t in ths]; [t.join() for t in ths]
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print(f"n={n} grid={grid}: {n * frames_per / dt:.1f} frames/s aggregate, {1000 * dt / frames_per:.3f} ms per frame per tracker")
