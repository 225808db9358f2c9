"""Prints the per-phase timing split of k_track for one workload (uses the library's timing mode)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from densemonoslam_b200 import synth
from densemonoslam_b200.odometry import RGBDOdometry

name = sys.argv[1] if len(sys.argv) > 1 else "vga"
icp_weight = float(sys.argv[2]) if len(sys.argv) > 2 else 10.0   # 100 = ICP only (RO.cpp:279)
cam = {"vga": synth.CAM_VGA, "kitti": synth.CAM_KITTI, "1080p": synth.CAM_1080P, "small": synth.scaled_camera(synth.CAM_VGA, 4)}[name]
model, live, T = synth.frame_pair(cam, seed=1234)
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
g = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
g.set_timing(int(os.environ.get('TL_MODE', '2')))
pose = np.eye(4, dtype=np.float32)
mv, mn, mi, ld, li = dev(model["vmap4"]), dev(model["nmap4"]), dev(model["rgba"]), dev(live["depth"]), dev(live["rgba"])
g.initFirstRGB(mi)
acc = {}
N = 30
for i in range(N + 5):
    g.initICPModel(mv, mn, 25.0, pose); g.initRGBModel(mi); g.initICP(ld, 25.0); g.initRGB(li)
    g.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], False, icp_weight, True, False, True)
    if i >= 5:
        for k, v in g.timing().items():
            acc[k] = acc.get(k, 0.0) + v / N
print(name, json.dumps({k: round(v * 1000, 1) for k, v in acc.items()}), "us; iters", g.iters, "so3", g.so3_iters)

tl_all = g.timeline()
tl = [e for e in tl_all if e[0] < 200]
stl = [e for e in tl_all if e[0] >= 200]
sagg = {}
for (a, ta), (b, tb) in zip(stl[:-1], stl[1:]):
    if b > a:
        v = sagg.setdefault(f"solver {a}->{b}", [0, 0.0]); v[0] += 1; v[1] += (tb - ta) / 1000.0
for k, (n, tot) in sagg.items():
    print(f"  {k:30s} n={n:3d} mean={tot / n:6.2f} us")
t0 = tl[0][1]
names = {10: "so3 start", 11: "so3 flushed", 12: "so3 barrier passed", 13: "so3 solved", 20: "A start L0", 21: "A start L1", 22: "A start L2",
         30: "A flushed", 31: "A barrier passed", 32: "A serial done", 40: "B flushed", 41: "B barrier passed", 43: "B scratch solve done", 42: "B solved", 99: "end"}
prev = t0
agg = {}
last_tag = None
for tag, t in tl:
    if last_tag is not None:
        key = f"{names.get(last_tag, last_tag)} -> {names.get(tag, tag)}"
        a = agg.setdefault(key, [0, 0.0]); a[0] += 1; a[1] += (t - prev) / 1000.0
    prev = t; last_tag = tag
for k, (n, tot) in list(agg.items())[:int(os.environ.get("TL_ROWS", "100"))]:
    print(f"  {k:55s} n={n:3d} mean={tot / n:7.2f} us total={tot:8.1f}")
