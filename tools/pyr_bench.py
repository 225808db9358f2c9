"""Times the pyramid-construction entry points (CUDA events, mean of N back-to-back calls on resident inputs): python tools/pyr_bench.py [workload] [N]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from densemonoslam_b200 import synth
from densemonoslam_b200.odometry import RGBDOdometry
name = sys.argv[1] if len(sys.argv) > 1 else "kitti"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 200
cam = {"vga": synth.CAM_VGA, "kitti": synth.CAM_KITTI, "1080p": synth.CAM_1080P}[name]
f = synth.kitti_sequence(cam, 1, seed=2021)[0] if name == "kitti" else None
if f is None:
    model, live, _ = synth.frame_pair(cam, seed=1234)
else:
    model, live = f["model"], f["live"]
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
mv, mn, mi, ld, li = dev(model["vmap4"]), dev(model["nmap4"]), dev(model["rgba"]), dev(live["depth"]), dev(live["rgba"])
g = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
pose = np.eye(4, dtype=np.float32)
calls = {"initICPModel (k_maps_pyramid)": lambda: g.initICPModel(mv, mn, 25.0, pose), "initRGBModel (k_populate)": lambda: g.initRGBModel(mi),
         "initICP depth (k_icp_depth)": lambda: g.initICP(ld, 25.0), "initRGB (k_populate)": lambda: g.initRGB(li), "initFirstRGB (k_populate, image only)": lambda: g.initFirstRGB(mi)}
px = cam.width * cam.height
for k, fn in calls.items():
    for _ in range(10): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(N): fn()
    e1.record(); torch.cuda.synchronize()
    print(f"{name} {k:42s} {e0.elapsed_time(e1) * 1000 / N:7.2f} us per call (includes the launch gap of back-to-back calls)")
