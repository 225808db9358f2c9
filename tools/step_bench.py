"""Times the four stand-alone GN operators (same device functions as the phases of k_track, no barriers)
at level 0 of one workload: the pure pixel-phase throughput of the GPU."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from densemonoslam_b200 import synth
from densemonoslam_b200.odometry import RGBDOdometry

name = sys.argv[1] if len(sys.argv) > 1 else "kitti"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
cam = {"vga": synth.CAM_VGA, "kitti": synth.CAM_KITTI, "1080p": synth.CAM_1080P}[name]
fr = synth.kitti_sequence(cam, 2, seed=2021)[1] if name == "kitti" else None
if fr is None:
    model, live, T = synth.frame_pair(cam, seed=1234); pose = np.eye(4, dtype=np.float32)
else:
    model, live, pose = fr["model"], fr["live"], fr["prior"]
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
g = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
mv, mn, mi, ld, li = dev(model["vmap4"]), dev(model["nmap4"]), dev(model["rgba"]), dev(live["depth"]), dev(live["rgba"])
g.initFirstRGB(mi); g.initICPModel(mv, mn, 25.0, pose); g.initRGBModel(mi); g.initICP(ld, 25.0); g.initRGB(li)
g.computeDerivativeImages()
R = pose[:3, :3].astype(np.float32); t = pose[:3, 3].astype(np.float32)
K = np.array([[cam.fx, 0, cam.cx], [0, cam.fy, cam.cy], [0, 0, 1]], np.float64)
krkinv = (K @ np.eye(3) @ np.linalg.inv(K)).astype(np.float32); kt = np.zeros(3, np.float32)
npx = cam.width * cam.height

def timed(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000.0 / reps

res = {}
res["icpStep"] = timed(lambda: g.icpStep(0, R, t, np.linalg.inv(R).astype(np.float32), t))
res["computeRgbResidual"] = timed(lambda: g.computeRgbResidual(0, kt, krkinv))
cnt, sig = g.computeRgbResidual(0, kt, krkinv)
res["rgbStep"] = timed(lambda: g.rgbStep(0, float(np.sqrt(max(cnt, 1)))))
res["computeDerivativeImages"] = timed(lambda: (g.computeDerivativeImages(), torch.cuda.synchronize()))
out, n = g.icpStep(0, R, t, np.linalg.inv(R).astype(np.float32), t)
print(name, "level 0, us per call incl. launch + result readback:", json.dumps({k: round(v, 1) for k, v in res.items()}), "icp inliers", n, "rgb terms", cnt)
print("  algorithmic GB/s: icp", round(48 * npx / res["icpStep"] / 1e3, 1), "rgbres+rgbstep", round(14 * npx / (res["computeRgbResidual"] + res["rgbStep"]) / 1e3, 1))
