"""(lives under tests/ because it times the CPU restatements of oracle/ beside the kernels)
Device timings (CUDA events, inputs resident) of the SURVEY 8f "next" rows beside their CPU restatements:
depth pre-filter (N2), FillIn / denseEnough (N1), surfel-splat prediction (N3), NID score (N4)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch
from densemonoslam_b200 import synth, splat, frontend
from densemonoslam_b200.nid import MutualInformation
from densemonoslam_b200.odometry import RGBDOdometry, filterDepth, metriciseDepth
from oracle import oracle as orc, splat_ref, nid_ref, frontend_ref

dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()


def gpu_us(fn, reps=50):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000.0 / reps


def cpu_ms(fn, reps=2):
    fn()
    t = time.perf_counter()
    for _ in range(reps): fn()
    return (time.perf_counter() - t) * 1000.0 / reps


out = {}
cam = synth.CAM_KITTI
fr = synth.kitti_sequence(cam, 2, seed=2021)[1]
model, live = fr["model"], fr["live"]
npx = cam.width * cam.height
c = (cam.cx, cam.cy, cam.fx, cam.fy)
d_raw = dev(live["depth"]); d_rgba = dev(live["rgba"]); d_mv = dev(model["vmap4"]); d_mn = dev(model["nmap4"]); d_mi = dev(model["rgba"])

us = gpu_us(lambda: filterDepth(d_raw, 25.0))
out["N2 depth_bilateral 1241x376"] = {"gpu_us": us, "algorithmic_GBps": 4 * npx / us / 1e3, "taps_per_s": 169 * npx / us * 1e6,
                                      "cpu_oracle_ms": cpu_ms(lambda: orc.depth_bilateral(live["depth"], 25.0))}
us = gpu_us(lambda: metriciseDepth(d_raw, 25.0))
out["N2 depth_metric 1241x376"] = {"gpu_us": us, "algorithmic_GBps": 6 * npx / us / 1e3}
filt = filterDepth(d_raw, 25.0)
us = gpu_us(lambda: frontend.fill_in(d_mv, d_mn, d_mi, filt, d_rgba, c))
out["N1 fill_in 1241x376"] = {"gpu_us": us, "algorithmic_GBps": (36 + 2 + 4 + 36) * npx / us / 1e3,
                              "cpu_oracle_ms": cpu_ms(lambda: frontend_ref.fill_in(model["vmap4"], model["nmap4"], model["rgba"], filt.cpu().numpy(), live["rgba"], c))}
t0 = time.perf_counter(); [frontend.dense_fraction(d_mi) for _ in range(50)]; out["N1 dense_fraction 1241x376 (incl. sync + read-back)"] = {"host_us": (time.perf_counter() - t0) * 1e6 / 50}

# N3: a surfel map of ~1.4 M surfels (three overlapping full-resolution views), predicted into the KITTI camera
views = [np.eye(4), synth.pose_from((0, 0.03, 0), (0.4, 0, 0.3)), synth.pose_from((0, -0.03, 0), (-0.4, 0, 0.6))]
scene = synth.street_scene(2021)
S = np.concatenate([splat_ref.surfels_from_view(v["vmap4"], v["nmap4"], v["rgba"], T, c, time=5)
                    for T in views for v in [synth.render(scene, cam, T, seed=3, max_depth=60.0, hole_frac=0.0)]])
dS = dev(S)
pose = synth.pose_from((0.002, 0.01, 0.0), (0.1, 0.0, 0.2)).astype(np.float32)
ti = splat.invert_pose(pose)
us = gpu_us(lambda: splat.combinedPredict(dS, pose, c, cam.height, cam.width, 25.0, 5.0, 6, 0, 6, 200, 1, t_inv=ti), reps=20)
img, ver, _, _ = splat.combinedPredict(dS, pose, c, cam.height, cam.width, 25.0, 5.0, 6, 0, 6, 200, 1, t_inv=ti)
out["N3 splat_predict 1241x376"] = {"surfels": int(len(S)), "gpu_us": us, "surfels_per_s": len(S) / us * 1e6, "algorithmic_GBps": (60 * len(S) + 44 * npx) / us / 1e3,
                                    "coverage": float((ver[..., 2] > 0).float().mean()),
                                    "cpu_oracle_ms": cpu_ms(lambda: splat_ref.predict(S, ti, c, cam.height, cam.width, 25.0, 5.0, 6, 0, 6, 200, 1), reps=1),
                                    "note": "includes the allocation of the four output tensors by the Python wrapper"}

# N4
g = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
g.initICP(d_raw, 25.0); g.initRGB(d_rgba)
mi = MutualInformation(cam.width, cam.height, 500, 64, 1)
us_kf = gpu_us(lambda: mi.setKeyFrame(d_mv, d_mi, d_mv, d_mi, 25.0))
t0 = time.perf_counter(); [mi.score(g, 25.0, 0.5) for _ in range(50)]; host_us = (time.perf_counter() - t0) * 1e6 / 50
o = orc.OracleOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy); o.initICP(live["depth"], 25.0); o.initRGB(live["rgba"])
out["N4 nid 1241x376 level 1"] = {"set_keyframe_gpu_us": us_kf, "score_host_us (2 histograms + entropies + sync)": host_us,
                                  "cpu_oracle_ms": cpu_ms(lambda: nid_ref.nid_score(model["vmap4"], model["rgba"], model["vmap4"], model["rgba"], 25.0,
                                                                                    o.buffer("nextImage", 1), o.buffer("nextDepth", 1), 1, 64, 500, 25.0, 0.5), reps=1)}
# whole front-end frame without GL: depth filter -> splat prediction -> FillIn / denseEnough -> frame-to-model track
# -> second prediction + inactive prediction -> model-to-model registration and its gates
from densemonoslam_b200.frontend import TrackingFrontEnd
S2 = dev(S)


def predictor(p, kind, tick):
    img, ver, nor, _ = splat.combinedPredict(S2, p, c, cam.height, cam.width, 25.0, 5.0, 6, 0, 6, 200, splat.ACTIVE if kind != 2 else splat.ACTIVE)
    return {"vertex": ver, "normal": nor, "image": img}


for loops in (0, 1):
    fe = TrackingFrontEnd(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, predictor, depthCutoff=25.0, closeLoops=loops)
    seq = synth.kitti_sequence(cam, 6, seed=2021)
    frames_dev = [(dev(f["live"]["rgba"]), dev(f["live"]["depth"]), f["prior"]) for f in seq]
    fe.process(*frames_dev[0])
    torch.cuda.synchronize()
    t0 = time.perf_counter(); n = 0
    for rep in range(8):
        for rgba, depth, prior in frames_dev[1:]:
            fe.process(rgba, depth, prior); n += 1
    ms = (time.perf_counter() - t0) * 1e3 / n
    out[f"front end frame 1241x376, 1.3 M surfels, closeLoops={loops}"] = {"host_ms_per_frame": ms, "frames_per_s": 1000.0 / ms}
print(json.dumps(out, indent=1))
