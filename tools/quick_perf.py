"""Tuning helper: one process, one set of resident frames, many launch shapes.

    python tools/quick_perf.py [workload] [steps] cfg cfg ...     cfg = <trackers>x<threads>[x<grid>]

Prints frames/s of the open-loop sequence (dms_sequence_run, inputs resident in HBM) per configuration, plus the mean
device time of the tracking kernel.  Not a bench line: bench.py is the contract."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from densemonoslam_b200 import synth
from densemonoslam_b200.sequence import SequenceTracker

name = sys.argv[1] if len(sys.argv) > 1 else "kitti"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 100
cfgs = sys.argv[3:] or ["1x1024", "8x512"]
cam = {"vga": synth.CAM_VGA, "kitti": synth.CAM_KITTI, "1080p": synth.CAM_1080P}[name]
ring = 8
if name == "kitti":
    frames = synth.kitti_sequence(cam, ring, seed=2021)
else:
    scene = synth.room_scene(2021); rng = np.random.RandomState(2021); T = np.eye(4); frames = []
    for f in range(ring):
        T = T @ synth.pose_from(rng.randn(3) * 0.004, rng.randn(3) * 0.006)
        prior = T @ synth.pose_from(rng.randn(3) * 0.006, rng.randn(3) * 0.01)
        frames.append({"model": synth.render(scene, cam, prior, seed=2021 + 2 * f, hole_frac=0.0), "live": synth.render(scene, cam, T, seed=2022 + 2 * f),
                       "prior": prior.astype(np.float32)})
dev = [{"v": torch.from_numpy(f["model"]["vmap4"]).cuda(), "n": torch.from_numpy(f["model"]["nmap4"]).cuda(), "mi": torch.from_numpy(f["model"]["rgba"]).cuda(),
        "d": torch.from_numpy(f["live"]["depth"]).cuda(), "li": torch.from_numpy(f["live"]["rgba"]).cuda(), "prior": f["prior"]} for f in frames]


HOST = os.environ.get("QP_HOST", "0") == "1"   # end-to-end shape: pinned host inputs, uploads inside the timed region
if HOST:
    dev = [{k: (v.cpu().pin_memory() if hasattr(v, "cpu") else v) for k, v in d.items()} for d in dev]


def frame_list(n):
    return [dev[i % ring] for i in range(n)]


for cfg in cfgs:
    parts = [int(x) for x in cfg.split("x")]
    trackers, threads = parts[0], parts[1]
    grid = parts[2] if len(parts) > 2 else 0
    seq = SequenceTracker(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, n_trackers=trackers, block_threads=threads if trackers > 1 or threads != 1024 else 0,
                          grid=grid, wait_mode=int(os.environ.get("QP_WAIT", "2")) if trackers > 1 else 0, timing=int(os.environ.get("QP_TIMING", "3")))
    first_prev = dev[-1]["li"]
    seq.run(seq.make_frames(frame_list(3 * trackers)), first_prev, HOST)
    fc = seq.make_frames(frame_list(steps))
    torch.cuda.synchronize()
    best = None; alls = []
    for rep in range(int(os.environ.get("QP_REPS", "3"))):
        t0 = time.perf_counter()
        res = seq.run(fc, first_prev, HOST)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        best = dt if best is None or dt < best else best; alls.append(round(steps / dt))
    k_ms = float(np.mean([r.track_ms for r in res]))
    print(f"{name} {"host" if HOST else "dev"} {cfg:>12s}: {steps / best:8.0f} frames/s  k_track {k_ms * 1000:7.1f} us  resident {seq.resident}  all {alls}", flush=True)
    del seq
