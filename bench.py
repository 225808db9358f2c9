#!/usr/bin/env python
"""Benchmark of the dense tracking hot path (BASELINE.json metric: odometry frames/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload kitti|vga|1080p|fast]

A step = one tracked frame of the named workload: initICPModel + initRGBModel + initICP(depth) +
initRGB + getIncrementalTransformation (ICP+RGB, 3-level pyramid, icpWeight 10, SO(3) on), i.e. the
call sequence of ElasticFusion::processFrame (EF.cpp:173-201).  Default workload = BASELINE.json
configs[1]: KITTI shape 1241x376, synthetic forward-driving sequence, open loop.

`value`     frames/s with the frame inputs already resident in HBM (a ring of distinct frames
            larger than L2), timed with CUDA events on the launching stream.
`e2e`       the same through the host-buffer C-ABI entry points (pinned host inputs, H2D copies and
            the D2H result read inside the timed region).
`roofline`  the tracking kernel k_track: algorithmic bytes per launch / mean launch duration
            (CUDA events around each launch, recorded inside the timed region).
`cpu_baseline` / --impl reference: the CPU oracle (an OpenMP transcription of the reference's
            equations - the reference has no CPU implementation of this path) on the host cores.
N > 1 (torchrun): one independent sensor sequence per GPU (the collaborative --n mode maps one
tracker per GPU); no data-path collective; value = frames of all ranks / max-over-ranks time.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

# Trackers sharing a GPU run on one stream each; streams beyond the number of hardware queues (8 by default) alias and
# serialise.  Must be set before CUDA starts.  (The library sizes its tracking kernels by this variable.)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLAGS = dict(rgbOnly=False, icpWeight=10.0, pyramid=True, fastOdom=False, so3=True)


def traffic_of(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of one k_track launch of this workload, from the committed ncu summary
    (profiles/r02_traffic.json, written by tools/ncu_summary.py from an `ncu --set full` capture); None without a capture."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))[workload]["k_track_dram_bytes"]
    except Exception:
        return None


def workload_def(name):
    from densemonoslam_b200 import synth
    if name == "kitti":
        return synth.CAM_KITTI, "KITTI-shape 1241x376 synthetic sequence, ICP+RGB, 3-level pyramid, so3 on, open loop (configs[1])"
    if name == "vga":
        return synth.CAM_VGA, "640x480 synthetic RGB-D pair sequence, ICP+RGB, 3-level pyramid, icpWeight=10, so3 on (configs[0])"
    if name == "1080p":
        return synth.CAM_1080P, "1920x1080 synthetic RGB-D stress sequence, full pyramid (configs[3])"
    if name == "fast":
        return synth.CAM_VGA, "640x480 fast odometry (single level, 3 iterations) (configs[2])"
    raise SystemExit(f"unknown workload {name}")


def make_frames(name, cam, n, seed):
    """n distinct frames: model buffers rendered at the prior pose, live buffers at the GT pose."""
    from densemonoslam_b200 import synth
    if name == "kitti":
        return synth.kitti_sequence(cam, n, seed=seed)
    scene = synth.room_scene(seed)
    frames = []
    rng = np.random.RandomState(seed)
    T = np.eye(4)
    for f in range(n):
        T = T @ synth.pose_from(rng.randn(3) * 0.004, rng.randn(3) * 0.006)
        prior = T @ synth.pose_from(rng.randn(3) * 0.006, rng.randn(3) * 0.01)
        model = synth.render(scene, cam, prior, seed=seed + 2 * f, hole_frac=0.0)
        live = synth.render(scene, cam, T, seed=seed + 2 * f + 1)
        frames.append({"model": model, "live": live, "prior": prior.astype(np.float32), "gt": T.astype(np.float32)})
    return frames


def level_pixels(cam):
    return [(cam.height >> l) * (cam.width >> l) for l in range(3)]


def algorithmic_bytes(cam, iters, so3_iters):
    """62 B per pixel per fused ICP+RGB GN iteration (48 ICP + 14 RGB), 2 B per pixel
    per so3 evaluation (level 2), derivative images 1 B read + 4 B written per pixel of every level."""
    n = level_pixels(cam)
    gn = sum(62 * n[l] * iters[l] for l in range(3))
    so3 = 2 * n[2] * so3_iters
    der = 5 * sum(n)
    return gn + so3 + der


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        # the first query waits until the trackers' launches of the first round are out: an NVML call takes driver locks, and
        # one that coincides with the launches at the start of a 4 ms timed region was seen to cost the region 0.8 ms
        time.sleep(float(os.environ.get("DMS_CLOCK_SAMPLER_DELAY_S", "0.0012")))
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.01)

    def result(self):
        self.stop_flag = True
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def run_oracle_frames(cam, frames, n_frames, budget_s, threads=None):
    """Times the CPU oracle on the same per-frame call sequence; returns (frames_done, seconds, threads)."""
    from oracle import oracle as orc
    # all host cores unless told otherwise (torchrun exports OMP_NUM_THREADS=1, which must not throttle the CPU arm)
    orc.set_num_threads(threads if threads else len(os.sched_getaffinity(0)))
    o = orc.OracleOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
    o.initFirstRGB(frames[0]["live"]["rgba"])
    done, t0 = 0, time.perf_counter()
    while done < n_frames:
        f = frames[done % len(frames)]
        p = f["prior"]
        o.initICPModel(f["model"]["vmap4"], f["model"]["nmap4"], p)
        o.initRGBModel(f["model"]["rgba"])
        o.initICP(f["live"]["depth"], 25.0)
        o.initRGB(f["live"]["rgba"])
        o.getIncrementalTransformation(p[:3, 3].copy(), p[:3, :3].copy(), FLAGS["rgbOnly"], FLAGS["icpWeight"], FLAGS["pyramid"],
                                       FLAGS["fastOdom"], FLAGS["so3"], keepLog=False)
        done += 1
        if time.perf_counter() - t0 > budget_s:
            break
    return done, time.perf_counter() - t0, orc.num_threads()


def run_legacy_frames(cam, frames, n_frames, budget_s):
    """Reported baseline: the reference's own legacy kernels (oracle/_ref, built unmodified from /root/reference for
    sm_100a) driven by the reference's host loop (blocking sync + download after every kernel), inputs resident on the device."""
    try:
        from oracle import ref_legacy
        if not ref_legacy.available():
            return None
        r = ref_legacy.LegacyOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
        n = min(len(frames), 8)
        for i in range(n):
            f = frames[i]
            r.upload(2 * i, v4=f["model"]["vmap4"], n4=f["model"]["nmap4"], rgba=f["model"]["rgba"])
            r.upload(2 * i + 1, rgba=f["live"]["rgba"], depth=f["live"]["depth"])
        r.initFirstRGB(1)

        def one(i):
            f = frames[i % n]; p = f["prior"]; a, b = 2 * (i % n), 2 * (i % n) + 1
            r.initICPModel(a, p); r.initRGBModel(a); r.initICP(b, 25.0); r.initRGB(b)
            return r.getIncrementalTransformation(p[:3, 3].copy(), p[:3, :3].copy(), FLAGS["rgbOnly"], FLAGS["icpWeight"], FLAGS["pyramid"], FLAGS["fastOdom"], FLAGS["so3"])
        for i in range(3):
            one(i)
        done, t0 = 0, time.perf_counter()
        while done < n_frames and time.perf_counter() - t0 < budget_s:
            one(done); done += 1
        secs = time.perf_counter() - t0
        return {"value": done / secs, "unit": "frames/s", "ms_per_frame": 1000.0 * secs / done, "sample": f"{done} frames",
                "note": "reference's own reduce.cu / cudafuncs.cu compiled for sm_100a + the reference's host loop (sync and D2H after every kernel); reported baseline, not the target"}
    except Exception as e:  # a baseline must never break the bench line
        return {"unavailable": repr(e)}


_REAL_STDOUT = None


def run_frontend_e2e(cam, steps, device):
    """Frame-to-model variant of the end-to-end leg: what a caller of the reference really uploads per
    frame is the sensor frame (RGBA8 + u16 depth, 6 B per pixel); the model view is predicted on the device from the surfel
    map (dms_splat_predict = IndexMap::combinedPredict) and the frame runs through the processFrame-shaped front end
    (dms_frontend_process: depth filter, prediction, FillIn / denseEnough, frame-to-model track with the previous pose as
    the prior).  One frame in flight by construction: frame f + 1 is predicted from the pose of frame f.  Returns the block
    reported as `e2e_frontend`."""
    import torch
    from densemonoslam_b200 import splat, synth
    from densemonoslam_b200.frontend import TrackingFrontEnd
    scene = synth.street_scene(2021)
    n_ring = 8
    poses, T = [], np.eye(4)
    for f in range(n_ring):  # slow forward motion with a little yaw: the three map views below cover all of it
        T = T @ synth.pose_from((0.0, np.deg2rad(0.15) * np.sin(f * 0.7), 0.0), (0.002 * np.cos(f), 0.0, 0.03))
        poses.append(T.copy())
    S = np.concatenate([synth.surfels_from_view(synth.render(scene, cam, poses[k], seed=77 + k, max_depth=60.0, hole_frac=0.0, zero_frac=0.0), poses[k], cam, time=1)
                        for k in (0, 3, 7)])
    d_S = torch.from_numpy(S).to(device)
    lives = [synth.render(scene, cam, poses[f], seed=300 + f, max_depth=60.0, depth_noise=0.004) for f in range(n_ring)]
    pin = [(torch.from_numpy(l["rgba"]).pin_memory(), torch.from_numpy(l["depth"]).pin_memory()) for l in lives]
    d_rgba = [torch.empty_like(pin[0][0], device=device) for _ in range(2)]
    d_depth = [torch.empty_like(pin[0][1], device=device) for _ in range(2)]
    c = (cam.cx, cam.cy, cam.fx, cam.fy)

    def predictor(pose, kind, tick):
        img, ver, nor, _ = splat.combinedPredict(d_S, pose, c, cam.height, cam.width, 25.0, 5.0, 2, 0, 2, 200, splat.ACTIVE)
        return {"vertex": ver, "normal": nor, "image": img}

    fe = TrackingFrontEnd(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, predictor, device=device.index or 0, depthCutoff=25.0, closeLoops=0)
    order = list(range(n_ring)) + list(range(n_ring - 2, 0, -1))   # there and back again: the motion stays continuous when the ring wraps

    def frame(i, first=False):
        k = order[i % len(order)]
        b = i & 1
        d_rgba[b].copy_(pin[k][0], non_blocking=True); d_depth[b].copy_(pin[k][1], non_blocking=True)
        return k, fe.process(d_rgba[b], d_depth[b], poses[k].astype(np.float32) if first else None)

    frame(0, True)
    for i in range(1, 4):
        frame(i)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    ok = 0
    for i in range(4, 4 + steps):
        k, r = frame(i)
        ok += int(r["tracked"] and r["trackingOk"])
    ev1.record(); torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    err = float(np.max(np.abs(r["pose"][:3, 3] - poses[k][:3, 3])))
    return {"value": steps / (ms * 1e-3), "unit": "frames/s", "ms_per_frame": ms / steps, "steps": steps,
            "h2d_bytes_per_step": int(pin[0][0].numel() * pin[0][0].element_size() + pin[0][1].numel() * pin[0][1].element_size()),
            "d2h_bytes_per_step": int(_lib_track_d2h()), "surfels": int(len(S)), "frames_tracked": ok,
            "last_frame_abs_translation_error_m": err,
            "note": "sensor frame uploaded from pinned host memory, model view predicted on the device from the surfel map, "
                    "front end with the previous pose as the prior; one frame in flight; the Python callback of the predictor is inside the timed region"}


def _lib_track_d2h():
    from densemonoslam_b200 import _lib
    return _lib.load().dms_track_d2h_bytes()


def _guard_stdout():
    """Libraries (NCCL's version banner, OpenMP) may write to fd 1; the contract is ONE JSON line on stdout.  Everything
    else goes to stderr, the JSON line to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(obj):
    _REAL_STDOUT.write(json.dumps(obj) + "\n")
    _REAL_STDOUT.flush()


def main():
    _guard_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="kitti")
    ap.add_argument("--ring", type=int, default=8, help="distinct frames kept resident (must exceed L2 in total)")
    ap.add_argument("--pipeline", type=int, default=0, help="trackers sharing the GPU concurrently (1 = strictly sequential frames; 0 = auto: 20, fewer when the run has fewer frames)")
    ap.add_argument("--threads", type=int, default=512, help="threads per block of the tracking kernel when trackers share the GPU (0 = library default, 768)")
    ap.add_argument("--grid", type=int, default=0, help="thread blocks per tracking kernel (default: SMs / pipeline depth)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--wait", type=int, default=-1, help="how pipelined trackers wait for a result: 0 spin, 1 blocking event, 2 sleep then poll + yield, 3 sleep then poll + 20 us sleeps; default: 2, or 3 when the ranks of this box have more than two tracker threads per host core")
    ap.add_argument("--driver", default="native", choices=["native", "python"], help="who runs the trackers' host threads when they share the GPU: the library's sequence runner or Python threads")
    ap.add_argument("--no-share-rgb", action="store_true", help="pipelined trackers reload the previous frame's image instead of taking its pyramid from the tracker that has it")
    ap.add_argument("--quick", action="store_true", help="tuning runs: only the device-resident leg")
    args = ap.parse_args()

    if args.wait < 0:
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
        # sleep-then-poll with short sleeps only when the box is heavily oversubscribed (more than two tracker threads per core);
        # twenty yielding pollers on sixteen cores were measured faster than twenty 20-us sleepers
        args.wait = 3 if local_world * max(1, args.pipeline if args.pipeline else min(20, max(2, args.steps))) > 2 * len(os.sched_getaffinity(0)) else 2
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    cam, wl_desc = workload_def(args.workload)
    if args.workload == "fast":
        FLAGS.update(pyramid=False, fastOdom=True)
    iters_nominal = [3 if FLAGS["fastOdom"] else 10, 5 if FLAGS["pyramid"] else 0, 4 if FLAGS["pyramid"] else 0]
    config = {"workload": wl_desc, "width": cam.width, "height": cam.height, "icpWeight": FLAGS["icpWeight"], "iterations": iters_nominal,
              "so3": FLAGS["so3"], "parallelism": f"{world} independent sensor tracker(s), one per GPU, no collective on the path"}
    config_exec = {}  # how this arm runs the workload (not part of the workload: both arms print the same `config`)

    # ------------------------------------------------------------------ reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return
        frames = make_frames(args.workload, cam, min(args.ring, 4), seed=2021)
        run_oracle_frames(cam, frames, max(1, min(args.warmup, 2)), 30.0)
        done, secs, threads = run_oracle_frames(cam, frames, args.steps, 150.0)
        fps = done / secs
        sample = f"{done} of {args.steps} frames (time-bounded), same per-frame call sequence"
        emit({
            "impl": "reference", "metric": "odometry_frames_per_s", "value": fps, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 / fps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "the reference has no CPU implementation of this path; this is the OpenMP transcription of its equations (oracle/cpu_odom.cpp)"})
        return

    # ------------------------------------------------------------------ our arm (GPU)
    import torch
    import torch.distributed as dist
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from densemonoslam_b200.odometry import RGBDOdometry
    from densemonoslam_b200 import _lib

    ring = max(2, args.ring)
    frames = make_frames(args.workload, cam, ring, seed=2021 + 1000 * rank)  # each rank = its own sensor
    per_frame_in = sum(frames[0][k][b].nbytes for k, b in (("model", "vmap4"), ("model", "nmap4"), ("model", "rgba"), ("live", "depth"), ("live", "rgba")))
    dev = []
    pinned = []
    for f in frames:
        d = {"v": torch.from_numpy(f["model"]["vmap4"]).cuda(), "n": torch.from_numpy(f["model"]["nmap4"]).cuda(),
             "mi": torch.from_numpy(f["model"]["rgba"]).cuda(), "d": torch.from_numpy(f["live"]["depth"]).cuda(),
             "li": torch.from_numpy(f["live"]["rgba"]).cuda()}
        dev.append(d)
        pinned.append({k: torch.from_numpy(a).pin_memory() for k, a in (("v", f["model"]["vmap4"]), ("n", f["model"]["nmap4"]),
                                                                          ("mi", f["model"]["rgba"]), ("d", f["live"]["depth"]), ("li", f["live"]["rgba"]))})
    config_exec["ring_frames"] = ring
    config_exec["cache"] = f"inputs cycle through {ring} distinct resident frames = {ring * per_frame_in / 1e6:.0f} MB (> 126 MB L2)"
    config_exec["warmup"] = "W frames per tracker, repeated until 0.25 s of untimed work have passed (clock ramp of a fresh process)"

    props = torch.cuda.get_device_properties(local_rank)
    sms = props.multi_processor_count

    share_rgb = not args.no_share_rgb

    def make_pipe(depth):
        """`depth` trackers sharing the GPU (SMs split evenly), each on its own stream.  Frame f goes to tracker
        f % depth; before it, the tracker's SO(3) reference image is (re)loaded from frame f-1 (initFirstRGB), so the
        results are those of the sequential single-tracker run."""
        hs, ss = [], []
        for k in range(depth):
            h = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, device=local_rank)
            threads = args.threads if depth > 1 else 0  # a tracker that owns the GPU keeps the library default (one 768-thread block per SM)
            if threads:
                h.set_block(threads)
            per_sm = max(1, 1024 // threads) if threads else 1
            h.set_grid(args.grid if args.grid else (0 if depth == 1 else (sms * per_sm) // min(depth, int(os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS", "8")))))  # as many kernels resident as there are hardware queues; further trackers queue
            h.set_wait(args.wait if depth > 1 else 0)   # trackers sharing the GPU: do not burn a host core per tracker while waiting
            h.set_timing(3)   # CUDA events around the kernel only: the production build of k_track (mode 1 = instrumented build with phase counters)
            hs.append(h); ss.append(torch.cuda.Stream(device=local_rank) if depth > 1 else torch.cuda.current_stream())
        return hs, ss

    def one_frame(h, st, i, host, prev_tracker=None, sync=None):
        """One tracked frame.  The SO(3) reference image of frame i is the live image of frame i-1: with one tracker it is
        already in place (RO.cpp:595-601); with trackers taking turns it comes from the pyramid of the tracker that
        processed frame i-1 (`prev_tracker`, dms_odom_init_first_rgb_from), or is loaded again when that is not possible
        (first frame of a run)."""
        f = frames[i % ring]
        src = pinned if host else dev
        cur, prev = src[i % ring], src[(i - 1) % ring]
        get = (lambda t: t.numpy()) if host else (lambda t: t)
        p = f["prior"]
        h.initICPModel(get(cur["v"]), get(cur["n"]), 25.0, p, stream=st)
        h.initRGBModel(get(cur["mi"]), stream=st)
        h.initICP(get(cur["d"]), 25.0, stream=st)
        if sync is not None:
            sync["taken"][i].wait()      # the frame that takes this tracker's previous image has issued its copy
        h.initRGB(get(cur["li"]), stream=st)
        if sync is not None:
            sync["issued"][i].set()
        if prev_tracker is not None:
            sync["issued"][i - 1].wait()  # frame i-1's initRGB call has returned
            h.initFirstRGBFrom(prev_tracker, stream=st)
            sync["taken"][i - 1 + len(sync["hs"])].set()
        else:
            h.initFirstRGB(get(prev["li"]), stream=st)
            if sync is not None:
                sync["taken"][i - 1 + len(sync["hs"])].set()
        return h.getIncrementalTransformation(p[:3, 3], p[:3, :3], FLAGS["rgbOnly"], FLAGS["icpWeight"], FLAGS["pyramid"], FLAGS["fastOdom"],
                                              FLAGS["so3"], stream=st)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_native(seq, steps, warmup, host):
        """The same timed region through the library's own sequence runner (dms_sequence_run): the trackers' host threads
        live in C++, so the interpreter lock is not on the path."""
        src = pinned if host else dev

        def frame_list(n):
            return [{"v": src[i % ring]["v"], "n": src[i % ring]["n"], "mi": src[i % ring]["mi"], "d": src[i % ring]["d"], "li": src[i % ring]["li"],
                     "prior": frames[i % ring]["prior"]} for i in range(n)]
        first_prev = src[(0 - 1) % ring]["li"]
        # warm-up: W frames per tracker, repeated until the GPU has been busy for a quarter of a second -- a fresh process finds
        # the GPU at its idle clock, and a 10 ms warm-up in front of a 4 ms timed region was seen to leave the region 5 - 20 %
        # slow on some boxes (clock ramp, first touch of the trackers' memory)
        wf = seq.make_frames(frame_list(max(3, warmup) * max(1, seq.options.n_trackers)))
        t_w = time.perf_counter()
        while True:
            seq.run(wf, first_prev, host)
            torch.cuda.synchronize()
            if time.perf_counter() - t_w > 0.25:
                break
        fc = seq.make_frames(frame_list(steps))
        rring = getattr(seq, "_ring", None)
        if rring is not None and rring.owner:
            torch.cuda.synchronize()
            rring.clear()    # the warm-up frames carry the same frame numbers; producers are idle until the barrier below
        barrier()
        sampler = ClockSampler(local_rank)
        sampler.start()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        res = seq.run(fc, first_prev, host)
        torch.cuda.synchronize()
        if rring is not None and rring.owner:
            # the consumer of the multi-sensor mode: every sensor's per-frame record must have arrived in this GPU's memory
            # (written there by the producing GPU behind each tracking kernel); part of the timed region
            n_chk = min(steps, rring.slots)
            t_end = time.perf_counter() + 10.0
            got = [0] * rring.n_sensors
            while time.perf_counter() < t_end:
                got = [rring.complete(sid, steps - n_chk, n_chk) for sid in range(rring.n_sensors)]
                if all(g == n_chk for g in got):
                    break
            timed_native.ring_report = {"records_expected_per_sensor": n_chk, "records_received": got, "record_bytes": 512,
                                        "transport": "stores by a kernel of the producing GPU into rank 0's GPU memory (CUDA IPC mapping, NVLink peer access); no NCCL, no host hop"}
        ev1.record()
        barrier()
        ms = ev0.elapsed_time(ev1)
        clocks = sampler.result()
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        # host-side timeline of the timed call (per rank; max and median over ranks): where a short region spends its time
        tl = torch.tensor(seq.host_timeline(), device="cuda", dtype=torch.float64)
        if world > 1:
            allv = [torch.zeros_like(tl) for _ in range(world)]
            dist.all_gather(allv, tl)
            allv = torch.stack(allv).cpu().numpy()
        else:
            allv = tl.cpu().numpy()[None]
        timed_native.host_timeline = {"unit": "us since the entry of dms_sequence_run", "keys": ["first_k_track_enqueued", "last_result_collected", "returned"],
                                      "max_over_ranks": [float(v) for v in allv.max(0)], "median_over_ranks": [float(v) for v in np.median(allv, 0)]}
        ktimes = [r.track_ms for r in res]
        iters_seen = [list(r.r.iters) for r in res]
        so3_seen = [r.r.so3_iters for r in res]
        pose = np.array(res[steps - 1].pose[:], np.float32).reshape(4, 4)
        timed_native.last_result = res[steps - 1].r
        timed_native.host_ms = [float(r.host_ms) for r in res]
        return ms, ktimes, iters_seen, so3_seen, clocks, (pose[:3, 3].copy(), pose[:3, :3].copy())

    def timed(pipe, steps, warmup, host):
        hs, ss = pipe
        depth = len(hs)
        t_w = time.perf_counter()
        while True:   # see timed_native: W frames per tracker, for at least a quarter of a second
            for i in range(max(3, warmup) * depth):
                one_frame(hs[i % depth], ss[i % depth], i, host)
            torch.cuda.synchronize()
            if time.perf_counter() - t_w > 0.25:
                break
        barrier()
        sampler = ClockSampler(local_rank)
        sampler.start()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        stats = [{"k": [], "it": [], "so3": [], "last": None, "lat": [], "phase": []} for _ in range(depth)]

        # host-side ordering between the trackers' threads: issued[i] = initRGB of frame i has returned; taken[j] = the
        # image tracker j % depth holds from frame j - depth has been handed on, so frame j may load its own
        sync = None
        if depth > 1 and share_rgb:
            sync = {"hs": hs, "issued": [threading.Event() for _ in range(steps + 1)], "taken": [threading.Event() for _ in range(steps + depth + 1)]}
            for j in range(depth):
                sync["taken"][j].set()   # the first frame of every tracker has nothing to wait for

        def work(k):
            torch.cuda.set_device(local_rank)
            for i in range(k, steps, depth):
                t_host = time.perf_counter()
                out = one_frame(hs[k], ss[k], i, host, hs[(i - 1) % depth] if (sync is not None and i > 0) else None, sync)
                stats[k]["lat"].append((time.perf_counter() - t_host) * 1e3)   # host-visible latency of one frame (result in hand)
                tm = hs[k].timing()
                stats[k]["k"].append(tm["track"]); stats[k]["it"].append(list(hs[k].iters)); stats[k]["so3"].append(hs[k].so3_iters)
                stats[k]["phase"].append(tm)
                if i == steps - 1:
                    stats[k]["last"] = out
        ev0.record()
        if depth == 1:
            work(0)
        else:
            ths = [threading.Thread(target=work, args=(k,)) for k in range(depth)]
            [t.start() for t in ths]; [t.join() for t in ths]
        torch.cuda.synchronize()
        ev1.record()
        barrier()
        ms = ev0.elapsed_time(ev1)
        clocks = sampler.result()
        if world > 1:
            t = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        ktimes = [x for s_ in stats for x in s_["k"]]
        iters_seen = [x for s_ in stats for x in s_["it"]]
        so3_seen = [x for s_ in stats for x in s_["so3"]]
        last = [s_["last"] for s_ in stats if s_["last"] is not None][0]
        timed.latencies = [x for s_ in stats for x in s_["lat"]]
        timed.phases = [x for s_ in stats for x in s_["phase"]]
        return ms, ktimes, iters_seen, so3_seen, clocks, last

    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak = float(json.load(open(peaks_path))["hbm_gbs"]); peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak = 6650.0; peak_src = "fallback (B200_PROFILING.md)"

    def roofline_of(ms, steps, ktimes, iters_seen, so3_seen, depth):
        k_ms = float(np.mean(ktimes))
        alg = float(np.mean([algorithmic_bytes(cam, it, s3) for it, s3 in zip(iters_seen, so3_seen)]))
        # `depth` = tracking kernels resident at once, each on its share of the block slots (further trackers queue): the
        # aggregate rate of the kernel is depth launches per k_ms
        achieved = depth * alg / (k_ms * 1e-3) / 1e9
        return {"bound": "hbm", "kernel": "k_track (persistent cooperative: derivatives + SO3 + 3-level GN incl. solves)", "achieved": achieved,
                "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic_of(args.workload), "peak_source": peak_src,
                "frac_step_anchored": alg / ((ms / steps) * 1e-3) / 1e9 / peak,  # algorithmic bytes of one frame's k_track / wall time per frame: no concurrency factor
                "algorithmic_bytes_per_launch": alg, "kernel_ms": k_ms, "concurrent_launches": depth,
                "kernel_share_of_step": (k_ms / depth) / (ms / steps),
                "us_per_gn_iteration": 1000.0 * k_ms / max(1, int(np.mean([sum(it) for it in iters_seen])))}

    # trackers sharing the GPU: 20 (measured, tools/quick_perf.py at 300 frames: 12 x 24 blocks 5500 frames/s, 20 x 14 blocks
    # 5900, 28 x 10 5950), also in the driver's 20-step run, where every tracker then gets one frame (20 x 14: 5350 - 5440
    # frames/s at 20 frames, 10 x 29: 5150; 8 GPUs: 38 050 against 35 760 - 36 930)
    depth = max(1, args.pipeline) if args.pipeline else min(20, max(2, args.steps))
    depth_e2e = max(1, args.pipeline) if args.pipeline else min(6, max(2, args.steps // 5))   # uploads inside the region: PCIe-bound, fewer trackers start sooner
    _thr = args.threads if args.threads else 768
    _queues = int(os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS", "8"))
    _grid = args.grid if args.grid else (sms * max(1, 1024 // _thr)) // min(depth, _queues)
    config_exec["pipeline"] = (f"{depth} trackers share the GPU ({_grid} blocks of {_thr} threads each, own stream); frame f runs on tracker f % {depth}"
                          if depth > 1 else "single tracker, frames strictly one after another")
    native = depth > 1 and args.driver == "native" and share_rgb
    seq = None
    if native:
        from densemonoslam_b200.sequence import SequenceTracker
        seq = SequenceTracker(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, device=local_rank, n_trackers=depth, block_threads=args.threads,
                              grid=args.grid, wait_mode=args.wait, rgbOnly=int(FLAGS["rgbOnly"]), icpWeight=FLAGS["icpWeight"], pyramid=int(FLAGS["pyramid"]),
                              fastOdom=int(FLAGS["fastOdom"]), so3=int(FLAGS["so3"]), depthCutoff=25.0, timing=3)
        config_exec["driver"] = "dms_sequence_run (native host threads)"
        pipe = None
        # per-frame result exchange of the multi-sensor mode: one ring in rank 0's GPU memory, every rank's trackers publish into it
        from densemonoslam_b200.ring import ResultRing
        ring_slots = 1024
        from densemonoslam_b200 import multi
        if rank == 0:
            rring = ResultRing(world, ring_slots, device=local_rank)
            multi.share_ring_handle(rring.export())
        else:
            rring = ResultRing(world, ring_slots, device=local_rank, handle=multi.share_ring_handle(None))
        rring.attach(seq, rank)
        ms, ktimes, iters_seen, so3_seen, clocks, last = timed_native(seq, args.steps, args.warmup, False)
    else:
        config_exec["driver"] = "python threads over the per-call C ABI"
        pipe = make_pipe(depth)
        ms, ktimes, iters_seen, so3_seen, clocks, last = timed(pipe, args.steps, args.warmup, False)
    value = world * args.steps / (ms * 1e-3)
    if native:
        config_exec["host_timeline"] = timed_native.host_timeline
        result_ring = getattr(timed_native, "ring_report", None)
    gt = frames[(args.steps - 1) % ring]["gt"]
    pose_err = float(np.max(np.abs(last[0] - gt[:3, 3])))
    roofline = roofline_of(ms, args.steps, ktimes, iters_seen, so3_seen, min(depth, _queues) if not args.grid else min(depth, max(1, (sms * max(1, 1024 // _thr)) // args.grid)))

    if args.quick:
        if rank == 0:
            emit({"value": value, "roofline": roofline, "config": config, "config_exec": config_exec, "clocks": clocks})
        return
    e2e_steps = min(args.steps, 300)
    if native:
        if depth_e2e != depth:
            del seq
            seq = SequenceTracker(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, device=local_rank, n_trackers=depth_e2e, block_threads=args.threads,
                                  grid=args.grid, wait_mode=args.wait, rgbOnly=int(FLAGS["rgbOnly"]), icpWeight=FLAGS["icpWeight"], pyramid=int(FLAGS["pyramid"]),
                                  fastOdom=int(FLAGS["fastOdom"]), so3=int(FLAGS["so3"]), depthCutoff=25.0, timing=3)
        config_exec["pipeline_e2e"] = f"{depth_e2e} trackers, each with two input stages on the device and a copy stream (frame f + {depth_e2e} uploads under frame f's kernels)"
        ms_e2e, _, _, _, _, last_e2e = timed_native(seq, e2e_steps, min(args.warmup, 5), True)
    else:
        ms_e2e, _, _, _, _, last_e2e = timed(pipe, e2e_steps, min(args.warmup, 5), True)
    pose_err_e2e = float(np.max(np.abs(last_e2e[0] - frames[(e2e_steps - 1) % ring]["gt"][:3, 3])))
    e2e_value = world * e2e_steps / (ms_e2e * 1e-3)

    out = {"metric": "odometry_frames_per_s", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": config, "config_exec": config_exec, "roofline": roofline, "result_ring": result_ring if native else None,
           "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": per_frame_in + (0 if (share_rgb and depth > 1) else frames[0]["live"]["rgba"].nbytes),
                   "d2h_bytes_per_step": int(_lib.load().dms_track_d2h_bytes()), "steps": e2e_steps, "last_frame_abs_translation_error_m": pose_err_e2e,
                   "note": "host-buffer C-ABI entry points (pinned inputs), pose result read back every frame"},
           "gpu_launches": (6 * args.steps + 1) if native else 6 * args.steps,  # per frame: k_populate x2 (x3 when the previous image is reloaded), k_maps_pyramid, k_icp_depth, k_track, k_ring_publish
           "clocks": clocks, "last_frame_abs_translation_error_m": pose_err,
           "lib": os.path.relpath(_lib.LIB_PATH, ROOT)}
    if world > 1:
        # the only exchange of the multi-sensor mode: per-frame result records, gathered on rank 0 outside the timed region
        from densemonoslam_b200 import multi
        lr = timed_native.last_result if native else pipe[0][0]
        rec = multi.pack_record(last[0], last[1], [lr.lastICPError, lr.lastICPCount, lr.lastRGBError, lr.lastRGBCount, lr.lastSO3Error, lr.lastSO3Count])
        full = multi.gather_records({rank: rec[None]}, world, 1, device="cuda")
        if rank == 0:
            out["gathered_result_records"] = list(full.shape)
    if depth > 1:
        # Latency regime: the same sequence on ONE tracker that owns the GPU, one frame in flight (what a frame-to-model
        # caller sees, ElasticFusion.cpp:173-201: the model view of frame f+1 exists only after frame f is fused).  Every
        # rank measures it on its own GPU; rank 0 reports its own numbers.
        del pipe, seq
        n1 = min(args.steps, 300)
        if native:
            # the library's own runner with one tracker (frames strictly one after another, its host side in C++): what a
            # C / C++ caller of the library sees; per-frame host latency = first init call of the frame -> result in hand
            from densemonoslam_b200.sequence import SequenceTracker
            seq1 = SequenceTracker(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, device=local_rank, n_trackers=1, block_threads=0, grid=0, wait_mode=0,
                                   rgbOnly=int(FLAGS["rgbOnly"]), icpWeight=FLAGS["icpWeight"], pyramid=int(FLAGS["pyramid"]), fastOdom=int(FLAGS["fastOdom"]),
                                   so3=int(FLAGS["so3"]), depthCutoff=25.0, timing=3)
            ms1, k1, it1, s31, _, _ = timed_native(seq1, n1, args.warmup, False)
            lat = np.sort(np.array(timed_native.host_ms))
            track_us = float(np.mean(k1)) * 1e3
            del seq1
            pipe1 = make_pipe(1)
        else:
            pipe1 = make_pipe(1)
            ms1, k1, it1, s31, _, _ = timed(pipe1, n1, args.warmup, False)
            lat = np.sort(np.array(timed.latencies))
            track_us = float(np.mean([p_["track"] for p_ in timed.phases])) * 1e3
        # phase breakdown: a short second run on the instrumented build of the kernel (about 3 % slower), outside the numbers above
        pipe1[0][0].set_timing(1)
        timed(pipe1, min(n1, 40), 3, False)
        ph = {k: float(np.mean([p_[k] for p_ in timed.phases])) * 1e3 for k in timed.phases[0]}
        ph["track_instrumented"] = ph.get("track"); ph["track"] = track_us
        gn_its = max(1, int(np.mean([sum(it) for it in it1])))
        out["latency"] = {"value": world * n1 / (ms1 * 1e-3), "unit": "frames/s", "ms_per_frame": ms1 / n1, "frames": n1,
                          "mode": "one tracker per GPU, one frame in flight, inputs resident in HBM; driven by " + ("dms_sequence_run with one tracker" if native else "Python over the per-call C ABI"),
                          "host_latency_ms": {"p50": float(lat[len(lat) // 2]), "p99": float(lat[min(len(lat) - 1, int(len(lat) * 0.99))]),
                                              "max": float(lat[-1]), "frames": int(len(lat))},
                          "track_us": ph.get("track"), "phase_us": ph,
                          "us_per_gn_iteration": (ph.get("icpStep", 0.0) + ph.get("rgbStep", 0.0) + ph.get("computeRgbResidual", 0.0)) / gn_its,
                          "roofline": roofline_of(ms1, n1, k1, it1, s31, 1)}

    if rank == 0 and world == 1 and args.workload == "kitti":
        try:
            out["e2e_frontend"] = run_frontend_e2e(cam, min(args.steps, 100), torch.device("cuda", local_rank))
        except Exception as e:  # reported, never fatal for the contract line
            out["e2e_frontend"] = {"unavailable": f"{type(e).__name__}: {e}"}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        host_frames = frames[: min(ring, 4)]
        run_oracle_frames(cam, host_frames, 1, 30.0)
        done, secs, threads = run_oracle_frames(cam, host_frames, 400, 15.0)
        out["cpu_baseline"] = {"value": done / secs, "unit": "frames/s", "cores": threads, "kind": "port",
                               "sample": f"{done} frames of the same workload in {secs:.1f} s (OpenMP oracle, all host threads)"}
        done1, secs1, _ = run_oracle_frames(cam, host_frames, 40, 6.0, threads=1)
        out["cpu_baseline"]["single_thread_value"] = done1 / secs1
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        legacy = run_legacy_frames(cam, frames, 200, 10.0)
        if legacy:
            out["legacy_cuda"] = legacy
    if rank == 0:
        emit(out)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
