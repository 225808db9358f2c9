"""Multi-GPU layout of the tracker: the path does not shard inside a frame (a 29-float all-reduce per
~10 us Gauss-Newton iteration would cost more than the iteration), it shards across sensors -- the
collaborative mode of the reference (MainController.cpp:226-233, one Context with its own trackers per
camera).  One process per GPU (torchrun), sensors dealt round-robin to ranks, no collective on the data
path; the only exchange is the small per-frame result record: written device-to-device into a ring in rank 0's GPU memory
(densemonoslam_b200/ring.py; set up with share_ring_handle below), or gathered after the run (gather_records).

Works with any torch.distributed backend (tests use gloo on CPU with a stand-in tracker).
"""
from __future__ import annotations

import numpy as np

RECORD_FLOATS = 3 + 9 + 6  # translation, rotation (row-major), lastICPError/Count, lastRGBError/Count, lastSO3Error/Count


def assign_sensors(n_sensors: int, world_size: int, rank: int) -> list[int]:
    """Sensors owned by `rank`: round-robin, so any n_sensors >= 1 and world_size >= 1 is balanced to within one."""
    if n_sensors < 0 or world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad sensor / rank layout")
    return list(range(rank, n_sensors, world_size))


def pack_record(t, R, fields) -> np.ndarray:
    return np.concatenate([np.asarray(t, np.float32).reshape(3), np.asarray(R, np.float32).reshape(9), np.asarray(fields, np.float32).reshape(6)])


def gather_records(local: dict[int, np.ndarray], n_sensors: int, n_frames: int, device="cpu"):
    """local: {sensor id: [n_frames, RECORD_FLOATS]} of this rank.  Returns on rank 0 the full
    [n_sensors, n_frames, RECORD_FLOATS] array (None elsewhere).  Uses one all_gather of a fixed-size tensor."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size() if dist.is_initialized() else 1
    rank = dist.get_rank() if dist.is_initialized() else 0
    per_rank = (n_sensors + world - 1) // world
    buf = torch.full((per_rank, n_frames, RECORD_FLOATS), float("nan"), dtype=torch.float32, device=device)
    for slot, sid in enumerate(assign_sensors(n_sensors, world, rank)):
        buf[slot] = torch.from_numpy(np.asarray(local[sid], np.float32)).to(device)
    if world == 1:
        allb = [buf]
    else:
        allb = [torch.empty_like(buf) for _ in range(world)]
        dist.all_gather(allb, buf)
    if rank != 0:
        return None
    out = np.full((n_sensors, n_frames, RECORD_FLOATS), np.nan, np.float32)
    for r in range(world):
        for slot, sid in enumerate(assign_sensors(n_sensors, world, r)):
            out[sid] = allb[r][slot].cpu().numpy()
    return out


def run_sensors(make_tracker, sensor_frames: dict[int, list], flags) -> dict[int, np.ndarray]:
    """Tracks every frame of every sensor owned by this rank; `make_tracker(sensor_id)` returns an object with the
    RGBDOdometry call surface.  Returns {sensor id: [n_frames, RECORD_FLOATS]}."""
    out = {}
    for sid, frames in sensor_frames.items():
        trk = make_tracker(sid)
        trk.initFirstRGB(frames[0]["live"]["rgba"])
        recs = []
        for f in frames:
            p = f["prior"]
            trk.initICPModel(f["model"]["vmap4"], f["model"]["nmap4"], 25.0, p) if _takes_cutoff(trk) else trk.initICPModel(f["model"]["vmap4"], f["model"]["nmap4"], p)
            trk.initRGBModel(f["model"]["rgba"])
            trk.initICP(f["live"]["depth"], 25.0)
            trk.initRGB(f["live"]["rgba"])
            res = trk.getIncrementalTransformation(p[:3, 3].copy(), p[:3, :3].copy(), *flags)
            t, R = res[0], res[1]
            fields = _fields(trk, res)
            recs.append(pack_record(t, R, fields))
        out[sid] = np.stack(recs)
    return out


def _takes_cutoff(trk) -> bool:
    return trk.__class__.__name__ == "RGBDOdometry"


def _fields(trk, res):
    if len(res) > 2:  # a tracker stand-in injected by the CPU tests returns (t, R, result struct)
        r = res[2]
        return [r.lastICPError, r.lastICPCount, r.lastRGBError, r.lastRGBCount, r.lastSO3Error, r.lastSO3Count]
    return [trk.lastICPError, trk.lastICPCount, trk.lastRGBError, trk.lastRGBCount, trk.lastSO3Error, trk.lastSO3Count]


def share_ring_handle(handle: bytes | None) -> bytes:
    """The result ring lives in rank 0's GPU memory: rank 0 passes its 64-byte CUDA IPC handle (ResultRing.export()), every
    rank gets it back (the producers map the ring with it).  One broadcast of 64 bytes at set-up, nothing per frame."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        if handle is None:
            raise ValueError("the only rank must own the ring")
        return bytes(handle)
    if (dist.get_rank() == 0) != (handle is not None):
        raise ValueError("exactly rank 0 passes a handle")
    box = [bytes(handle) if handle is not None else None]
    dist.broadcast_object_list(box, src=0)
    if not isinstance(box[0], (bytes, bytearray)) or len(box[0]) != 64:
        raise ValueError("bad ring handle")
    return bytes(box[0])
