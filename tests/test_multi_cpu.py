"""N > 1 host logic on CPU: world_size-2 gloo, sensors dealt round-robin, per-frame records gathered on rank 0.
The tracker is a stand-in (the CPU oracle) -- only the sharding / gather logic is under test here."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

from conftest import ROOT

FLAGS = (False, 10.0, True, False, True, False)
N_SENSORS, N_FRAMES = 3, 2


def _sensor_frames(sid):
    from densemonoslam_b200 import synth
    cam = synth.scaled_camera(synth.CAM_VGA, 8)
    frames = []
    for f in range(N_FRAMES):
        model, live, T = synth.frame_pair(cam, seed=100 * sid + f, rvec=(0.002, -0.003, 0.001), tvec=(0.004, -0.002, 0.003))
        frames.append({"model": model, "live": live, "prior": np.eye(4, dtype=np.float32)})
    return cam, frames


def _make_tracker(sid):
    from densemonoslam_b200 import synth
    from oracle import oracle as orc
    cam = synth.scaled_camera(synth.CAM_VGA, 8)
    return orc.OracleOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)


def _worker(rank, world, port, outfile):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from densemonoslam_b200 import multi
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = multi.assign_sensors(N_SENSORS, world, rank)
    local = multi.run_sensors(_make_tracker, {sid: _sensor_frames(sid)[1] for sid in mine}, FLAGS)
    full = multi.gather_records(local, N_SENSORS, N_FRAMES)
    # set-up of the result ring: rank 0's 64-byte handle reaches every rank unchanged
    handle = bytes(range(64))
    got = multi.share_ring_handle(handle if rank == 0 else None)
    assert got == handle
    if rank == 0:
        np.save(outfile, full)
    dist.barrier()
    dist.destroy_process_group()


def test_assign_sensors_layout():
    from densemonoslam_b200 import multi
    for n in (0, 1, 3, 8, 11):
        for w in (1, 2, 4, 8):
            got = sorted(s for r in range(w) for s in multi.assign_sensors(n, w, r))
            assert got == list(range(n))
            sizes = [len(multi.assign_sensors(n, w, r)) for r in range(w)]
            assert max(sizes) - min(sizes) <= 1


def test_two_rank_gloo_gather_equals_single_process(tmp_path):
    from densemonoslam_b200 import multi
    out = str(tmp_path / "gathered.npy")
    mp.spawn(_worker, args=(2, 29517, out), nprocs=2, join=True)
    full = np.load(out)
    assert full.shape == (N_SENSORS, N_FRAMES, multi.RECORD_FLOATS) and not np.isnan(full[:, :, :12]).any()
    ref = multi.run_sensors(_make_tracker, {sid: _sensor_frames(sid)[1] for sid in range(N_SENSORS)}, FLAGS)
    for sid in range(N_SENSORS):
        assert np.array_equal(full[sid], ref[sid], equal_nan=True), sid
