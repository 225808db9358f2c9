"""Result ring (include/dmsodom/ring_c.h): two sequence runners (two sensors) publish every tracked frame into one ring on
the device; the consumer finds every record complete, in its slot, and identical to what the runner returned to its host."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_ring_receives_every_frame_of_two_sensors(synth):
    import torch
    from densemonoslam_b200.ring import ResultRing
    from densemonoslam_b200.sequence import SequenceTracker
    cam = synth.scaled_camera(synth.CAM_VGA, 2)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    n_frames, slots = 7, 4          # fewer slots than frames: the ring wraps
    ring = ResultRing(2, slots)
    results = {}
    for sensor, n_trackers in ((0, 3), (1, 1)):
        scene = synth.room_scene(40 + sensor); rng = np.random.RandomState(sensor); T = np.eye(4); fl = []
        for f in range(n_frames):
            T = T @ synth.pose_from(rng.randn(3) * 0.004, rng.randn(3) * 0.006)
            prior = (T @ synth.pose_from(rng.randn(3) * 0.004, rng.randn(3) * 0.008)).astype(np.float32)
            m, l = synth.render(scene, cam, prior, seed=2 * f, hole_frac=0.0), synth.render(scene, cam, T, seed=2 * f + 1)
            fl.append({"v": dev(m["vmap4"]), "n": dev(m["nmap4"]), "mi": dev(m["rgba"]), "d": dev(l["depth"]), "li": dev(l["rgba"]), "prior": prior})
        seq = SequenceTracker(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, n_trackers=n_trackers, block_threads=512, grid=20, wait_mode=2)
        ring.attach(seq, sensor)
        results[sensor] = seq.run(seq.make_frames(fl), dev(synth.render(scene, cam, np.eye(4), seed=99)["rgba"]), False)
        torch.cuda.synchronize()
        del seq
    for sensor in (0, 1):
        assert ring.complete(sensor, 0, n_frames) == slots            # only the last `slots` frames are still in the ring
        assert ring.complete(sensor, n_frames - slots, slots) == slots
        for frame in range(n_frames - slots, n_frames):
            rec, res = ring.read(sensor, frame), results[sensor][frame]
            pose = np.array(res.pose[:], np.float32).reshape(4, 4)
            assert rec["sensor"] == sensor and rec["frame"] == frame
            assert np.array_equal(rec["R"], pose[:3, :3]) and np.array_equal(rec["t"], pose[:3, 3])
            assert np.array_equal(rec["stats"], np.array([res.r.lastICPError, res.r.lastICPCount, res.r.lastRGBError, res.r.lastRGBCount,
                                                          res.r.lastSO3Error, res.r.lastSO3Count], np.float32), equal_nan=True)
            assert np.array_equal(rec["A"].reshape(36), np.array(res.r.lastA[:])) and np.array_equal(rec["b"], np.array(res.r.lastb[:]))
            assert rec["iters"] == list(res.r.iters)
    with pytest.raises(Exception):
        ring.read(0, 0)   # lapped
