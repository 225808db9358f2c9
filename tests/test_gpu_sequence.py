"""The native sequence runner (include/dmsodom/sequence_c.h) against one tracker called frame by frame."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n_trackers,host", [(4, False), (1, False), (3, False), (1, True), (3, True), (6, True)])
def test_sequence_runner_equals_sequential_tracker(n_trackers, host):
    import torch
    from densemonoslam_b200 import synth
    from densemonoslam_b200.odometry import RGBDOdometry
    from densemonoslam_b200.sequence import SequenceTracker
    cam = synth.scaled_camera(synth.CAM_VGA, 2)
    scene = synth.room_scene(4)
    rng = np.random.RandomState(1)
    T = np.eye(4)
    frames = []
    for f in range(9):
        T = T @ synth.pose_from(rng.randn(3) * 0.004, rng.randn(3) * 0.006)
        prior = (T @ synth.pose_from(rng.randn(3) * 0.004, rng.randn(3) * 0.008)).astype(np.float32)
        frames.append({"model": synth.render(scene, cam, prior, seed=2 * f, hole_frac=0.0), "live": synth.render(scene, cam, T, seed=2 * f + 1), "prior": prior})
    first_prev = synth.render(scene, cam, np.eye(4), seed=99)["rgba"]
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    put = (lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()) if host else dev
    fl = [{"v": put(f["model"]["vmap4"]), "n": put(f["model"]["nmap4"]), "mi": put(f["model"]["rgba"]), "d": put(f["live"]["depth"]),
           "li": put(f["live"]["rgba"]), "prior": f["prior"]} for f in frames]
    seq = SequenceTracker(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, n_trackers=n_trackers, block_threads=512, grid=20, wait_mode=2, timing=1)
    res = seq.run(seq.make_frames(fl), put(first_prev), host)
    assert seq.resident == min(n_trackers, 296 // 20) or n_trackers == 1

    g = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
    g.initFirstRGB(dev(first_prev))
    for f, r in zip(frames, res):
        p = f["prior"]
        g.initICPModel(dev(f["model"]["vmap4"]), dev(f["model"]["nmap4"]), 25.0, p); g.initRGBModel(dev(f["model"]["rgba"]))
        g.initICP(dev(f["live"]["depth"]), 25.0); g.initRGB(dev(f["live"]["rgba"]))
        t, R = g.getIncrementalTransformation(p[:3, 3], p[:3, :3], False, 10.0, True, False, True)
        pose = np.array(r.pose[:], np.float32).reshape(4, 4)
        assert np.array_equal(pose[:3, 3], t) and np.array_equal(pose[:3, :3], R)
        assert (r.r.lastICPCount, r.r.lastRGBCount, r.r.lastSO3Count) == (g.lastICPCount, g.lastRGBCount, g.lastSO3Count)
        assert list(r.r.iters) == g.iters and r.track_ms > 0
