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


def test_image_hand_off_between_trackers_under_stress():
    """The SO(3) reference image of frame f travels on the device from the tracker of frame f-1 (dms_odom_init_first_rgb_from)
    while that tracker is already loading its next frame.  compute-sanitizer's racecheck is not available on this pool, so the
    hand-off is hammered instead: a tiny frame (kernels of a few microseconds, so host-side ordering mistakes would show), many
    frames, several tracker counts, host and device inputs, the same runner used for repeated runs - every frame's SO(3) count and
    pose must equal the strictly sequential run, which they only do if every tracker saw exactly the previous frame's image."""
    import torch
    from densemonoslam_b200 import synth
    from densemonoslam_b200.odometry import RGBDOdometry
    from densemonoslam_b200.sequence import SequenceTracker
    cam = synth.scaled_camera(synth.CAM_VGA, 8)   # 80 x 60
    scene = synth.room_scene(9)
    rng = np.random.RandomState(3)
    T = np.eye(4); frames = []
    for f in range(48):
        T = T @ synth.pose_from(rng.randn(3) * 0.004, rng.randn(3) * 0.006)
        prior = (T @ synth.pose_from(rng.randn(3) * 0.003, rng.randn(3) * 0.005)).astype(np.float32)
        frames.append({"model": synth.render(scene, cam, prior, seed=2 * f, hole_frac=0.0), "live": synth.render(scene, cam, T, seed=2 * f + 1), "prior": prior})
    first_prev = synth.render(scene, cam, np.eye(4), seed=99)["rgba"]
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    g = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
    g.initFirstRGB(dev(first_prev))
    want = []
    for f in frames:
        p = f["prior"]
        g.initICPModel(dev(f["model"]["vmap4"]), dev(f["model"]["nmap4"]), 25.0, p); g.initRGBModel(dev(f["model"]["rgba"]))
        g.initICP(dev(f["live"]["depth"]), 25.0); g.initRGB(dev(f["live"]["rgba"]))
        t, R = g.getIncrementalTransformation(p[:3, 3], p[:3, :3], False, 10.0, True, False, True)
        want.append((t.copy(), R.copy(), g.lastSO3Count, g.lastICPCount, g.lastRGBCount))
    for n_trackers in (2, 5, 8):
        for host in (False, True):
            put = (lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()) if host else dev
            fl = [{"v": put(f["model"]["vmap4"]), "n": put(f["model"]["nmap4"]), "mi": put(f["model"]["rgba"]), "d": put(f["live"]["depth"]),
                   "li": put(f["live"]["rgba"]), "prior": f["prior"]} for f in frames]
            seq = SequenceTracker(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, n_trackers=n_trackers, block_threads=512, grid=8, wait_mode=2)
            fc, fp = seq.make_frames(fl), put(first_prev)
            for rep in range(3):
                res = seq.run(fc, fp, host)
                for i, (r, w) in enumerate(zip(res, want)):
                    pose = np.array(r.pose[:], np.float32).reshape(4, 4)
                    assert np.array_equal(pose[:3, 3], w[0]) and np.array_equal(pose[:3, :3], w[1]), (n_trackers, host, rep, i)
                    assert (r.r.lastSO3Count, r.r.lastICPCount, r.r.lastRGBCount) == w[2:], (n_trackers, host, rep, i)
            del seq
