"""Edge cases of the path on the GPU, each against the CPU oracle: ragged sizes, empty inputs, pitched inputs,
rotated poses, the 0.3 m guard, shared-GPU trackers."""
import threading

import numpy as np
import pytest

from conftest import assert_f32_equal
from test_gpu_parity import FLAG_ROWS, both, check_pyramids, compare_track, load_frame_to_model, to_dev

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available()
    return torch


@pytest.mark.parametrize("wh", [(323, 241), (97, 65), (64, 48)])
def test_ragged_sizes(torch_cuda, orc, synth, wh):
    """Widths / heights that are not multiples of 4, 32 or the pyramid factor (levels W>>1, W>>2)."""
    w, h = wh
    cam = synth.Camera(w, h, 0.8 * w, 0.8 * w, w / 2 - 0.5, h / 2 + 0.25)
    model, live, T = synth.frame_pair(cam, seed=3, rvec=(0.004, -0.006, 0.003), tvec=(0.008, -0.004, 0.006))
    g, o = both(orc, synth, cam)
    pose = synth.pose_from((0.3, -0.2, 0.1), (0.5, -0.25, 1.0)).astype(np.float32)
    load_frame_to_model(torch_cuda, g, o, model, live, pose)
    g.initFirstRGB(to_dev(torch_cuda, model["rgba"])); o.initFirstRGB(model["rgba"])
    check_pyramids(g, o)
    compare_track(g, o, pose[:3, 3].copy(), pose[:3, :3].copy(), FLAG_ROWS["frameToModel"])


def test_empty_inputs_return_the_prior_pose(torch_cuda, orc, synth, pair_small):
    """No depth, empty model: zero correspondences everywhere, A = 0 and b = 0.  Eigen's ldlt().solve (RO.cpp:554-566)
    zeroes the components of vanishing pivots, so the step is zero and the prior pose comes back (not NaN); the error
    fields are 0/0 = NaN as in the reference."""
    cam, model, live, T = pair_small
    g, o = both(orc, synth, cam)
    m = {k: np.zeros_like(v) for k, v in model.items()}
    l = {k: np.zeros_like(v) for k, v in live.items()}
    pose = synth.pose_from((0.1, -0.2, 0.05), (0.3, 0.1, -0.2)).astype(np.float32)
    load_frame_to_model(torch_cuda, g, o, m, l, pose)
    check_pyramids(g, o)
    tg, Rg = g.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], *FLAG_ROWS["modelToModel"])
    to_, Ro, res = o.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], *FLAG_ROWS["modelToModel"])
    assert g.lastICPCount == res.lastICPCount == 0 and g.lastRGBCount == res.lastRGBCount == 0
    assert np.array_equal(tg, to_) and np.array_equal(Rg, Ro)
    assert np.all(np.isfinite(tg)) and np.all(np.isfinite(Rg))
    assert np.max(np.abs(tg - pose[:3, 3])) <= 1e-6 and np.max(np.abs(Rg - pose[:3, :3])) <= 1e-6
    assert np.array_equal(g.lastA, np.zeros((6, 6))) and np.array_equal(np.array(res.A[:]), np.zeros(36))


def test_pitched_device_inputs(torch_cuda, orc, synth, pair_small):
    """Row pitch larger than the row (views into wider tensors), as a mapped texture can have."""
    torch = torch_cuda
    cam, model, live, T = pair_small
    g, o = both(orc, synth, cam)

    def pitched(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
        wide = torch.zeros((t.shape[0], t.shape[1] + 24) + tuple(t.shape[2:]), dtype=t.dtype, device="cuda")
        wide[:, : t.shape[1]] = t
        return wide[:, : t.shape[1]]
    pose = np.eye(4, dtype=np.float32)
    g.initICPModel(pitched(model["vmap4"]), pitched(model["nmap4"]), 25.0, pose); g.initRGBModel(pitched(model["rgba"]))
    g.initICP(pitched(live["depth"]), 25.0); g.initRGB(pitched(live["rgba"]))
    o.initICPModel(model["vmap4"], model["nmap4"], pose); o.initRGBModel(model["rgba"]); o.initICP(live["depth"], 25.0); o.initRGB(live["rgba"])
    check_pyramids(g, o)


def test_translation_guard(torch_cuda, orc, synth):
    """RO.cpp:589-593: an update of more than 0.3 m with an RGB term returns the input pose."""
    cam = synth.scaled_camera(synth.CAM_VGA, 4)
    model, live, T = synth.frame_pair(cam, seed=3, tvec=(0.9, 0.0, 0.0), rvec=(0, 0, 0))
    g, o = both(orc, synth, cam)
    pose = np.eye(4, dtype=np.float32)
    load_frame_to_model(torch_cuda, g, o, model, live, pose)
    t, R = compare_track(g, o, pose[:3, 3], pose[:3, :3], FLAG_ROWS["modelToModel"])
    assert np.linalg.norm(t) <= 0.3 + 1e-6


def test_depth_cutoff_and_thresholds(torch_cuda, orc, synth, pair_small):
    """Non-default distThresh and a depth cutoff inside the scene."""
    cam, model, live, T = pair_small
    g, o = both(orc, synth, cam, distThresh=0.03)
    pose = np.eye(4, dtype=np.float32)
    cv = lambda a: to_dev(torch_cuda, a)
    g.initICPModel(cv(model["vmap4"]), cv(model["nmap4"]), 25.0, pose); o.initICPModel(model["vmap4"], model["nmap4"], pose)
    g.initRGBModel(cv(model["rgba"])); o.initRGBModel(model["rgba"])
    g.initICP(cv(live["depth"]), 3.0); o.initICP(live["depth"], 3.0)
    g.initRGB(cv(live["rgba"])); o.initRGB(live["rgba"])
    check_pyramids(g, o)
    compare_track(g, o, pose[:3, 3], pose[:3, :3], FLAG_ROWS["modelToModel"])


def test_trackers_sharing_the_gpu_are_independent_and_exact(torch_cuda, orc, synth, pair_small):
    """Two handles, own streams, each limited to half the SMs, driven from two host threads at once:
    every result equals the oracle's (the reduction does not depend on the grid)."""
    torch = torch_cuda
    cam, model, live, T = pair_small
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    pose = np.eye(4, dtype=np.float32)
    o = orc.OracleOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
    o.initFirstRGB(model["rgba"]); o.initICPModel(model["vmap4"], model["nmap4"], pose); o.initRGBModel(model["rgba"]); o.initICP(live["depth"], 25.0); o.initRGB(live["rgba"])
    to_, Ro, res = o.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], *FLAG_ROWS["frameToModel"])
    from densemonoslam_b200.odometry import RGBDOdometry
    dev = {k: to_dev(torch, v) for k, v in (("v", model["vmap4"]), ("n", model["nmap4"]), ("mi", model["rgba"]), ("d", live["depth"]), ("li", live["rgba"]))}
    outs = [None, None]

    def work(k):
        h = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
        h.set_grid(sms // 2)
        s = torch.cuda.Stream()
        for _ in range(5):
            h.initFirstRGB(dev["mi"], stream=s); h.initICPModel(dev["v"], dev["n"], 25.0, pose, stream=s); h.initRGBModel(dev["mi"], stream=s)
            h.initICP(dev["d"], 25.0, stream=s); h.initRGB(dev["li"], stream=s)
            outs[k] = h.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], *FLAG_ROWS["frameToModel"], stream=s) + (h.lastICPCount, h.lastRGBCount)
    ths = [threading.Thread(target=work, args=(k,)) for k in range(2)]
    [t.start() for t in ths]; [t.join() for t in ths]
    for t, R, ci, cr in outs:
        assert np.array_equal(t, to_) and np.array_equal(R, Ro) and ci == res.lastICPCount and cr == res.lastRGBCount


@pytest.mark.parametrize("camname", ["vga", "kitti"])
def test_depth_prefilter_bit_exact(torch_cuda, orc, synth, camname):
    """N2: bilateral depth filter + metricise (depth_bilateral.frag:31-75, depth_metric.frag:29-39), the step before initICP(depth)."""
    from densemonoslam_b200.odometry import filterDepth, metriciseDepth
    cam = synth.CAM_VGA if camname == "vga" else synth.CAM_KITTI
    scene = synth.street_scene(5) if camname == "kitti" else None
    model, live, T = synth.frame_pair(cam, seed=31, scene=scene, max_depth=60.0, depth_noise=0.01)
    raw = live["depth"].copy()
    raw[::37, ::29] = 120     # below the 300 mm validity floor
    maxD = 8.0 if camname == "kitti" else 3.5
    d = to_dev(torch_cuda, raw)
    assert np.array_equal(filterDepth(d, maxD).cpu().numpy(), orc.depth_bilateral(raw, maxD))
    assert_f32_equal(metriciseDepth(d, maxD).cpu().numpy(), orc.depth_metric(raw, maxD), "metric depth")


def test_oversubscribed_trackers_serialise_and_agree(torch_cuda, synth):
    """More co-operative grids than the GPU can hold at once (3 x the full grid; mixed block sizes): the launches queue
    instead of dead-locking at their grid barriers, and every tracker returns the same bits."""
    import threading
    from densemonoslam_b200.odometry import RGBDOdometry
    cam = synth.scaled_camera(synth.CAM_VGA, 2)
    model, live, T = synth.frame_pair(cam, seed=3)
    D = {k: to_dev(torch_cuda, a) for k, a in (("v", model["vmap4"]), ("n", model["nmap4"]), ("mi", model["rgba"]), ("d", live["depth"]), ("li", live["rgba"]))}
    pose = np.eye(4, dtype=np.float32)

    def work(res, k, grid, threads):
        h = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
        if threads:
            h.set_block(threads)
        h.set_grid(grid)
        s = torch_cuda.cuda.Stream()
        h.initFirstRGB(D["mi"], stream=s)
        for _ in range(10):
            h.initICPModel(D["v"], D["n"], 25.0, pose, stream=s); h.initRGBModel(D["mi"], stream=s); h.initICP(D["d"], 25.0, stream=s); h.initRGB(D["li"], stream=s)
            t, R = h.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], False, 10.0, True, False, True, stream=s)
        res[k] = (t, R, h.lastICPCount, h.lastRGBCount)

    for cfgs in ([(0, 0)] * 3, [(0, 0), (148, 512), (60, 256), (0, 512)]):
        res = {}
        ths = [threading.Thread(target=work, args=(res, k, g, th)) for k, (g, th) in enumerate(cfgs)]
        [t.start() for t in ths]
        [t.join(timeout=120) for t in ths]
        assert len(res) == len(cfgs), "a tracker did not finish"
        for r in res.values():
            assert np.array_equal(r[0], res[0][0]) and np.array_equal(r[1], res[0][1]) and r[2:] == res[0][2:]


def test_first_rgb_from_another_tracker(torch_cuda, synth):
    """dms_odom_init_first_rgb_from: the SO(3) reference image taken from the pyramid of the tracker that processed the
    previous frame gives the same track, bit for bit, as loading that image with initFirstRGB."""
    from densemonoslam_b200.odometry import RGBDOdometry
    cam = synth.scaled_camera(synth.CAM_VGA, 2)
    m0, l0, _ = synth.frame_pair(cam, seed=5)
    m1, l1, _ = synth.frame_pair(cam, seed=6, rvec=(0.004, 0.002, -0.003), tvec=(0.006, 0.004, -0.002))
    pose = np.eye(4, dtype=np.float32)
    d = lambda a: to_dev(torch_cuda, a)

    def frame(h, m, l):
        h.initICPModel(d(m["vmap4"]), d(m["nmap4"]), 25.0, pose); h.initRGBModel(d(m["rgba"])); h.initICP(d(l["depth"]), 25.0); h.initRGB(d(l["rgba"]))

    a, b, c = (RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy) for _ in range(3))
    # tracker a processes frame 0 (and tracks, which swaps its image pyramids); b and c process frame 1
    a.initFirstRGB(d(m0["rgba"])); frame(a, m0, l0)
    a.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], False, 10.0, True, False, True)
    frame(b, m1, l1); b.initFirstRGBFrom(a)
    frame(c, m1, l1); c.initFirstRGB(d(l0["rgba"]))
    assert np.array_equal(b.download("lastNextImage", 0), c.download("lastNextImage", 0))
    assert np.array_equal(b.download("lastNextImage", 2), c.download("lastNextImage", 2))
    tb, Rb = b.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], False, 10.0, True, False, True)
    tc, Rc = c.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], False, 10.0, True, False, True)
    assert np.array_equal(tb, tc) and np.array_equal(Rb, Rc) and b.lastSO3Count == c.lastSO3Count


@pytest.mark.parametrize("grid,threads", [(8, 512), (37, 512), (5, 1024)])
def test_full_track_parity_in_the_shared_gpu_launch_shape(torch_cuda, orc, synth, grid, threads):
    """The launch shapes trackers use when they share the GPU (few blocks, 512 threads): every warp walks many segments and
    phase B takes the photometric work list (sparse terms) -- integer outputs bit-exact at every iteration, as with the
    full grid."""
    cam = synth.CAM_KITTI
    frames = synth.kitti_sequence(cam, 2, seed=2021)
    g, o = both(orc, synth, cam)
    g.set_block(threads); g.set_grid(grid)
    g.initFirstRGB(to_dev(torch_cuda, frames[0]["live"]["rgba"])); o.initFirstRGB(frames[0]["live"]["rgba"])
    for f in frames:
        prior = f["prior"]
        load_frame_to_model(torch_cuda, g, o, f["model"], f["live"], prior)
        compare_track(g, o, prior[:3, 3].copy(), prior[:3, :3].copy(), FLAG_ROWS["frameToModel"])


@pytest.mark.parametrize("row", ["frameToModel", "frameToModel_fastOdom_nopyr", "modelToModel", "rgbOnly", "icpOnly_w100", "referenceFrame"])
@pytest.mark.parametrize("shape", [(0, 0), (14, 512)])
def test_production_build_of_the_tracking_kernel(torch_cuda, orc, synth, row, shape):
    """The parity tests above run the instrumented build of k_track (they read its per-iteration trace).  The build a
    caller gets -- no trace, timing off or events only -- is a separate instantiation of the same body: its final pose,
    counts, errors and last system must equal the oracle's as well, alone on the GPU and in the launch shape of trackers
    that share it."""
    from densemonoslam_b200.odometry import RGBDOdometry, DEFAULT_ANGLE_THRESH
    cam = synth.CAM_KITTI if shape[0] else synth.CAM_VGA
    frames = synth.kitti_sequence(cam, 2, seed=2021)
    o = orc.OracleOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, 0.10, DEFAULT_ANGLE_THRESH)
    for mode in (0, 3):
        g = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy, 0.10, DEFAULT_ANGLE_THRESH)
        g.set_timing(mode)
        if shape[0]:
            g.set_block(shape[1]); g.set_grid(shape[0])
        g.initFirstRGB(to_dev(torch_cuda, frames[0]["live"]["rgba"])); o.initFirstRGB(frames[0]["live"]["rgba"])
        for f in frames:
            prior = f["prior"]
            load_frame_to_model(torch_cuda, g, o, f["model"], f["live"], prior)
            compare_track(g, o, prior[:3, 3].copy(), prior[:3, :3].copy(), FLAG_ROWS[row], check_trace=False)
            if mode == 3:
                tm = g.timing()
                assert tm["track"] > 0 and tm["icpStep"] == 0 and tm["rgbStep"] == 0   # events only: no phase counters
