"""Pins the CPU oracle against the reference's OWN kernels: Cuda/reduce.cu and Cuda/cudafuncs.cu compiled
unmodified from /root/reference into oracle/_ref/libref_legacy.so (oracle/ref_build.sh) and run here on the GPU.
Exact where the arithmetic is division-free (intensity, derivative images); bounded mismatch counts where the
legacy build's --prec-div=false / rsqrtf / --ftz approximations can flip a threshold or a truncation."""
import json
import os

import numpy as np
import pytest

import ref_compare as rc
from conftest import ROOT

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ref():
    from oracle import ref_legacy
    if not ref_legacy.available():
        pytest.skip("oracle/_ref/libref_legacy.so not built (needs /root/reference at build time)")
    assert ref_legacy.lib().ref_device_ok()
    return ref_legacy


@pytest.mark.parametrize("name", ["vga", "small"])
def test_oracle_matches_reference_kernels(orc, synth, ref, name):
    cam = synth.CAM_VGA if name == "vga" else synth.scaled_camera(synth.CAM_VGA, 4)
    c = rc.case_inputs(synth, cam, seed=1234 if name == "vga" else 7)
    o = rc.oracle_stages(orc, cam, c)
    r = rc.legacy_stages(ref, cam, c)
    rep = rc.compare(o, r)
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    json.dump(rep, open(os.path.join(out_dir, f"legacy_parity_{name}.json"), "w"), indent=1)
    if name == "small":  # golden vectors for the CPU-only suite (tests/golden/, made by this test on the GPU box)
        np.savez_compressed(os.path.join(out_dir, "golden_legacy_small.npz"), **{k: np.asarray(v) for k, v in r.items()})
    rc.check(rep)


def test_intensity_all_colours_match_reference(orc, ref):
    """All 2^24 colours: pins the FMA contraction order of bgr2IntensityKernel (cf.cu:651-654)."""
    v = np.arange(1 << 24, dtype=np.uint32)
    rgba = np.stack([v & 255, (v >> 8) & 255, (v >> 16) & 255, np.full_like(v, 255)], axis=-1).astype(np.uint8).reshape(4096, 4096, 4)
    assert np.array_equal(orc.bgr_to_intensity(rgba), ref.bgr_to_intensity(rgba))


def test_full_track_against_reference_loop(orc, synth, ref):
    """Whole getIncrementalTransformation: the reference's host loop over its own kernels vs the oracle.
    The legacy build's fast-math flips a few threshold pixels per iteration, so poses agree to ~1e-4, not bit-exactly."""
    cam = synth.CAM_VGA
    model, live, T = synth.frame_pair(cam, seed=1234)
    pose = np.eye(4, dtype=np.float32)
    o = orc.OracleOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
    r = ref.LegacyOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
    r.upload(0, v4=model["vmap4"], n4=model["nmap4"], rgba=model["rgba"])
    r.upload(1, rgba=live["rgba"], depth=live["depth"])
    o.initFirstRGB(model["rgba"]); r.initFirstRGB(0)
    o.initICPModel(model["vmap4"], model["nmap4"], pose); r.initICPModel(0, pose)
    o.initRGBModel(model["rgba"]); r.initRGBModel(0)
    o.initICP(live["depth"], 25.0); r.initICP(1, 25.0)
    o.initRGB(live["rgba"]); r.initRGB(1)
    flags = (False, 10.0, True, False, True, False)
    to, Ro, res = o.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], *flags)
    tr, Rr = r.getIncrementalTransformation(pose[:3, 3], pose[:3, :3], *flags)
    rep = {"t_oracle": to.tolist(), "t_reference": tr.tolist(), "max_dt": float(np.max(np.abs(to - tr))), "max_dR": float(np.max(np.abs(Ro - Rr))),
           "icp_count": [res.lastICPCount, float(r.out[1])], "rgb_count": [res.lastRGBCount, float(r.out[3])], "so3_count": [res.lastSO3Count, float(r.out[5])],
           "icp_error": [res.lastICPError, float(r.out[0])], "rgb_error": [res.lastRGBError, float(r.out[2])]}
    json.dump(rep, open(os.path.join(ROOT, "gpurun_out", "legacy_track_vga.json"), "w"), indent=1)
    assert rep["max_dt"] < 2e-4 and rep["max_dR"] < 2e-4, rep
    assert abs(res.lastICPCount - r.out[1]) <= 2e-3 * r.out[1], rep
    assert abs(res.lastRGBCount - r.out[3]) <= 5e-3 * max(r.out[3], 1), rep
    assert np.max(np.abs(tr - T[:3, 3])) < 5e-3


@pytest.mark.parametrize("nb_img,nb_depth", [(64, 500), (32, 100)])
def test_nid_restatement_against_reference_kernels(orc, synth, ref, nb_img, nb_depth):
    """N4: oracle/nid_ref.py vs the reference's computeNIDImg / computeNIDDepth (cf.cu:1513-1650, 1794-1915) on the same
    pyramids, in-range depths only (the reference's histogram has no bounds check)."""
    from oracle import nid_ref
    cam = synth.scaled_camera(synth.CAM_VGA, 2)
    scene = synth.room_scene(21)
    kf = synth.render(scene, cam, np.eye(4), seed=1, max_depth=50.0)
    old = synth.render(scene, cam, synth.pose_from((0.01, 0.0, 0.0), (0.03, 0.0, 0.0)), seed=2, max_depth=50.0)
    live = synth.render(scene, cam, synth.pose_from((0.02, -0.01, 0.0), (0.05, 0.02, 0.01)), seed=3, max_depth=50.0)
    cutoff = 3.5
    ik, dk = nid_ref.keyframe_pyramids(kf["vmap4"], kf["rgba"], cutoff, 1)
    iko, dko = nid_ref.keyframe_pyramids(old["vmap4"], old["rgba"], cutoff, 1)
    ic, dc = nid_ref.keyframe_pyramids(live["vmap4"], live["rgba"], cutoff, 1)
    n = ic.size
    max_mm = 6000.0   # every finite depth (< 3.5 m) stays inside the histogram
    r_img = ref.nid_img(ik, iko, dk, dko, ic, nb_img)
    r_dep = ref.nid_depth(dk, dko, dc, nb_depth, max_mm)
    o_img = float(nid_ref.nid_from_hist(nid_ref.hist_img(ik, iko, dk, dko, ic, nb_img), n))
    o_dep = float(nid_ref.nid_from_hist(nid_ref.hist_depth(dk, dko, dc, nb_depth, max_mm), n))
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    json.dump({"nid_img": {"reference": r_img, "oracle": o_img}, "nid_depth": {"reference": r_dep, "oracle": o_dep}},
              open(os.path.join(out_dir, f"legacy_nid_{nb_img}_{nb_depth}.json"), "w"), indent=1)
    # same integer histograms; host entropy loops in fp32 with glibc log2f on both sides
    assert abs(r_img - o_img) <= 2e-6 * abs(r_img), (r_img, o_img)
    assert abs(r_dep - o_dep) <= 2e-6 * abs(r_dep), (r_dep, o_dep)
