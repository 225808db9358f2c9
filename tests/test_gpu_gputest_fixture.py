"""Real sensor data through the path: the reference's own fixtures elasticfusion/GPUTest/{1c,1d,2c,2d}.png (a desk scene,
17 % depth holes, depth quantised at /5) prepared with GPUTest's recipe (tests/gputest_fixture.py).

  * CUDA path vs CPU oracle, bit-exact, in the frame-to-model call order and in GPUTest's own call shape
    (initICPModel; initICP(depth); getIncrementalTransformation(false, 10, false, false, true) with NO RGB init, i.e. on
    zero-filled image pyramids: GPUTest.cpp:260-278);
  * oracle vs the reference's own kernels (oracle/_ref) stage by stage on the same data.
"""
import json
import os

import numpy as np
import pytest

import gputest_fixture as gf
import ref_compare as rc
from conftest import ROOT
from test_gpu_parity import FLAG_ROWS, both, check_pyramids, compare_track, load_frame_to_model, to_dev, torch_cuda  # noqa: F401

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def real_pair(synth):
    return gf.load_pair(synth)


def test_frame_to_model_on_real_frames(torch_cuda, orc, synth, real_pair):
    cam, model, live = real_pair
    g, o = both(orc, synth, cam)
    pose = np.eye(4, dtype=np.float32)
    load_frame_to_model(torch_cuda, g, o, model, live, pose)
    g.initFirstRGB(to_dev(torch_cuda, model["rgba"])); o.initFirstRGB(model["rgba"])
    check_pyramids(g, o)
    t, R = compare_track(g, o, pose[:3, 3], pose[:3, :3], FLAG_ROWS["frameToModel"])
    assert g.lastICPCount > 50000 and g.lastRGBCount > 5000          # a real desk scene: most of the frame takes part
    assert 1e-4 < np.linalg.norm(t) < 0.05                              # consecutive sensor frames: millimetres to centimetres


def test_gputest_call_shape_on_real_frames(torch_cuda, orc, synth, real_pair):
    cam, model, live = real_pair
    g, o = both(orc, synth, cam)
    pose = np.eye(4, dtype=np.float32)
    cv = lambda a: to_dev(torch_cuda, a)
    for _ in range(2):  # GPUTest repeats the same three calls (GPUTest.cpp:260-278)
        g.initICPModel(cv(model["vmap4"]), cv(model["nmap4"]), 20.0, pose); o.initICPModel(model["vmap4"], model["nmap4"], pose)
        g.initICP(cv(live["depth"]), 20.0); o.initICP(live["depth"], 20.0)
        compare_track(g, o, pose[:3, 3], pose[:3, :3], FLAG_ROWS["gputest"])
    assert g.iters == [10, 0, 0] and g.lastRGBCount == 0                 # image pyramids never initialised: no photometric term


def test_oracle_matches_reference_kernels_on_real_frames(orc, synth, real_pair):
    from oracle import ref_legacy
    if not ref_legacy.available():
        pytest.skip("oracle/_ref/libref_legacy.so not built (needs /root/reference at build time)")
    cam, model, live = real_pair
    c = rc.case_inputs_from(synth, cam, model, live)
    rep = rc.compare(rc.oracle_stages(orc, cam, c), rc.legacy_stages(ref_legacy, cam, c))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rep, open(os.path.join(ROOT, "gpurun_out", "legacy_parity_gputest_png.json"), "w"), indent=1)
    rc.check(rep, exact_counts=False)
