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
