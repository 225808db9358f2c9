"""CPU-only: the oracle against golden vectors produced by the reference's own legacy kernels
(tests/golden/golden_legacy_small.npz, written by tests/test_legacy_ref.py on the B200 box)."""
import os

import numpy as np
import pytest

import ref_compare as rc
from conftest import ROOT

GOLDEN = os.path.join(ROOT, "tests", "golden", "golden_legacy_small.npz")


@pytest.mark.skipif(not os.path.exists(GOLDEN), reason="golden vectors not generated yet")
def test_oracle_matches_reference_golden_vectors(orc, synth):
    cam = synth.scaled_camera(synth.CAM_VGA, 4)
    c = rc.case_inputs(synth, cam, seed=7)
    o = rc.oracle_stages(orc, cam, c)
    g = dict(np.load(GOLDEN))
    rc.check(rc.compare(o, g))
