"""Packs the reference's only real-data fixtures -- elasticfusion/GPUTest/{1c,1d,2c,2d}.png, the inputs of its launch-shape
tuner (GPUTest.cpp:166-205) -- into tests/golden/gputest_frames.npz so that the GPU box (which has no /root/reference) can
read them.  Pixel data only, unchanged: 8-bit RGB and 16-bit depth as stored in the PNGs.

    python tests/golden/make_gputest_fixture.py [/root/reference/elasticfusion/GPUTest]
"""
import os
import sys

import numpy as np
from PIL import Image

src = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/elasticfusion/GPUTest"
out = {}
for name in ("1c", "2c"):
    a = np.asarray(Image.open(os.path.join(src, name + ".png")))
    assert a.shape == (480, 640, 3) and a.dtype == np.uint8, (name, a.shape, a.dtype)
    out["rgb" + name[0]] = a
for name in ("1d", "2d"):
    a = np.asarray(Image.open(os.path.join(src, name + ".png")))
    assert a.shape == (480, 640), (name, a.shape, a.dtype)
    out["depth" + name[0]] = a.astype(np.uint16)
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "gputest_frames.npz"), **out)
print({k: (v.shape, str(v.dtype), int(v.max())) for k, v in out.items()})
