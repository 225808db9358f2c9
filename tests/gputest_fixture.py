"""The reference's real-data fixtures (elasticfusion/GPUTest/{1c,1d,2c,2d}.png, packed by tests/golden/make_gputest_fixture.py)
turned into tracker inputs with GPUTest's own recipe: K = 528/528/320/240 (GPUTest.cpp:150-152), depth / 5 -> millimetres
(:50-56), vertex / normal maps from forward differences, invalid -> (0, 0, 0, 1) (:69-129)."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def _rgba(rgb):
    a = np.full(rgb.shape[:2] + (4,), 255, np.uint8)
    a[..., :3] = rgb                     # RGBA internal format fed from a 3-byte RGB buffer: .x = first byte
    return np.ascontiguousarray(a)


def _vertices(raw, fx, fy, cx, cy):
    """loadVertices (GPUTest.cpp:77-129): float arithmetic in the order of getVertex; border rows / columns stay empty."""
    h, w = raw.shape
    f32 = np.float32
    d = raw.astype(f32) / f32(5000.0)
    col = np.arange(w, dtype=f32)[None, :].repeat(h, 0)
    row = np.arange(h, dtype=f32)[:, None].repeat(w, 1)
    vx = (col - f32(cx)) * d * (f32(1.0) / f32(fx))
    vy = (row - f32(cy)) * d * (f32(1.0) / f32(fy))
    v = np.stack([vx, vy, d], -1).astype(f32)
    ok = np.zeros((h, w), bool)
    r = raw
    ok[1:-1, 1:-1] = (r[1:-1, 1:-1] > 0) & (r[2:, 1:-1] > 0) & (r[1:-1, 2:] > 0) & (r[:-2, 1:-1] > 0) & (r[1:-1, :-2] > 0)
    del_x = np.zeros_like(v); del_y = np.zeros_like(v)
    del_x[:, :-1] = v[:, 1:] - v[:, :-1]
    del_y[:-1, :] = v[1:, :] - v[:-1, :]
    n = np.cross(del_x, del_y).astype(f32)
    nn = np.sqrt((n * n).sum(-1, dtype=f32)).astype(f32)
    with np.errstate(invalid="ignore", divide="ignore"):
        n = (n / nn[..., None]).astype(f32)
    v4 = np.zeros((h, w, 4), f32); n4 = np.zeros((h, w, 4), f32)
    v4[ok, :3] = v[ok]; n4[ok, :3] = n[ok]
    v4[1:-1, 1:-1, 3] = 1.0; n4[1:-1, 1:-1, 3] = 1.0
    return v4, n4


def load_pair(synth):
    """(cam, model, live): model = frame 1 as GPUTest loads it (vertex / normal maps from 1d.png, image 1c.png), live = frame 2
    (depth 2d.png / 5, image 2c.png; vertex maps of frame 2 for the callers that register map against map)."""
    z = np.load(os.path.join(HERE, "golden", "gputest_frames.npz"))
    cam = synth.Camera(640, 480, 528.0, 528.0, 320.0, 240.0)
    mv, mn = _vertices(z["depth1"], cam.fx, cam.fy, cam.cx, cam.cy)
    lv, ln = _vertices(z["depth2"], cam.fx, cam.fy, cam.cx, cam.cy)
    model = {"vmap4": mv, "nmap4": mn, "rgba": _rgba(z["rgb1"]), "depth": np.ascontiguousarray(z["depth1"] // 5).astype(np.uint16)}
    live = {"vmap4": lv, "nmap4": ln, "rgba": _rgba(z["rgb2"]), "depth": np.ascontiguousarray(z["depth2"] // 5).astype(np.uint16)}
    return cam, model, live
