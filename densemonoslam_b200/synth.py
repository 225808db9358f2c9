"""Synthetic RGB-D / model-prediction inputs for the dense tracker (tests and bench).

The reference feeds `RGBDOdometry` from OpenGL textures: a surfel splat of the model
(vertex / normal / colour maps, `Shaders/combo_splat.frag:36-59`) and the bilateral-filtered
sensor depth (`Shaders/depth_bilateral.frag:31-75`).  There is no graphics stack here, so
this module ray-casts an analytic scene (planes, spheres, boxes) with a procedural texture
and emits exactly those buffer formats, headless:

* ``rgba``   uint8  [H, W, 4]   colour image (alpha 255), 0 = empty
* ``depth``  uint16 [H, W]      millimetres, 0 = hole            (R16UI,  `Context.h:158-169`)
* ``vmap4``  float32 [H, W, 4]  camera-frame (x, y, z, conf), z == 0 empty   (RGBA32F)
* ``nmap4``  float32 [H, W, 4]  camera-frame (nx, ny, nz, radius)

Vertex formula follows `combo_splat.frag` / `GPUTest.cpp:62-67`:
``v = ((u - cx) z / fx, (v - cy) z / fy, z)``.
"""
from __future__ import annotations

import dataclasses
import numpy as np


@dataclasses.dataclass
class Camera:
    width: int
    height: int
    fx: float
    fy: float
    cx: float
    cy: float


# GPUTest.cpp:150-152
CAM_VGA = Camera(640, 480, 528.0, 528.0, 320.0, 240.0)
# logs/kitti/KITTI_RGBD_template_params.yaml:9-12
CAM_KITTI = Camera(1241, 376, 718.856, 718.856, 607.1928, 185.2157)
CAM_1080P = Camera(1920, 1080, 1400.0, 1400.0, 960.0, 540.0)


def scaled_camera(cam: Camera, div: int) -> Camera:
    return Camera(cam.width // div, cam.height // div, cam.fx / div, cam.fy / div, cam.cx / div, cam.cy / div)


def rodrigues(r) -> np.ndarray:
    r = np.asarray(r, dtype=np.float64)
    th = np.linalg.norm(r)
    if th < 1e-15:
        return np.eye(3)
    k = r / th
    Kx = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.cos(th) * np.eye(3) + (1 - np.cos(th)) * np.outer(k, k) + np.sin(th) * Kx


def pose_from(rvec, tvec) -> np.ndarray:
    T = np.eye(4)
    T[:3, :3] = rodrigues(rvec)
    T[:3, 3] = np.asarray(tvec, dtype=np.float64)
    return T


class Scene:
    """Analytic scene: list of primitives in world coordinates."""

    def __init__(self, planes=(), spheres=(), boxes=(), tex_scale=6.0, seed=0):
        self.planes = [(np.asarray(n, float) / np.linalg.norm(n), float(d)) for n, d in planes]  # n.p = d
        self.spheres = [(np.asarray(c, float), float(r)) for c, r in spheres]
        self.boxes = [(np.asarray(lo, float), np.asarray(hi, float)) for lo, hi in boxes]
        self.tex_scale = tex_scale
        self.seed = seed

    def intersect(self, o, d, forward=None, max_range=None):
        """o [3], d [...,3] unit rays -> (t, normal) nearest hit, t=inf where none.
        `forward`/`max_range` let boxes behind the camera or out of range be skipped."""
        shape = d.shape[:-1]
        best_t = np.full(shape, np.inf)
        best_n = np.zeros(shape + (3,))
        for n, dist in self.planes:
            denom = d @ n
            with np.errstate(divide="ignore", invalid="ignore"):
                t = (dist - o @ n) / denom
            ok = (np.abs(denom) > 1e-9) & (t > 1e-3) & (t < best_t)
            best_t = np.where(ok, t, best_t)
            nn = np.where((denom > 0)[..., None], -n, n)
            best_n = np.where(ok[..., None], nn, best_n)
        for c, r in self.spheres:
            oc = o - c
            b = d @ oc
            cc = oc @ oc - r * r
            disc = b * b - cc
            with np.errstate(invalid="ignore"):
                sq = np.sqrt(np.maximum(disc, 0))
            t = -b - sq
            ok = (disc > 0) & (t > 1e-3) & (t < best_t)
            best_t = np.where(ok, t, best_t)
            p = o + d * np.where(ok, t, 0)[..., None]
            nn = (p - c) / r
            best_n = np.where(ok[..., None], nn, best_n)
        for lo, hi in self.boxes:
            if forward is not None:
                c = 0.5 * (lo + hi) - o
                rad = 0.5 * np.linalg.norm(hi - lo)
                along = float(c @ forward)
                if along + rad < 0 or (max_range is not None and along - rad > max_range):
                    continue
            with np.errstate(divide="ignore", invalid="ignore"):
                t0 = (lo - o) / d
                t1 = (hi - o) / d
            tn = np.minimum(t0, t1)
            tf = np.maximum(t0, t1)
            tnear = np.max(tn, axis=-1)
            tfar = np.min(tf, axis=-1)
            ok = (tnear < tfar) & (tnear > 1e-3) & (tnear < best_t)
            axis = np.argmax(tn, axis=-1)
            nn = np.zeros(shape + (3,))
            sgn = -np.sign(np.take_along_axis(d, axis[..., None], axis=-1))[..., 0]
            for a in range(3):
                nn[..., a] = np.where(axis == a, sgn, 0.0)
            best_t = np.where(ok, tnear, best_t)
            best_n = np.where(ok[..., None], nn, best_n)
        return best_t, best_n

    def texture(self, p):
        """Procedural colour in [0,1]^3 as a function of world position."""
        s = self.tex_scale
        x, y, z = p[..., 0] * s, p[..., 1] * s, p[..., 2] * s
        base = (
            0.5
            + 0.22 * np.sin(1.7 * x + 0.5 * np.sin(0.9 * y))
            + 0.16 * np.sin(2.3 * y + 1.1 * z + 0.3)
            + 0.12 * np.sin(3.1 * z - 1.3 * x + 1.7)
        )
        # hashed fine-grain noise (continuous in p so it is view consistent)
        h = np.sin(12.9898 * x + 78.233 * y + 37.719 * z) * 43758.5453
        noise = (h - np.floor(h)) - 0.5
        f = 8.0
        check = np.sign(np.sin(f * x * 0.31) * np.sin(f * y * 0.29) * np.sin(f * z * 0.33))
        g = base + 0.06 * noise + 0.07 * check
        r = g + 0.05 * np.sin(0.7 * x)
        b = g - 0.05 * np.sin(0.6 * z)
        return np.clip(np.stack([r, g, b], axis=-1), 0.0, 1.0)


def room_scene(seed=0) -> Scene:
    """Textured room: 5 planes + 2 spheres, depths ~0.8-4.5 m from the origin camera."""
    # y-down camera convention: floor at y=+1.2, ceiling y=-1.3, back wall z=4.2, walls x=-2.4 / x=2.6
    planes = [((0, 1, 0), 1.2), ((0, 1, 0), -1.3), ((0, 0, 1), 4.2), ((1, 0, 0), -2.4), ((1, 0, 0), 2.6)]
    spheres = [((-0.6, 0.5, 2.2), 0.55), ((0.9, -0.1, 3.0), 0.7)]
    return Scene(planes=planes, spheres=spheres, boxes=[((-1.9, 0.4, 2.9), (-1.1, 1.2, 3.8))], tex_scale=20.0, seed=seed)


def street_scene(seed=0) -> Scene:
    """KITTI-like street: ground plane, two walls, parked boxes on both sides; depths ~2-40 m."""
    planes = [((0, 1, 0), 1.65), ((1, 0, 0), -6.0), ((1, 0, 0), 6.5), ((0, 0, 1), 4000.0)]
    boxes = []
    rng = np.random.RandomState(seed + 17)
    for i in range(400):
        z0 = 2.0 + i * 4.5 + rng.uniform(0, 1.5)
        side = -1 if i % 2 == 0 else 1
        x0 = side * rng.uniform(1.8, 3.2)
        w, h, l = rng.uniform(1.2, 1.9), rng.uniform(1.3, 2.6), rng.uniform(2.5, 4.0)
        boxes.append(((x0 - w / 2, 1.65 - h, z0), (x0 + w / 2, 1.65, z0 + l)))
    return Scene(planes=planes, boxes=boxes, tex_scale=8.0, seed=seed)


def render(scene: Scene, cam: Camera, T_wc: np.ndarray, *, max_depth=50.0, seed=0, zero_frac=0.03, hole_frac=0.04,
           depth_noise=0.0):
    """Ray-cast `scene` from camera pose T_wc (camera -> world).  Returns dict of buffers."""
    H, W = cam.height, cam.width
    rng = np.random.RandomState(seed)
    u, v = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64))
    dirs_c = np.stack([(u - cam.cx) / cam.fx, (v - cam.cy) / cam.fy, np.ones_like(u)], axis=-1)
    zscale = np.linalg.norm(dirs_c, axis=-1)
    dirs_c_unit = dirs_c / zscale[..., None]
    R, t = T_wc[:3, :3], T_wc[:3, 3]
    d_w = dirs_c_unit @ R.T
    tt, n_w = scene.intersect(t, d_w, forward=R[:, 2], max_range=max_depth * 1.2)
    hit = np.isfinite(tt)
    tt_safe = np.where(hit, tt, 0.0)
    p_w = t + d_w * tt_safe[..., None]
    z = tt_safe / zscale  # camera-frame depth
    hit &= z < max_depth
    col = scene.texture(p_w)
    rgb = np.clip(np.rint(16 + col * (240 - 16)), 0, 255).astype(np.uint8)
    rgb[~hit] = 0
    if zero_frac > 0:
        rgb[rng.rand(H, W) < zero_frac] = 0
    rgba = np.concatenate([rgb, np.full((H, W, 1), 255, np.uint8)], axis=-1)

    z32 = z.astype(np.float32)
    n_c = (n_w @ R).astype(np.float32)  # R^T n
    vmap4 = np.zeros((H, W, 4), np.float32)
    fxi, fyi = np.float32(1.0 / cam.fx), np.float32(1.0 / cam.fy)
    vmap4[..., 0] = (u.astype(np.float32) - np.float32(cam.cx)) * z32 * fxi
    vmap4[..., 1] = (v.astype(np.float32) - np.float32(cam.cy)) * z32 * fyi
    vmap4[..., 2] = z32
    vmap4[..., 3] = 1.0
    nmap4 = np.zeros((H, W, 4), np.float32)
    nmap4[..., :3] = n_c
    nmap4[..., 3] = 0.005
    empty = ~hit
    vmap4[empty] = 0
    nmap4[empty] = 0

    zz = z.copy()
    if depth_noise > 0:
        # predicted-depth style error: a smooth multiplicative bias field plus a little white noise
        ph = rng.uniform(0, 2 * np.pi, 2)
        bias = np.sin(u * (2 * np.pi / W) * 1.5 + ph[0]) * np.sin(v * (2 * np.pi / H) * 1.1 + ph[1])
        zz = zz * (1.0 + depth_noise * bias + 0.1 * depth_noise * rng.randn(H, W))
    dmm = np.clip(np.rint(zz * 1000.0), 0, 65535)
    dmm[~hit] = 0
    if hole_frac > 0:
        dmm[rng.rand(H, W) < hole_frac] = 0
    depth = dmm.astype(np.uint16)
    return {"rgba": np.ascontiguousarray(rgba), "depth": depth, "vmap4": vmap4, "nmap4": nmap4,
            "pose": T_wc.astype(np.float32)}


def vertices_from_depth(depth_mm_raw: np.ndarray, cam: Camera, scale=5000.0):
    """GPUTest.cpp:69-129 loadVertices: vertex + forward-difference normal maps from a raw
    TUM-style depth PNG (values / `scale` = metres); invalid -> (0,0,0,1)."""
    H, W = depth_mm_raw.shape
    d = depth_mm_raw.astype(np.float32) / np.float32(scale)
    u, v = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32))
    fxi, fyi = np.float32(1.0) / np.float32(cam.fx), np.float32(1.0) / np.float32(cam.fy)
    P = np.stack([(u - np.float32(cam.cx)) * d * fxi, (v - np.float32(cam.cy)) * d * fyi, d], axis=-1)
    valid = np.zeros((H, W), bool)
    r = depth_mm_raw
    valid[1:-1, 1:-1] = (r[1:-1, 1:-1] > 0) & (r[2:, 1:-1] > 0) & (r[1:-1, 2:] > 0) & (r[:-2, 1:-1] > 0) & (r[1:-1, :-2] > 0)
    delx = np.zeros_like(P)
    dely = np.zeros_like(P)
    delx[:, :-1] = P[:, 1:] - P[:, :-1]
    dely[:-1, :] = P[1:, :] - P[:-1, :]
    n = np.cross(delx, dely)
    nn = np.linalg.norm(n, axis=-1, keepdims=True)
    with np.errstate(invalid="ignore", divide="ignore"):
        n = (n / nn).astype(np.float32)
    vmap4 = np.zeros((H, W, 4), np.float32)
    nmap4 = np.zeros((H, W, 4), np.float32)
    vmap4[..., 3] = 1.0
    nmap4[..., 3] = 1.0
    vmap4[valid, :3] = P[valid]
    nmap4[valid, :3] = n[valid]
    # GPUTest leaves the 1-pixel border untouched (zero-initialised Img) -> all zeros there
    vmap4[0, :, 3] = vmap4[-1, :, 3] = vmap4[:, 0, 3] = vmap4[:, -1, 3] = 0
    nmap4[0, :, 3] = nmap4[-1, :, 3] = nmap4[:, 0, 3] = nmap4[:, -1, 3] = 0
    return vmap4, nmap4


def frame_pair(cam: Camera, scene: Scene | None = None, *, seed=1234, rvec=(0.01, -0.015, 0.008), tvec=(0.02, -0.01, 0.015),
               max_depth=50.0, depth_noise=0.0):
    """Config-1 style pair: model view at identity pose, live view moved by (rvec, tvec).

    Returns (model, live, T_true) where T_true is the live camera pose in the model
    camera's frame (what the tracker should recover when started from identity)."""
    scene = scene or room_scene(seed)
    T0 = np.eye(4)
    T1 = pose_from(rvec, tvec)
    model = render(scene, cam, T0, seed=seed, max_depth=max_depth, hole_frac=0.0)
    live = render(scene, cam, T1, seed=seed + 1, max_depth=max_depth, depth_noise=depth_noise)
    return model, live, T1


def kitti_sequence(cam: Camera, n_frames: int, *, seed=2021, step=0.8, yaw_deg=0.2, prior_sigma_t=0.03, prior_sigma_r_deg=0.3,
                   depth_noise=0.004):
    """Config-2 style sequence: forward driving through `street_scene`.

    Yields per frame: model buffers rendered at the perturbed prior pose, live buffers at the
    ground-truth pose, the prior pose (what the caller passes as trans/rot) and the GT pose."""
    scene = street_scene(seed)
    rng = np.random.RandomState(seed)
    T = np.eye(4)
    frames = []
    for f in range(n_frames):
        dT = pose_from((0.0, np.deg2rad(yaw_deg) * np.sin(f * 0.05), 0.0), (0.0, 0.0, step))
        T = T @ dT
        prior = T @ pose_from(rng.randn(3) * np.deg2rad(prior_sigma_r_deg), rng.randn(3) * prior_sigma_t)
        model = render(scene, cam, prior, seed=seed + 2 * f, max_depth=60.0, hole_frac=0.0, zero_frac=0.0)
        live = render(scene, cam, T, seed=seed + 2 * f + 1, max_depth=60.0, depth_noise=depth_noise)
        frames.append({"model": model, "live": live, "prior": prior.astype(np.float32), "gt": T.astype(np.float32)})
    return frames


SURFEL_FLOATS = 15  # include/dmsodom/splat_c.h: pos.xyz conf | colour (encoded rgb, -, initTime, -) | times[3] | normal.xyz radius


def surfels_from_view(view: dict, T_wc: np.ndarray, cam: Camera, *, time=1, conf=10.0) -> np.ndarray:
    """A surfel map to predict model views from (bench.py's front-end leg): one surfel per valid pixel of a rendered view,
    position and normal moved to the world by `T_wc`, radius as Shaders/surfels.glsl:18-35 (getRadius), colour packed as
    Shaders/color.glsl:18-25 (encodeColor), layout of Vertex.cpp:49-50."""
    v = view["vmap4"][..., :3].reshape(-1, 3).astype(np.float64)
    n = view["nmap4"][..., :3].reshape(-1, 3).astype(np.float64)
    ok = np.isfinite(v).all(1) & np.isfinite(n).all(1) & (v[:, 2] > 0) & (np.abs(n).sum(1) > 0)
    v, n = v[ok], n[ok]
    c = view["rgba"].reshape(-1, 4)[ok].astype(np.int64)
    R, t = np.asarray(T_wc, np.float64)[:3, :3], np.asarray(T_wc, np.float64)[:3, 3]
    mean_focal = (1.0 / abs(cam.fx) + 1.0 / abs(cam.fy)) / 2.0
    radius = v[:, 2] * mean_focal * 1.41421356237
    radius = np.minimum(2.0 * radius, radius / np.maximum(np.abs(n[:, 2]), 1e-6))
    S = np.zeros((len(v), SURFEL_FLOATS), np.float32)
    S[:, 0:3] = v @ R.T + t
    S[:, 3] = conf
    S[:, 4] = ((c[:, 0] << 16) + (c[:, 1] << 8) + c[:, 2]).astype(np.float32)
    S[:, 6] = time
    S[:, 8:11] = -3.0
    S[:, 8] = time
    S[:, 11:14] = n @ R.T
    S[:, 14] = radius
    return S
