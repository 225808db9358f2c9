"""TEST INFRASTRUCTURE ONLY.  CPU restatement (numpy, fp32 ufuncs = IEEE, unfused, left to right) of
IndexMap::combinedPredict (IndexMap.cpp:253-368): Shaders/splat.vert:42-95, Shaders/combo_splat.frag:36-66,
Shaders/color.glsl:27-34, with the rasterisation rules stated in include/dmsodom/splat_c.h (point sprite covers the
pixel centres within size/2 of its centre, centre outside the viewport = clipped, 24-bit depth, GL_LESS, draw order
breaks ties).  Parity unpinned against the GL pipeline itself (it cannot run on this box)."""
import numpy as np

f32 = np.float32
SURFEL_FLOATS = 15


def _dot(ax, ay, az, bx, by, bz):
    return (ax * bx + ay * by) + az * bz


def _normalize(x, y, z):
    inv = f32(1.0) / np.sqrt(_dot(x, y, z, x, y, z))
    return x * inv, y * inv, z * inv


def vertex_stage(S, t_inv, cam, rows, cols, maxDepth, confThreshold, time, timeIdx, maxTime, timeDelta, active):
    """splat.vert:42-95 for all surfels at once."""
    cx, cy, fx, fy = (f32(v) for v in cam)
    t = np.asarray(t_inv, f32).reshape(4, 4)
    px, py, pz, conf = S[:, 0], S[:, 1], S[:, 2], S[:, 3]
    vt = S[:, 8 + timeIdx]
    with np.errstate(all="ignore"):
        hx = _dot(t[0, 0], t[0, 1], t[0, 2], px, py, pz) + t[0, 3]
        hy = _dot(t[1, 0], t[1, 1], t[1, 2], px, py, pz) + t[1, 3]
        hz = _dot(t[2, 0], t[2, 1], t[2, 2], px, py, pz) + t[2, 3]
        t3 = vt == f32(-3.0)
        actv = bool(active)
        reject = ~((not actv) & t3) & ((hz > f32(maxDepth)) | (hz < 0) | (conf < f32(confThreshold)) | (actv & t3) |
                                        (~t3 & ((f32(time) - vt) > f32(timeDelta))) | (vt > f32(maxTime)))
        nx, ny, nz = _normalize(_dot(t[0, 0], t[0, 1], t[0, 2], S[:, 11], S[:, 12], S[:, 13]),
                                _dot(t[1, 0], t[1, 1], t[1, 2], S[:, 11], S[:, 12], S[:, 13]),
                                _dot(t[2, 0], t[2, 1], t[2, 2], S[:, 11], S[:, 12], S[:, 13]))
        rad = S[:, 14]
        ux, uy, uz = _normalize(ny - nz, -nx, nx)
        r2 = f32(1.41421356)
        x1 = ((ux * rad) * r2, (uy * rad) * r2, (uz * rad) * r2)
        y1 = (ny * x1[2] - x1[1] * nz, nz * x1[0] - x1[2] * nx, nx * x1[1] - x1[0] * ny)
        ix, iy = [], []
        for d in (x1, y1, tuple(-c for c in y1), tuple(-c for c in x1)):
            qx, qy, qz = hx + d[0], hy + d[1], hz + d[2]
            ix.append(((fx * qx) / qz) + cx); iy.append(((fy * qy) / qz) + cy)
        # fminf / fmaxf semantics: a NaN operand is ignored
        xmin = np.fmin(np.fmin(np.fmin(ix[0], ix[1]), ix[2]), ix[3]); xmax = np.fmax(np.fmax(np.fmax(ix[0], ix[1]), ix[2]), ix[3])
        ymin = np.fmin(np.fmin(np.fmin(iy[0], iy[1]), iy[2]), iy[3]); ymax = np.fmax(np.fmax(np.fmax(iy[0], iy[1]), iy[2]), iy[3])
        size = np.fmax(f32(0), np.fmax(np.abs(xmax - xmin), np.abs(ymax - ymin)))
        xw = ((fx * hx) / hz) + cx; yw = ((fy * hy) / hz) + cy
        visible = ~reject & (xw >= 0) & (xw <= f32(cols)) & (yw >= 0) & (yw <= f32(rows)) & (size > 0)
    return dict(pos=(hx, hy, hz), conf=conf, n=(nx, ny, nz), rad=rad, xw=xw, yw=yw, size=size, visible=visible, colour=S[:, 4], initTime=S[:, 6])


def fragment(v, sel, i, j, cam, maxDepth):
    """combo_splat.frag:36-66 for surfels `sel` at pixels (i, j) (arrays of the same length)."""
    cx, cy, fx, fy = (f32(c) for c in cam)
    with np.errstate(all="ignore"):
        fxc = i.astype(f32) + f32(0.5); fyc = j.astype(f32) + f32(0.5)
        lx, ly, lz = _normalize((fxc - cx) / fx, (fyc - cy) / fy, np.ones_like(fxc))
        px, py, pz = (c[sel] for c in v["pos"]); nx, ny, nz = (c[sel] for c in v["n"]); rad = v["rad"][sel]
        sc = _dot(px, py, pz, nx, ny, nz) / _dot(lx, ly, lz, nx, ny, nz)
        cpx, cpy, cpz = sc * lx, sc * ly, sc * lz
        dx, dy, dz = cpx - px, cpy - py, cpz - pz
        keep = ~(_dot(dx, dy, dz, dx, dy, dz) > rad * rad)
        depth = (cpz / (f32(2.0) * f32(maxDepth))) + f32(0.5)
        keep &= (depth >= 0) & (depth <= 1)
        d24 = np.where(keep, depth.astype(np.float64) * 16777215.0 + 0.5, 0).astype(np.uint64)
    return keep, cpz, d24


def predict(S, t_inv, cam, rows, cols, maxDepth, confThreshold, time, timeIdx, maxTime, timeDelta, active):
    """Returns image [H,W,4] u8, vertex [H,W,4] f32, normal [H,W,4] f32, time [H,W] u32."""
    S = np.ascontiguousarray(S, f32).reshape(-1, SURFEL_FLOATS)
    v = vertex_stage(S, t_inv, cam, rows, cols, maxDepth, confThreshold, time, timeIdx, maxTime, timeDelta, active)
    vis = np.nonzero(v["visible"])[0]
    zbuf = np.full(rows * cols, np.iinfo(np.uint64).max, np.uint64)
    if len(vis):
        with np.errstate(all="ignore"):
            h = v["size"][vis] * f32(0.5)
            x0 = np.maximum(0, np.ceil((v["xw"][vis] - h) - f32(0.5)).astype(np.int64)); x1 = np.minimum(cols - 1, np.floor((v["xw"][vis] + h) - f32(0.5)).astype(np.int64))
            y0 = np.maximum(0, np.ceil((v["yw"][vis] - h) - f32(0.5)).astype(np.int64)); y1 = np.minimum(rows - 1, np.floor((v["yw"][vis] + h) - f32(0.5)).astype(np.int64))
        w = np.maximum(x1 - x0 + 1, 0); hh = np.maximum(y1 - y0 + 1, 0)
        for oy in range(int(hh.max(initial=0))):
            for ox in range(int(w.max(initial=0))):
                m = (ox < w) & (oy < hh)
                if not m.any():
                    continue
                sel = vis[m]; i = x0[m] + ox; j = y0[m] + oy
                keep, _, d24 = fragment(v, sel, i, j, cam, maxDepth)
                if keep.any():
                    key = (d24[keep] << np.uint64(32)) | sel[keep].astype(np.uint64)
                    np.minimum.at(zbuf, (j[keep] * cols + i[keep]), key)
    image = np.zeros((rows * cols, 4), np.uint8); vertex = np.zeros((rows * cols, 4), f32); normal = np.zeros((rows * cols, 4), f32)
    tim = np.zeros(rows * cols, np.uint32)
    pix = np.nonzero(zbuf != np.iinfo(np.uint64).max)[0]
    if len(pix):
        sel = (zbuf[pix] & np.uint64(0xffffffff)).astype(np.int64)
        j, i = pix // cols, pix % cols
        _, z, _ = fragment(v, sel, i, j, cam, maxDepth)
        cx, cy, fx, fy = (f32(c) for c in cam)
        c = v["colour"][sel].astype(np.int64)
        image[pix] = np.stack([(c >> 16) & 0xFF, (c >> 8) & 0xFF, c & 0xFF, np.full_like(c, 255)], -1).astype(np.uint8)
        fxc = i.astype(f32) + f32(0.5); fyc = j.astype(f32) + f32(0.5)
        vertex[pix] = np.stack([((fxc - cx) * z) * (f32(1.0) / fx), ((fyc - cy) * z) * (f32(1.0) / fy), z, v["conf"][sel]], -1)
        normal[pix] = np.stack([v["n"][0][sel], v["n"][1][sel], v["n"][2][sel], v["rad"][sel]], -1)
        tim[pix] = v["initTime"][sel].astype(np.uint32)
    return image.reshape(rows, cols, 4), vertex.reshape(rows, cols, 4), normal.reshape(rows, cols, 4), tim.reshape(rows, cols)


def surfels_from_view(vmap4, nmap4, rgba, pose, cam, time=1, sensor=0, conf=10.0):
    """Test helper: one surfel per valid pixel of a rendered view (position / normal moved to the world by `pose`), radius
    as Shaders/surfels.glsl:18-35 (getRadius), colour encoded as color.glsl:18-25."""
    cx, cy, fx, fy = cam
    v = vmap4[..., :3].reshape(-1, 3).astype(np.float64); n = nmap4[..., :3].reshape(-1, 3).astype(np.float64)
    ok = np.isfinite(v).all(1) & np.isfinite(n).all(1) & (v[:, 2] > 0) & (np.abs(n).sum(1) > 0)
    v, n = v[ok], n[ok]
    c = rgba.reshape(-1, 4)[ok].astype(np.int64)
    R, t = np.asarray(pose, np.float64)[:3, :3], np.asarray(pose, np.float64)[:3, 3]
    meanFocal = ((1.0 / abs(fx)) + (1.0 / abs(fy))) / 2.0
    radius = (v[:, 2] * meanFocal) * 1.41421356237
    rad = np.minimum(2.0 * radius, radius / np.maximum(np.abs(n[:, 2]), 1e-6))
    S = np.zeros((len(v), SURFEL_FLOATS), f32)
    S[:, 0:3] = (v @ R.T + t); S[:, 3] = conf
    S[:, 4] = ((c[:, 0] << 16) + (c[:, 1] << 8) + c[:, 2]).astype(f32); S[:, 6] = time
    S[:, 8:11] = -3.0; S[:, 8 + sensor] = time
    S[:, 11:14] = n @ R.T; S[:, 14] = rad
    return S
