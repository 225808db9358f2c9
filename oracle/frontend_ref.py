"""TEST INFRASTRUCTURE ONLY.  CPU restatement (numpy + the C oracle) of the tracking part of
ElasticFusion::processFrame -- ElasticFusion.cpp:99-268 and :397-442 -- with the GLSL passes that feed it:
FillIn (Shaders/FillIn.cpp:65-190, fill_vertex.frag:37-52, fill_normal.frag:38-51, geometry.glsl:53-77,
fill_rgb.frag:29-36), Resize::image + denseEnough (Shaders/Resize.cpp:67-98, EF.cpp:84-97), rodrigues2
(EF.cpp:941-986).  Parity unpinned against the GL shaders themselves (they cannot run on this box)."""
import numpy as np

from . import oracle as orc

f32 = np.float32


def fill_in(mv, mn, mi, depth, live, cam, pass_geom=False, pass_img=False):
    """cam = (cx, cy, fx, fy).  Every fp32 operation is one numpy float32 ufunc (IEEE, no fusion)."""
    H, W = depth.shape
    cx, cy, ifx, ify = f32(cam[0]), f32(cam[1]), f32(1.0) / f32(cam[2]), f32(1.0) / f32(cam[3])
    with np.errstate(all="ignore"):
        def vertex(dx, dy):  # geometry.glsl:53-58 sampled at (x+dx, y+dy) with clamp-to-edge
            xs = np.arange(W, dtype=np.int64) + dx; ys = np.arange(H, dtype=np.int64) + dy
            z = depth[np.clip(ys, 0, H - 1)][:, np.clip(xs, 0, W - 1)].astype(f32) / f32(1000.0)
            X = ((xs.astype(f32)[None, :] - cx) * z) * ifx
            Y = ((ys.astype(f32)[:, None] - cy) * z) * ify
            return X, Y, z
        vx, vy, vz = vertex(0, 0)
        take = (mv[..., 2] == 0) | bool(pass_geom)
        ov = mv.copy()
        ov[take] = np.stack([vx, vy, vz, np.ones_like(vz)], -1)[take]
        ax, ay, az = vertex(1, 0); bx, by, bz = vertex(0, 1)
        dxx, dxy, dxz = ax - vx, ay - vy, az - vz
        dyx, dyy, dyz = bx - vx, by - vy, bz - vz
        c0 = dxy * dyz - dxz * dyy; c1 = dxz * dyx - dxx * dyz; c2 = dxx * dyy - dxy * dyx
        inv = f32(1.0) / np.sqrt((c0 * c0 + c1 * c1) + c2 * c2)
        n = np.stack([c0 * inv, c1 * inv, c2 * inv, np.ones_like(c0)], -1).astype(f32)
        taken = (mn[..., 2] == 0) | bool(pass_geom)
        on = mn.copy()
        on[taken] = n[taken]
    takei = ((mi[..., 0] == 0) & (mi[..., 1] == 0) & (mi[..., 2] == 0)) | bool(pass_img)
    oi = mi.copy()
    oi[takei] = live[takei]
    return ov, on, oi


def dense_fraction(img, sample=20):
    H, W = img.shape[:2]
    dh, dw = max(H // sample, 1), max(W // sample, 1)
    sx = np.minimum(((2 * np.arange(dw) + 1) * W) // (2 * dw), W - 1)
    sy = np.minimum(((2 * np.arange(dh) + 1) * H) // (2 * dh), H - 1)
    s = img[sy][:, sx]
    cnt = int(((s[..., 0] > 0) & (s[..., 1] > 0) & (s[..., 2] > 0)).sum())
    return f32(cnt) / f32(dh * dw)


def rodrigues2(R):
    """EF.cpp:941-986."""
    U, _, Vt = np.linalg.svd(np.asarray(R, np.float64))
    Rq = (U @ Vt).astype(f32).astype(np.float64)
    rx, ry, rz = Rq[2, 1] - Rq[1, 2], Rq[0, 2] - Rq[2, 0], Rq[1, 0] - Rq[0, 1]
    s = np.sqrt((rx * rx + ry * ry + rz * rz) * 0.25)
    c = min(1.0, max(-1.0, (np.trace(Rq) - 1) * 0.5))
    theta = np.arccos(c)
    if s < 1e-5:
        if c > 0:
            rx = ry = rz = 0.0
        else:
            rx = np.sqrt(max((Rq[0, 0] + 1) * 0.5, 0.0))
            ry = np.sqrt(max((Rq[1, 1] + 1) * 0.5, 0.0)) * (-1.0 if Rq[0, 1] < 0 else 1.0)
            rz = np.sqrt(max((Rq[2, 2] + 1) * 0.5, 0.0)) * (-1.0 if Rq[0, 2] < 0 else 1.0)
            if abs(rx) < abs(ry) and abs(rx) < abs(rz) and (Rq[1, 2] > 0) != (ry * rz > 0):
                rz = -rz
            theta /= np.sqrt(rx * rx + ry * ry + rz * rz)
            rx *= theta; ry *= theta; rz *= theta
    else:
        vth = theta / (2 * s)
        rx *= vth; ry *= vth; rz *= vth
    return np.array([rx, ry, rz], f32)


def velocity_weighting(currPose, lastPose, weightMultiplier=1.0):
    """EF.cpp:251-266."""
    diff = (np.linalg.inv(np.asarray(currPose, np.float64)) @ np.asarray(lastPose, np.float64)).astype(f32)
    w = max(f32(np.linalg.norm(diff[:3, 3].astype(np.float64))), f32(np.linalg.norm(rodrigues2(diff[:3, :3]).astype(np.float64))))
    largest, minWeight = f32(0.01), f32(0.5)
    if w > largest:
        w = largest
    return float(max(f32(1.0) - (w / largest), minWeight) * f32(weightMultiplier))


class FrontEndRef:
    """processFrame's tracking state machine over two OracleOdometry instances (Context.h:319-320)."""

    def __init__(self, width, height, cx, cy, fx, fy, predictor, depthCutoff=3.0, icpWeight=10.0, rgbOnly=False, pyramid=True, fastOdom=False, so3=True,
                 frameToFrameRGB=False, reloc=False, closeLoops=True, hybridTracking=True, weightMultiplier=1.0):
        self.cam = (cx, cy, fx, fy)
        self.f2m = orc.OracleOdometry(width, height, cx, cy, fx, fy)
        self.m2m = orc.OracleOdometry(width, height, cx, cy, fx, fy)
        self.predictor = predictor
        self.o = dict(depthCutoff=depthCutoff, icpWeight=icpWeight, rgbOnly=rgbOnly, pyramid=pyramid, fastOdom=fastOdom, so3=so3,
                      frameToFrameRGB=frameToFrameRGB, reloc=reloc, closeLoops=closeLoops, hybridTracking=hybridTracking, weightMultiplier=weightMultiplier)
        self.tick, self.lost, self.trackingCount, self.lastFrameRecovery = 1, False, 0, False  # Context.h:43
        self.currPose = np.eye(4, dtype=f32)
        self.fill = None

    def process(self, rgba, depth_raw, inPose=None):
        o = self.o
        out = {"tick": self.tick, "tracked": False, "loopChecked": False, "loopAccepted": False, "shouldFillIn": False}
        filt = orc.depth_bilateral(depth_raw, o["depthCutoff"])  # EF.cpp:116
        self.filtered = filt
        if self.tick == 1:  # EF.cpp:131-152
            self.currPose = np.array(inPose if inPose is not None else np.eye(4), f32)
            self.f2m.initFirstRGB(rgba)
            out["trackingOk"] = True; out["weighting"] = float(f32(o["weightMultiplier"]))
        else:
            lastPose = self.currPose.copy()
            trackingOk = True
            if inPose is not None:
                self.currPose = np.array(inPose, f32)
            view = self.predictor(self.currPose, 0, self.tick)
            fv, fn, fi = fill_in(view["vertex"], view["normal"], view["image"], filt, rgba, self.cam, self.lost, self.lost or o["frameToFrameRGB"])
            self.fill = (fv, fn, fi)
            frac = dense_fraction(view["image"])
            shouldFillIn = not (frac > f32(0.95))
            out.update(denseFraction=float(frac), shouldFillIn=shouldFillIn, tracked=True)
            if o["hybridTracking"]:
                t = self.f2m
                t.initICPModel(fv if shouldFillIn else view["vertex"], fn if shouldFillIn else view["normal"], self.currPose)
                t.initRGBModel(fi if (shouldFillIn or o["frameToFrameRGB"]) else view["image"])
                t.initICP(filt, o["depthCutoff"]); t.initRGB(rgba)
                trans, rot, res = t.getIncrementalTransformation(self.currPose[:3, 3].copy(), self.currPose[:3, :3].copy(), o["rgbOnly"], o["icpWeight"],
                                                                 o["pyramid"], o["fastOdom"], o["so3"])
                out["frameToModel"] = res
                trackingOk = (not o["reloc"]) or res.lastICPError < f32(1e-04)
                if o["reloc"]:
                    covOk = bool(np.all(~(np.diag(orc.covariance(np.array(res.A[:]))) > 1e-04)))
                    if not self.lost:
                        if not covOk:
                            trackingOk = False
                        if not trackingOk:
                            self.trackingCount += 1
                            if self.trackingCount > 10:
                                self.lost = True
                        else:
                            self.trackingCount = 0
                    elif self.lastFrameRecovery:
                        if not covOk:
                            trackingOk = False
                        if trackingOk:
                            self.lost = False; self.trackingCount = 0
                        self.lastFrameRecovery = False
                self.currPose[:3, 3] = trans; self.currPose[:3, :3] = rot
            out["trackingOk"] = trackingOk
            out["weighting"] = velocity_weighting(self.currPose, lastPose, o["weightMultiplier"])
            if not self.lost and o["closeLoops"]:
                cur = self.predictor(self.currPose, 1, self.tick) or view
                old = self.predictor(self.currPose, 2, self.tick)
                if old is not None:
                    m = self.m2m
                    m.initICPModel(old["vertex"], old["normal"], self.currPose); m.initRGBModel(old["image"])
                    m.initICPMaps(cur["vertex"], cur["normal"]); m.initRGB(cur["image"])
                    trans, rot, res = m.getIncrementalTransformation(self.currPose[:3, 3].copy(), self.currPose[:3, :3].copy(), False, 10.0, o["pyramid"],
                                                                     o["fastOdom"], False)
                    cov = np.diag(orc.covariance(np.array(res.A[:])))
                    covOk = bool(np.all(~(cov > 8e-05)))
                    est = np.eye(4, dtype=f32); est[:3, 3] = trans; est[:3, :3] = rot
                    out.update(loopChecked=True, loopPose=est, loopCovDiag=cov, modelToModel=res,
                               loopAccepted=bool(covOk and res.lastICPCount > 15000 and res.lastICPError < f32(0.0003)))
        out.update(pose=self.currPose.copy(), lost=self.lost, trackingCount=self.trackingCount, lastFrameRecovery=self.lastFrameRecovery)
        if not self.lost:
            self.tick += 1
        return out
