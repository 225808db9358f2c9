"""Diagnostic: locates the pixels whose ICP inlier decision differs between the product path in legacy-numerics mode and the
reference's own icpStep kernel (which has no per-pixel output) by bisection on the set of live pixels that are not NaN."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from densemonoslam_b200 import synth
from densemonoslam_b200.odometry import RGBDOdometry, DEFAULT_ANGLE_THRESH
from oracle import ref_legacy as ref
import ref_compare as rc

cam = synth.CAM_KITTI
f = synth.kitti_sequence(cam, 1, seed=2021)[0]
c = rc.case_inputs_from(synth, cam, f["model"], f["live"])
g = RGBDOdometry(cam.width, cam.height, cam.cx, cam.cy, cam.fx, cam.fy)
g.set_numerics("legacy")
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
g.initICPModel(dev(c["model"]["vmap4"]), dev(c["model"]["nmap4"]), 25.0, c["pose"])
g.initICP(dev(c["live"]["depth"]), 25.0)
v, n = g.download("vmap_curr", 0), g.download("nmap_curr", 0)
mv, mn = g.download("vmap_g_prev", 0), g.download("nmap_g_prev", 0)
intr = (cam.fx, cam.fy, cam.cx, cam.cy)
_, cnt, mask = g.icpStep(0, c["Rc"], c["tc"], c["Rprev_inv"], c["tprev"], want_mask=True)
H, W = mask.shape

def ref_count(rows):
    vv = v.copy(); keep = np.zeros(H, bool); keep[rows[0]:rows[1]] = True
    vv[0][~keep] = np.nan
    return ref.icp_step(c["Rc"], c["tc"], vv, n, c["Rprev_inv"], c["tprev"], intr, mv, mn, 0.10, DEFAULT_ANGLE_THRESH)[3]

def ref_count_px(y, x0, x1):
    vv = v.copy(); vv[0][:] = np.nan; vv[0][y, x0:x1] = v[0][y, x0:x1]
    return ref.icp_step(c["Rc"], c["tc"], vv, n, c["Rprev_inv"], c["tprev"], intr, mv, mn, 0.10, DEFAULT_ANGLE_THRESH)[3]

print("total ours", cnt, "ref", ref_count((0, H)))
bad_rows = []
def rec(a, b):
    r = ref_count((a, b)); o = int(mask[a:b].sum())
    if r == o: return
    if b - a == 1: bad_rows.append(a); return
    m = (a + b) // 2; rec(a, m); rec(m, b)
rec(0, H)
print("rows with a difference", bad_rows)
for y in bad_rows:
    def recx(a, b):
        r = ref_count_px(y, a, b); o = int(mask[y, a:b].sum())
        if r == o: return
        if b - a == 1:
            print("pixel", y, a, "ours", int(mask[y, a]), "ref", r); report(y, a); return
        m = (a + b) // 2; recx(a, m); recx(m, b)
    def report(y, x):
        f32 = np.float32
        Rc = c["Rc"].astype(f32); tc = c["tc"].astype(f32); Ri = c["Rprev_inv"].astype(f32); tp = c["tprev"].astype(f32)
        vc = v[:, y, x]; nc = n[:, y, x]
        vg = (Rc.astype(np.float64) @ vc + tc); vcp = Ri.astype(np.float64) @ (vg - tp)
        ux = vcp[0] * cam.fx / vcp[2] + cam.cx; uy = vcp[1] * cam.fy / vcp[2] + cam.cy
        print("  vcurr", vc, "ncurr", nc, "proj (float64)", ux, uy)
        ix, iy = int(round(ux)), int(round(uy))
        for (jx, jy) in {(ix, iy), (int(np.floor(ux)), iy), (int(np.ceil(ux)), iy), (ix, int(np.floor(uy))), (ix, int(np.ceil(uy)))}:
            if 0 <= jx < W and 0 <= jy < H:
                vp = mv[:, jy, jx].astype(np.float64); npv = mn[:, jy, jx].astype(np.float64)
                ng = Rc.astype(np.float64) @ nc
                print("   at", jx, jy, "dist", np.linalg.norm(vp - vg), "sine", np.linalg.norm(np.cross(ng, npv)), "(thresholds 0.10,", DEFAULT_ANGLE_THRESH, ")")
    recx(0, W)
