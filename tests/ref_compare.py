"""Shared by tests/test_legacy_ref.py (GPU, live legacy kernels) and tests/test_golden_cpu.py (CPU, committed
outputs of the legacy kernels): runs every stage of the path through a stage provider and returns its outputs."""
import numpy as np


def unpack28(p):
    """27 products in JtJJtrSE3 order -> (A 6x6, b 6) as reduce.cu:415-424 does."""
    A = np.zeros((6, 6), np.float32); b = np.zeros(6, np.float32); k = 0
    for i in range(6):
        for j in range(i, 7):
            if j == 6:
                b[i] = p[k]
            else:
                A[i, j] = A[j, i] = p[k]
            k += 1
    return A, b


def unpack10(p):
    A = np.zeros((3, 3), np.float32); b = np.zeros(3, np.float32); k = 0
    for i in range(3):
        for j in range(i, 4):
            if j == 3:
                b[i] = p[k]
            else:
                A[i, j] = A[j, i] = p[k]
            k += 1
    return A, b


def case_inputs(synth, cam, seed=1234):
    model, live, T = synth.frame_pair(cam, seed=seed)
    return case_inputs_from(synth, cam, model, live)


def case_inputs_from(synth, cam, model, live):
    """The poses / warps of a comparison case for a given frame pair (synthetic or the reference's GPUTest fixtures)."""
    pose = synth.pose_from((0.0, 0.01, 0.0), (0.05, 0.0, -0.02)).astype(np.float32)
    Tc = (pose.astype(np.float64) @ synth.pose_from((0.004, -0.006, 0.003), (0.008, -0.004, 0.006))).astype(np.float32)
    K = np.array([[cam.fx, 0, cam.cx], [0, cam.fy, cam.cy], [0, 0, 1]], np.float64)
    Rt = synth.pose_from((0.003, -0.004, 0.002), (0.006, -0.003, 0.004))
    Rs = synth.rodrigues((0.01, -0.02, 0.005))
    K2 = K.copy(); K2[:2] /= 4
    return dict(model=model, live=live, pose=pose, Rc=Tc[:3, :3].copy(), tc=Tc[:3, 3].copy(),
                Rprev_inv=np.linalg.inv(pose[:3, :3].astype(np.float64)).astype(np.float32), tprev=pose[:3, 3].copy(),
                krkinv=(K @ Rt[:3, :3] @ np.linalg.inv(K)).astype(np.float32), kt=(K @ Rt[:3, 3]).astype(np.float32),
                so3_H=(K2 @ Rs @ np.linalg.inv(K2)).astype(np.float32), so3_kinv=np.linalg.inv(K2).astype(np.float32), so3_krlr=(K2 @ Rs).astype(np.float32))


def oracle_stages(orc, cam, c):
    """Every stage with the CPU oracle.  Returns a flat dict of arrays / scalars."""
    from densemonoslam_b200.odometry import DEFAULT_ANGLE_THRESH
    m, l = c["model"], c["live"]
    out = {}
    I0 = orc.bgr_to_intensity(l["rgba"]); I1 = orc.pyrdown_u8_gauss(I0); I2 = orc.pyrdown_u8_gauss(I1)
    M0 = orc.bgr_to_intensity(m["rgba"]); M1 = orc.pyrdown_u8_gauss(M0); M2 = orc.pyrdown_u8_gauss(M1)
    out.update(I0=I0, I1=I1, I2=I2)
    d1 = orc.pyrdown_u16(l["depth"]); out.update(d1=d1, d2=orc.pyrdown_u16(d1))
    v = orc.create_vmap(l["depth"], cam.fx, cam.fy, cam.cx, cam.cy, 25.0); n = orc.create_nmap(v)
    out.update(vmap=v, nmap=n)
    mv, mn = orc.copy_maps(m["vmap4"], m["nmap4"])
    mv1, mn1 = orc.resize_map(mv, False), orc.resize_map(mn, True)
    mv2, mn2 = orc.resize_map(mv1, False), orc.resize_map(mn1, True)
    R, t = c["pose"][:3, :3], c["pose"][:3, 3]
    tv0, tn0 = orc.transform_maps(mv, mn, R, t); tv1, tn1 = orc.transform_maps(mv1, mn1, R, t); tv2, tn2 = orc.transform_maps(mv2, mn2, R, t)
    out.update(mv0=tv0, mn0=tn0, mv1=tv1, mn1=tn1, mv2=tv2, mn2=tn2)
    zd = orc.vertices_to_depth(m["vmap4"], 6.0); zd1 = orc.pyrdown_f32_gauss(zd)
    out.update(zd=zd, zd1=zd1)
    dx, dy = orc.derivatives(I0); out.update(dx=dx, dy=dy)
    intr = (cam.fx, cam.fy, cam.cx, cam.cy)
    p, inl = orc.icp_step(c["Rc"], c["tc"], v, n, c["Rprev_inv"], c["tprev"], intr, tv0, tn0, 0.10, DEFAULT_ANGLE_THRESH)
    A, b = unpack28(p); out.update(icp_A=A, icp_b=b, icp_res=np.float32(p[27]), icp_inliers=np.int64(inl))
    zl = orc.vertices_to_depth(l["vmap4"], 6.0)
    corr, cnt, sig = orc.rgb_residual(1600.0, dx, dy, zd, zl, M0, I0, 0.07, c["kt"], c["krkinv"])
    p2, _ = orc.rgb_step(corr, float(np.float32(np.sqrt(np.float32(cnt)))), zd, intr, dx, dy)
    A2, b2 = unpack28(p2)
    out.update(rgb_valid=corr["valid"].copy(), rgb_u0=corr["u0"].copy(), rgb_v0=corr["v0"].copy(), rgb_diff=corr["diff"].copy(),
               rgb_count=np.int64(cnt), rgb_sigma=np.int64(sig), rgb_A=A2, rgb_b=b2)
    p3, c3 = orc.so3_step(M2, I2, c["so3_H"], c["so3_kinv"], c["so3_krlr"])
    A3, b3 = unpack10(p3); out.update(so3_A=A3, so3_b=b3, so3_res=np.float32(p3[9]), so3_count=np.int64(c3))
    return out


def legacy_stages(ref, cam, c):
    """The same stages with the reference's own kernels (oracle/_ref)."""
    from densemonoslam_b200.odometry import DEFAULT_ANGLE_THRESH
    m, l = c["model"], c["live"]
    out = {}
    I0 = ref.bgr_to_intensity(l["rgba"]); I1 = ref.pyrdown_u8_gauss(I0); I2 = ref.pyrdown_u8_gauss(I1)
    M0 = ref.bgr_to_intensity(m["rgba"]); M1 = ref.pyrdown_u8_gauss(M0); M2 = ref.pyrdown_u8_gauss(M1)
    out.update(I0=I0, I1=I1, I2=I2)
    d1 = ref.pyrdown_u16(l["depth"]); out.update(d1=d1, d2=ref.pyrdown_u16(d1))
    v, n = ref.vmap_nmap(l["depth"], cam.fx, cam.fy, cam.cx, cam.cy, 25.0)
    out.update(vmap=v, nmap=n)
    mvs, mns = ref.model_maps(m["vmap4"], m["nmap4"], c["pose"])
    out.update(mv0=mvs[0], mn0=mns[0], mv1=mvs[1], mn1=mns[1], mv2=mvs[2], mn2=mns[2])
    zd = ref.vertices_to_depth(m["vmap4"], 6.0); zd1 = ref.pyrdown_f32_gauss(zd)
    out.update(zd=zd, zd1=zd1)
    dx, dy = ref.derivatives(I0); out.update(dx=dx, dy=dy)
    intr = (cam.fx, cam.fy, cam.cx, cam.cy)
    A, b, res, inl = ref.icp_step(c["Rc"], c["tc"], v, n, c["Rprev_inv"], c["tprev"], intr, mvs[0], mns[0], 0.10, DEFAULT_ANGLE_THRESH)
    out.update(icp_A=A, icp_b=b, icp_res=np.float32(res), icp_inliers=np.int64(inl))
    zl = ref.vertices_to_depth(l["vmap4"], 6.0)
    corr, cnt, sig, A2, b2 = ref.rgb(1600.0, dx, dy, zd, zl, M0, I0, 0.07, c["kt"], c["krkinv"], intr)
    out.update(rgb_valid=corr["valid"].copy(), rgb_u0=np.where(corr["valid"] > 0, corr["u0"], 0).astype(np.int16),
               rgb_v0=np.where(corr["valid"] > 0, corr["v0"], 0).astype(np.int16),
               rgb_diff=np.where(corr["valid"] > 0, corr["diff"], 0).astype(np.int16), rgb_count=np.int64(cnt), rgb_sigma=np.int64(sig), rgb_A=A2, rgb_b=b2)
    A3, b3, r3, c3 = ref.so3_step(M2, I2, c["so3_H"], c["so3_kinv"], c["so3_krlr"])
    out.update(so3_A=A3, so3_b=b3, so3_res=np.float32(r3), so3_count=np.int64(c3))
    return out


def product_stages(g, torch, cam, c):
    """The same stages through the PRODUCT path (C ABI of libdmsodom.so via the Python mirror `g`): pyramids downloaded
    after the init_* calls, the four step operators for the reductions."""
    from densemonoslam_b200.odometry import CORR_DTYPE  # noqa: F401
    m, l = c["model"], c["live"]
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    out = {}
    # model side: maps in the global frame, depth / image pyramids of the model view
    g.initICPModel(dev(m["vmap4"]), dev(m["nmap4"]), 25.0, c["pose"])
    g.initRGBModel(dev(m["rgba"]))
    for i in range(3):
        out[f"mv{i}"] = g.download("vmap_g_prev", i); out[f"mn{i}"] = g.download("nmap_g_prev", i)
    out.update(zd=g.download("lastDepth", 0), zd1=g.download("lastDepth", 1))
    # live side from the live vertex map (nextDepth = its z channel), derivative images, photometric stage
    g.initICP(dev(l["vmap4"]), dev(l["nmap4"]), 25.0)
    g.initRGB(dev(l["rgba"]))
    out.update(I0=g.download("nextImage", 0), I1=g.download("nextImage", 1), I2=g.download("nextImage", 2))
    g.computeDerivativeImages()
    out.update(dx=g.download("dIdx", 0), dy=g.download("dIdy", 0))
    cnt, sig = g.computeRgbResidual(0, c["kt"], c["krkinv"])
    corr = g.download("corr", 0)
    p2, _ = g.rgbStep(0, float(np.float32(np.sqrt(np.float32(cnt)))))
    A2, b2 = unpack28(p2)
    out.update(rgb_valid=corr["valid"].copy(), rgb_u0=corr["u0"].copy(), rgb_v0=corr["v0"].copy(), rgb_diff=corr["diff"].copy(),
               rgb_count=np.int64(cnt), rgb_sigma=np.int64(sig), rgb_A=A2, rgb_b=b2)
    # SO(3): model image (as previous frame) against the live image at level 2
    g.initFirstRGB(dev(m["rgba"]))
    p3, c3 = g.so3Step(2, c["so3_H"], c["so3_kinv"], c["so3_krlr"])
    A3, b3 = unpack10(p3); out.update(so3_A=A3, so3_b=b3, so3_res=np.float32(p3[9]), so3_count=np.int64(c3))
    # live side from the sensor depth: u16 pyramid, vertex / normal maps, ICP step
    g.initICP(dev(l["depth"]), 25.0)
    out.update(d1=g.download("depth_tmp", 1), d2=g.download("depth_tmp", 2), vmap=g.download("vmap_curr", 0), nmap=g.download("nmap_curr", 0))
    p, inl = g.icpStep(0, c["Rc"], c["tc"], c["Rprev_inv"], c["tprev"])
    A, b = unpack28(p); out.update(icp_A=A, icp_b=b, icp_res=np.float32(p[27]), icp_inliers=np.int64(inl))
    return out


def check_identical_integers(rep):
    """Product path in legacy-numerics mode vs the reference's kernels: every integer output identical, counts identical,
    the float outputs within the stated tolerance (the summation order of the reduced systems differs)."""
    for k in INT_KEYS:
        assert rep[k]["mismatch"] == 0, (k, rep[k])
    for k in COUNT_KEYS:
        assert rep[k]["oracle"] == rep[k]["reference"], (k, rep[k])
    for k in MAP_KEYS:
        assert rep[k]["nan_pattern_mismatch"] == 0, (k, rep[k])
    for k in F32_KEYS:
        assert rep[k]["nan_pattern_mismatch"] == 0, (k, rep[k])
    for k in SYS_KEYS:
        assert rep[k]["max_rel_to_norm"] <= 1e-4, (k, rep[k])


INT_KEYS = ["I0", "I1", "I2", "d1", "d2", "dx", "dy", "rgb_valid", "rgb_u0", "rgb_v0", "rgb_diff"]
MAP_KEYS = ["vmap", "nmap", "mv0", "mn0", "mv1", "mn1", "mv2", "mn2"]
F32_KEYS = ["zd", "zd1"]
SYS_KEYS = ["icp_A", "icp_b", "rgb_A", "rgb_b", "so3_A", "so3_b"]
COUNT_KEYS = ["icp_inliers", "rgb_count", "rgb_sigma", "so3_count"]


def compare(o, r):
    """Mismatch statistics oracle vs reference outputs."""
    rep = {}
    for k in INT_KEYS:
        a, b = o[k].astype(np.int64), r[k].astype(np.int64)
        rep[k] = {"n": int(a.size), "mismatch": int(np.count_nonzero(a != b)), "max_abs": int(np.max(np.abs(a - b)))}
    for k in MAP_KEYS:
        a, b = o[k], r[k]
        nan_a, nan_b = np.isnan(a[0]), np.isnan(b[0])
        ok = ~nan_a & ~nan_b
        d = np.abs(a[:, ok] - b[:, ok]); den = np.maximum(np.abs(b[:, ok]), 1e-3)
        rep[k] = {"n": int(a[0].size), "nan_pattern_mismatch": int(np.count_nonzero(nan_a != nan_b)), "max_rel": float(np.max(d / den)) if d.size else 0.0,
                  "max_abs": float(np.max(d)) if d.size else 0.0,
                  "bit_mismatch": int(np.count_nonzero(a[:, ok].view(np.uint32) != b[:, ok].view(np.uint32)))}
    for k in F32_KEYS:
        a, b = o[k], r[k]
        ok = ~np.isnan(a) & ~np.isnan(b)
        rep[k] = {"n": int(a.size), "nan_pattern_mismatch": int(np.count_nonzero(np.isnan(a) != np.isnan(b))),
                  "max_rel": float(np.max(np.abs(a[ok] - b[ok]) / np.maximum(np.abs(b[ok]), 1e-6))) if ok.any() else 0.0,
                  "bit_mismatch": int(np.count_nonzero(a[ok].view(np.uint32) != b[ok].view(np.uint32)))}
    for k in SYS_KEYS:
        a, b = o[k].astype(np.float64), r[k].astype(np.float64)
        rep[k] = {"max_rel_to_norm": float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-30))}
    for k in COUNT_KEYS:
        rep[k] = {"oracle": int(o[k]), "reference": int(r[k])}
    for k in ("icp_res", "so3_res"):
        rep[k] = {"oracle": float(o[k]), "reference": float(r[k])}
    return rep


def check(rep, exact_counts=True):
    """The gates of the oracle-vs-reference comparison (fast-math differences of the legacy build are bounded).
    exact_counts: inlier / correspondence / SO(3) counts must be IDENTICAL (what was measured on the synthetic cases); the sum
    of squared differences may move by the few correspondence records that differ."""
    # exact: intensity conversion, derivative images (same FMA contraction, no division involved)
    for k in ("I0", "dx", "dy"):
        assert rep[k]["mismatch"] == 0, (k, rep[k])
    # u8 / u16 pyramids: the legacy build divides with --prec-div=false, so a quotient that is an exact integer can
    # come out one ulp low and truncate to the value below (measured on B200: 0.03 % of the u8, 0.6-0.8 % of the u16 pixels)
    for k in ("I1", "I2"):
        assert rep[k]["mismatch"] <= 2e-3 * rep[k]["n"] and rep[k]["max_abs"] <= 1, (k, rep[k])
    for k in ("d1", "d2"):
        assert rep[k]["mismatch"] <= 1.5e-2 * rep[k]["n"] and rep[k]["max_abs"] <= 1, (k, rep[k])
    for k in MAP_KEYS:
        assert rep[k]["nan_pattern_mismatch"] == 0, (k, rep[k])
        if k in ("vmap", "mv0", "mv1", "mv2"):
            assert rep[k]["max_rel"] <= 2e-6, (k, rep[k])   # d/1000.f with the approximate division: 1-2 ulp
        else:
            assert rep[k]["max_abs"] <= 3e-4, (k, rep[k])   # unit normals: rsqrtf + cross products of 1-ulp-different vertices
    for k in F32_KEYS:
        assert rep[k]["nan_pattern_mismatch"] == 0 and rep[k]["max_rel"] <= 1e-6, (k, rep[k])
    # correspondence masks and counts: identical up to a handful of threshold-boundary pixels
    n = rep["rgb_valid"]["n"]
    assert rep["rgb_valid"]["mismatch"] <= max(3, 2e-5 * n), rep["rgb_valid"]
    for k in ("rgb_u0", "rgb_v0", "rgb_diff"):
        assert rep[k]["mismatch"] <= max(6, 1e-4 * n), (k, rep[k])
    for k in COUNT_KEYS:
        a, b = rep[k]["oracle"], rep[k]["reference"]
        if k == "rgb_sigma":
            assert abs(a - b) <= 255 * 255 * (rep["rgb_diff"]["mismatch"] + rep["rgb_valid"]["mismatch"]), (k, rep[k])
        elif exact_counts:
            assert a == b, (k, rep[k])
        else:
            assert abs(a - b) <= max(3, 2e-4 * abs(b)), (k, rep[k])
    for k in SYS_KEYS:
        assert rep[k]["max_rel_to_norm"] <= 1e-4, (k, rep[k])
