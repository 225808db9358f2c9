"""TEST INFRASTRUCTURE ONLY.  CPU restatement of the NID keyframing score:
cudafuncs.cu:906-917 (bin_idx), :1086-1157 (joint histogram kernels), :1513-1650 (computeNIDImg),
:1794-1915 (computeNIDDepth); KeyFrame.h:69-160; MutualInformation.cpp:154-212; ElasticFusion.cpp:651-662.
Entropies follow the reference's host loops literally (fp32, sequential, log2f)."""
import numpy as np

from . import oracle as orc

f32 = np.float32


def pick_active(dk, dko):
    # `d != __int_as_float(0x7fffffff)` is true for every d (NaN != NaN), so only the first branch of
    # cf.cu:1097-1112 is live: active iff dmap_kf <= dmap_kf_old (false when either is NaN)
    with np.errstate(invalid="ignore"):
        return dk <= dko


def hist_img(img_kf, img_kf_old, dmap_kf, dmap_kf_old, img_curr, nb):
    bw = 256 // nb
    a = np.where(pick_active(dmap_kf, dmap_kf_old), img_kf, img_kf_old).astype(np.int64) // bw
    b = img_curr.astype(np.int64) // bw
    ok = (a < nb) & (b < nb)  # the reference writes out of bounds otherwise
    h = np.zeros((nb, nb), np.int64)
    np.add.at(h, (b[ok], a[ok]), 1)
    return h


def _to_int(x):
    # CUDA float -> int conversion (cvt.rzi): NaN -> 0, saturating
    x = np.where(np.isnan(x), f32(0), x).astype(np.float64)
    return np.clip(np.trunc(x), -2**31, 2**31 - 1).astype(np.int64)


def hist_depth(dmap_kf, dmap_kf_old, dmap_curr, nb, max_depth_mm):
    bw = int(f32(max_depth_mm) / f32(nb))
    with np.errstate(invalid="ignore"):
        a = np.where(pick_active(dmap_kf, dmap_kf_old), dmap_kf, dmap_kf_old).astype(f32) * f32(1000.0)
        b = dmap_curr.astype(f32) * f32(1000.0)
        ba = _to_int(a / f32(bw)); bb = _to_int(b / f32(bw))
    ok = (ba >= 0) & (bb >= 0) & (ba < nb) & (bb < nb)
    h = np.zeros((nb, nb), np.int64)
    np.add.at(h, (bb[ok], ba[ok]), 1)
    return h


def nid_from_hist(h, num_points):
    """cf.cu:1556-1615, literally: fp32 histogram / num_points, sequential fp32 sums."""
    if num_points == 0:
        return f32(1.0)
    nb = h.shape[0]
    p = h.astype(f32) / f32(num_points)
    PB = np.zeros(nb, f32); PA = np.zeros(nb, f32)
    for a in range(nb):        # sequential fp32 accumulation, row by row
        PB = PB + p[a, :]
    for b in range(nb):
        PA = PA + p[:, b]
    def plogp(v):
        v = v.astype(f32)
        with np.errstate(divide="ignore", invalid="ignore"):
            return np.where(v == 0, f32(0), v * np.log2(v, dtype=f32)).astype(f32)
    joint = f32(0)
    for t in plogp(p).reshape(-1):
        if t != 0:
            joint = f32(joint + t)
    kf = f32(0)
    for t in plogp(PB):
        kf = f32(kf + t)
    cf = f32(0)
    for t in plogp(PA):
        cf = f32(cf + t)
    cf, kf, joint = f32(-cf), f32(-kf), f32(-joint)
    mutual = f32(f32(kf + cf) - joint)
    with np.errstate(divide="ignore", invalid="ignore"):
        return f32(f32(joint - mutual) / joint)


def keyframe_pyramids(vertex4, rgba, cutoff, level):
    """KeyFrame.h:69-160 (intensity + verticesToDepth) and MutualInformation.cpp:168-178 (pyramids)."""
    img = orc.bgr_to_intensity(rgba)
    d = orc.vertices_to_depth(vertex4, cutoff)
    for _ in range(level):
        img = orc.pyrdown_u8_gauss(img); d = orc.pyrdown_f32_gauss(d)
    return img, d


def nid_score(kf_vertex4, kf_rgba, old_vertex4, old_rgba, cutoff, curr_img, curr_depth, level, nb_img, nb_depth, max_depth_m, lam):
    ik, dk = keyframe_pyramids(kf_vertex4, kf_rgba, cutoff, level)
    iko, dko = keyframe_pyramids(old_vertex4, old_rgba, cutoff, level)
    hi = hist_img(ik, iko, dk, dko, curr_img, nb_img)
    hd = hist_depth(dk, dko, curr_depth, nb_depth, max_depth_m * 1000)
    n = curr_img.shape[0] * curr_img.shape[1]
    ni, nd = nid_from_hist(hi, n), nid_from_hist(hd, n)
    return ni, nd, f32(f32(f32(lam) * nd) + f32(f32(f32(1.0) - f32(lam)) * ni)), hi, hd
