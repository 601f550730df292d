"""Independent numpy restatement of the reference's depth projection, written from the reference text
(src/lidar/pcd_projection.cc, src/lidar/pcd_projection.h) and NOT from oracle/pcd_oracle.c or csrc/proj.cu.

    sub-map keys        pcd_projection.h:67-73, pcd_projection.cc:223-255   (round(coord / len), float divide, C round)
    frustum             pcd_projection.cc:258-297, pcd_projection.h:124-158 (five planes of the quad pyramid, float32)
    cell selection      pcd_projection.cc:499-559                           (cell centre idx*len inside all five planes)
    projection / splat  pcd_projection.cc:305-468, :561-594                 (float32 transform, fp64 pinhole + OpenCV
                                                                             distortion, round, depth-dependent half widths
                                                                             incl. the b_y quirk of :397, min range, strict <)
    feature pixels      pcd_projection.cc:35, :122                          ((xy * scale).cast<int>(), in-bounds)

Every float32 step is a separate numpy float32 operation (one rounding each, no contraction).  Expressions the
reference spells out are taken left to right; Eigen's own reductions over three elements (the coefficients of a
Matrix3f * Vector3f product, Vector3f::norm) follow Eigen/src/Core/Redux.h: redux_novec_unroller<Start, Length> splits
at HalfLength = Length / 2, so Length 3 is a0 + (a1 + a2).  (The first draft of this file summed left to right and
disagreed with the oracle by one ulp of the range on 0.3 % of the pixels -- the one place where two readings differ.)
The reference's loop is sequential in (cell order, insertion order) with a strict `>` replace test, i.e. the FIRST
visited point wins a tie; `project` returns both that winner and the lowest-index winner (north_star's tie rule) so a
fixture can assert that they coincide.
"""
import numpy as np

F = np.float32


def quat_to_rot(q):
    """Eigen::Quaterniond(w, x, y, z).toRotationMatrix() (Eigen/src/Geometry/Quaternion.h), fp64, no normalisation."""
    w, x, y, z = (float(v) for v in q)
    tx, ty, tz = 2.0 * x, 2.0 * y, 2.0 * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    return np.array([[1.0 - (tyy + tzz), txy - twz, txz + twy],
                     [txy + twz, 1.0 - (txx + tzz), tyz - twx],
                     [txz - twy, tyz + twx, 1.0 - (txx + tyy)]], dtype=np.float64)


def _dot3(r, v):
    """row . vector in float32, Eigen's unrolled redux order for three terms: a0 + (a1 + a2)."""
    return F(F(r[0] * v[0]) + F(F(r[1] * v[1]) + F(r[2] * v[2])))


def _matvec(R, v):
    return np.array([_dot3(R[0], v), _dot3(R[1], v), _dot3(R[2], v)], dtype=F)


def _plane(a, b, c):
    """QuadPyramid::GetPlane (pcd_projection.h:136-143)."""
    ab = (a - b).astype(F)
    ac = (a - c).astype(F)
    n = np.array([F(F(ab[1] * ac[2]) - F(ab[2] * ac[1])), F(F(ab[2] * ac[0]) - F(ab[0] * ac[2])), F(F(ab[0] * ac[1]) - F(ab[1] * ac[0]))], dtype=F)
    d = F(-F(F(F(n[0] * a[0]) + F(n[1] * a[1])) + F(n[2] * a[2])))
    return np.array([n[0], n[1], n[2], d], dtype=F)


def pyramid(qvec, tvec, cam, width, height, opt):
    """SearchSubMap (pcd_projection.cc:258-297): returns (planes [5,4] f32, vertices [5,3] f32: apex, corner_1..4)."""
    s = float(opt["depth_image_scale"])
    img_h, img_w = int(height * s), int(width * s)
    fx, fy, cx, cy = (float(cam[k]) * s for k in range(4))                  # LImage holds SCALED intrinsics (:45-48)
    R = quat_to_rot(qvec)
    rot_wc = R.astype(F).T.copy()
    t_wc = _matvec(-rot_wc, np.asarray(tvec, np.float64).astype(F))
    x_min, x_max = F(-cx / fx), F((img_w - cx) / fx)
    y_min, y_max = F(-cy / fy), F((img_h - cy) / fy)
    cm = F(opt["choose_meter"])
    corners = []
    for xb, yb in ((x_max, y_max), (x_max, y_min), (x_min, y_min), (x_min, y_max)):
        v = np.array([xb, yb, F(1.0)], dtype=F)                             # center_v + corner
        corners.append((t_wc + (_matvec(rot_wc, v) * cm).astype(F)).astype(F))
    c1, c2, c3, c4 = corners
    planes = np.stack([_plane(c1, c4, c3), _plane(t_wc, c1, c2), _plane(t_wc, c2, c3), _plane(t_wc, c3, c4), _plane(t_wc, c4, c1)])
    return planes, np.stack([t_wc, c1, c2, c3, c4])


def c_round(x):
    """C round(): half away from zero."""
    x = np.asarray(x, dtype=np.float64)
    return np.where(x >= 0, np.floor(x + 0.5), -np.floor(-x + 0.5))


def submap_keys(cloud, opt):
    """GetKeyType (pcd_projection.h:67-73): round(float coord / float len) per axis -> int keys [n, 3] (x, y, z)."""
    k = np.empty((len(cloud), 3), dtype=np.int64)
    for a, ln in enumerate((opt["submap_length"], opt["submap_height"], opt["submap_width"])):
        k[:, a] = c_round((cloud[:, a].astype(F) / F(ln)).astype(F)).astype(np.int64)
    return k


def select_cells(keys_unique, planes, verts, bbox, opt):
    """SearchImageMap (pcd_projection.cc:499-559): which occupied cells (rows of keys_unique) are selected."""
    lens = (F(opt["submap_length"]), F(opt["submap_height"]), F(opt["submap_width"]))
    lo, hi = [], []
    for a in range(3):
        vals = np.concatenate([verts[:, a], np.asarray(bbox[a], dtype=F)]).astype(F)
        lo.append(int(c_round(F(vals.min() / lens[a]))) - 1)
        hi.append(int(c_round(F(vals.max() / lens[a]))) + 1)
    k = keys_unique
    inside = np.ones(len(k), dtype=bool)
    for a in range(3):
        inside &= (k[:, a] >= lo[a]) & (k[:, a] <= hi[a])
    x = (k[:, 0].astype(F) * lens[0]).astype(F)
    y = (k[:, 1].astype(F) * lens[1]).astype(F)
    z = (k[:, 2].astype(F) * lens[2]).astype(F)
    for p in planes:
        v = ((((p[0] * x).astype(F) + (p[1] * y).astype(F)).astype(F) + (p[2] * z).astype(F)).astype(F) + p[3]).astype(F)
        inside &= v <= 0.0
    return inside


def map_bbox(cloud):
    """BuildSubMap's running bbox (pcd_projection.cc:229-234); max starts at numeric_limits<float>::min() (> 0)."""
    fmin = np.finfo(F).tiny
    return [(F(min(np.finfo(F).max, cloud[:, a].min())), F(max(fmin, cloud[:, a].max()))) for a in range(3)]


def splat_coef(cam, opt):
    """The function-local statics a_x, b_x, a_y, b_y of pcd_projection.cc:391-397, as set by the FIRST camera."""
    s = float(opt["depth_image_scale"])
    fx, fy = float(cam[0]), float(cam[1])
    max_x = float(opt["max_proj_scale"]) * (fx / 3039.0) * (s / 0.2)
    max_y = float(opt["max_proj_scale"]) * (fy / 3039.0) * (s / 0.2)
    min_x = float(opt["min_proj_scale"]) * (fx / 3039.0) * (s / 0.2)
    min_y = float(opt["min_proj_scale"]) * (fy / 3039.0) * (s / 0.2)
    cm = float(F(opt["choose_meter"]))
    a_x = (max_x - min_x) / (float(opt["min_proj_dist"]) - cm)
    b_x = min_x - a_x * cm
    a_y = (max_y - min_y) / (float(opt["min_proj_dist"]) - cm)
    b_y = float(opt["min_proj_scale"]) - a_y * cm                          # :397 uses the UNSCALED min_proj_scale
    return np.array([a_x, b_x, a_y, b_y])


def distort_opencv(u, v, cam):
    """DistortOpenCV (pcd_projection.cc:561-594), fp64, expression order as written."""
    fx, fy, cx, cy, k1, k2, p1, p2 = (float(c) for c in cam[:8])
    x = (u - cx) / fx
    y = (v - cy) / fy
    r2 = x * x + y * y
    ra = 1.0 + k1 * r2 + k2 * r2 * r2
    tx = 2.0 * p1 * x * y + p2 * (r2 + 2.0 * x * x)
    ty = p1 * (r2 + 2.0 * y * y) + 2.0 * p2 * x * y
    return (x * ra * 1.0 + tx) * fx + cx, (y * ra * 1.0 + ty) * fy + cy


def project(cloud, qvec, tvec, cam, width, height, opt, coef, feat_mask=None):
    """ImageMapProj (pcd_projection.cc:305-468) over the selected cells, sequential semantics.
    Returns dict(range f32 [h, w] (+inf = empty), first i64 [h, w] winner by visit order (-1 empty),
    lowest i64 [h, w] lowest original index among the points of minimal range, n_sel cells, n_pts points visited)."""
    cloud = np.asarray(cloud, dtype=F)
    s = float(opt["depth_image_scale"])
    img_h, img_w = int(height * s), int(width * s)
    planes, verts = pyramid(qvec, tvec, cam, width, height, opt)
    keys = submap_keys(cloud, opt)
    uniq, inv = np.unique(keys, axis=0, return_inverse=True)               # lexicographic (x, y, z) = std::map order
    inv = inv.reshape(-1)
    sel = select_cells(uniq, planes, verts, map_bbox(cloud), opt)
    order = np.lexsort((np.arange(len(cloud)), inv))                       # cell order, insertion order inside a cell
    order = order[sel[inv[order]]]
    P = cloud[order]
    R = quat_to_rot(qvec).astype(F)
    t = np.asarray(tvec, np.float64).astype(F)
    x, y, z = P[:, 0], P[:, 1], P[:, 2]
    pc = [(((R[i, 0] * x).astype(F) + ((R[i, 1] * y).astype(F) + (R[i, 2] * z).astype(F)).astype(F)).astype(F) + t[i]).astype(F) for i in range(3)]
    fx, fy, cx, cy = (float(cam[k]) for k in range(4))                     # UNSCALED intrinsics (:352-355)
    with np.errstate(all="ignore"):
        keep = ~(pc[2] < 0)
        u = fx * (pc[0] / pc[2]).astype(F).astype(np.float64) + cx
        v = fy * (pc[1] / pc[2]).astype(F).astype(np.float64) + cy
        ud, vd = distort_opencv(u, v, cam)
        u0 = c_round(ud * s)
        v0 = c_round(vd * s)
    dist = pc[2].astype(np.float64)
    max_x = float(opt["max_proj_scale"]) * (fx / 3039.0) * (s / 0.2)
    max_y = float(opt["max_proj_scale"]) * (fy / 3039.0) * (s / 0.2)
    a_x, b_x, a_y, b_y = (float(c) for c in coef)
    near = (dist >= float(opt["min_lidar_proj_dist"])) & (dist <= float(opt["min_proj_dist"]))
    far = dist > float(opt["min_proj_dist"])
    keep &= near | far                                                      # dist < min_lidar_proj_dist: skipped (:400)
    with np.errstate(all="ignore"):
        sx = np.where(near, np.trunc(max_x), np.trunc(a_x * dist + b_x))
        sy = np.where(near, np.trunc(max_y), np.trunc(a_y * dist + b_y))
    keep &= np.isfinite(u0) & np.isfinite(v0) & (sx >= 0) & (sy >= 0)
    keep &= (u0 + sx >= 0) & (u0 - sx < img_w) & (v0 + sy >= 0) & (v0 - sy < img_h)
    rng = np.sqrt(((pc[0] * pc[0]).astype(F) + ((pc[1] * pc[1]).astype(F) + (pc[2] * pc[2]).astype(F)).astype(F)).astype(F)).astype(F)
    idx = np.nonzero(keep)[0]
    u0 = u0[idx].astype(np.int64); v0 = v0[idx].astype(np.int64)
    sx = sx[idx].astype(np.int64); sy = sy[idx].astype(np.int64)
    rb = rng[idx].view(np.uint32).astype(np.uint64)                        # range >= 0: bit order = value order
    key_first = (rb << np.uint64(32)) | idx.astype(np.uint64)              # low word: visit order
    key_low = (rb << np.uint64(32)) | order[idx].astype(np.uint64)         # low word: original index
    EMPTY = np.iinfo(np.uint64).max
    zf = np.full(img_h * img_w, EMPTY, dtype=np.uint64)
    zl = np.full(img_h * img_w, EMPTY, dtype=np.uint64)
    fm = None if feat_mask is None else np.asarray(feat_mask, dtype=bool).reshape(-1)
    for du in range(-int(sx.max(initial=-1)), int(sx.max(initial=-1)) + 1):
        for dv in range(-int(sy.max(initial=-1)), int(sy.max(initial=-1)) + 1):
            m = (abs(du) <= sx) & (abs(dv) <= sy)
            uu = u0 + du; vv = v0 + dv
            m &= (uu >= 0) & (uu < img_w) & (vv >= 0) & (vv < img_h)
            if not m.any():
                continue
            pix = vv[m] * img_w + uu[m]
            if fm is not None:
                f = fm[pix]
                pix = pix[f]
                np.minimum.at(zf, pix, key_first[m][f]); np.minimum.at(zl, pix, key_low[m][f])
            else:
                np.minimum.at(zf, pix, key_first[m]); np.minimum.at(zl, pix, key_low[m])
    has = zf != EMPTY
    rng_img = np.full(img_h * img_w, np.inf, dtype=F)
    rng_img[has] = (zf[has] >> np.uint64(32)).astype(np.uint32).view(F)
    first = np.full(img_h * img_w, -1, dtype=np.int64)
    first[has] = order[(zf[has] & np.uint64(0xffffffff)).astype(np.int64)]
    lowest = np.full(img_h * img_w, -1, dtype=np.int64)
    lowest[has] = (zl[has] & np.uint64(0xffffffff)).astype(np.int64)
    return dict(range=rng_img.reshape(img_h, img_w), first=first.reshape(img_h, img_w), lowest=lowest.reshape(img_h, img_w),
                packed_lowest=zl.reshape(img_h, img_w), n_sel=int(sel.sum()), n_pts=int(len(order)))


def feature_pixels(xy, scale, width, height):
    """(point2D.XY() * scale).cast<int>() and the bounds test of pcd_projection.cc:35-36; -1 for rejected keypoints."""
    xy = np.asarray(xy, dtype=np.float64)
    img_h, img_w = int(height * scale), int(width * scale)
    u = np.trunc(xy[:, 0] * scale).astype(np.int64)
    v = np.trunc(xy[:, 1] * scale).astype(np.int64)
    ok = (u >= 0) & (u < img_w) & (v >= 0) & (v < img_h)
    return np.where(ok, v * img_w + u, -1)
