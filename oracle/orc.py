"""ctypes binding of the CPU oracle (oracle/pcd_oracle.c).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  PARITY UNPINNED: see the header of pcd_oracle.c.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liborc.so")
_lib = None


class ProjOptions(C.Structure):
    _fields_ = [("depth_image_scale", C.c_double), ("max_proj_scale", C.c_int), ("min_proj_scale", C.c_int),
                ("min_proj_dist", C.c_double), ("submap_length", C.c_float), ("submap_width", C.c_float),
                ("submap_height", C.c_float), ("choose_meter", C.c_float), ("min_lidar_proj_dist", C.c_double)]

    @classmethod
    def default(cls, **kw):
        o = cls(0.2, 10, 2, 2.0, 1.0, 1.0, 1.0, 40.0, 0.5)
        for k, v in kw.items():
            setattr(o, k, v)
        return o


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = C.CDLL(LIB_PATH)
        _lib.orc_kdtree_build.restype = C.c_void_p
        _lib.orc_submap_build.restype = C.c_void_p
        _lib.orc_direction_trans.restype = C.c_int64
        _lib.orc_submap_cells.restype = C.c_int64
        _lib.orc_select_cells.restype = C.c_int64
        _lib.orc_filter_outliers.restype = C.c_int64
    return _lib


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _f64(a, cols=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a.reshape(-1, cols) if cols else a


def direction_trans(aos8):
    a = np.ascontiguousarray(aos8, dtype=np.float32).reshape(-1, 8)
    out = np.empty_like(a)
    m = lib().orc_direction_trans(_p(a), C.c_int64(len(a)), _p(out))
    return out[:m].copy()


def nn_brute(cloud, q):
    cloud = np.ascontiguousarray(cloud, dtype=np.float32).reshape(-1, 8)
    q = _f64(q, 3)
    idx = np.empty(len(q), dtype=np.int32); d2 = np.empty(len(q), dtype=np.float32)
    lib().orc_nn_brute(_p(cloud), C.c_int64(len(cloud)), _p(q), C.c_int64(len(q)), _p(idx), _p(d2))
    return idx, d2


def search_range(opt_num, max_r=1.5, min_r=0.2, drop=0.1):
    on = np.ascontiguousarray(opt_num, dtype=np.int32)
    out = np.empty(len(on))
    lib().orc_search_range(_p(on), C.c_int64(len(on)), C.c_double(max_r), C.c_double(min_r), C.c_double(drop), _p(out))
    return out


class KdTree:
    """Restated FLANN KDTreeSingleIndex (leaf 15): the timed CPU baseline and fast exact oracle."""

    def __init__(self, cloud, leaf_max=15):
        self.cloud = np.ascontiguousarray(cloud, dtype=np.float32).reshape(-1, 8)
        self._h = C.c_void_p(lib().orc_kdtree_build(_p(self.cloud), C.c_int64(len(self.cloud)), C.c_int(leaf_max)))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_kdtree_free(self._h)
            self._h = None

    def nn(self, q, threads=0):
        q = _f64(q, 3)
        idx = np.empty(len(q), dtype=np.int32); d2 = np.empty(len(q), dtype=np.float32)
        lib().orc_kdtree_nn(self._h, _p(q), C.c_int64(len(q)), _p(idx), _p(d2), C.c_int(threads))
        return idx, d2

    def search_nn(self, p):
        p = _f64(p).ravel(); l = np.empty(6); idx = C.c_int32()
        ok = lib().orc_search_nn(self._h, _p(p), _p(l), C.byref(idx))
        return bool(ok), l, idx.value

    def associate(self, p, range_=None, variant=0, threads=0):
        p = _f64(p, 3); n = len(p)
        r = None if range_ is None else _f64(np.broadcast_to(range_, (n,))).ravel()
        res = {"idx": np.empty(n, np.int32), "accepted": np.empty(n, np.uint8), "type": np.empty(n, np.uint8),
               "rgb": np.empty((n, 3), np.uint8), "xyz": np.empty((n, 3)), "abcd": np.empty((n, 4)),
               "dist": np.empty(n), "angle": np.empty(n)}
        lib().orc_associate(self._h, _p(p), _p(r), C.c_int64(n), C.c_int(variant), _p(res["idx"]), _p(res["accepted"]),
                            _p(res["type"]), _p(res["rgb"]), _p(res["xyz"]), _p(res["abcd"]), _p(res["dist"]), _p(res["angle"]),
                            C.c_int(threads))
        return res


def records_from_idx(cloud, nn_idx, p, range_=None, variant=0):
    cloud = np.ascontiguousarray(cloud, dtype=np.float32).reshape(-1, 8)
    p = _f64(p, 3); n = len(p)
    nn_idx = np.ascontiguousarray(nn_idx, dtype=np.int32)
    r = None if range_ is None else _f64(np.broadcast_to(range_, (n,))).ravel()
    res = {"accepted": np.empty(n, np.uint8), "type": np.empty(n, np.uint8), "rgb": np.empty((n, 3), np.uint8),
           "xyz": np.empty((n, 3)), "abcd": np.empty((n, 4)), "dist": np.empty(n), "angle": np.empty(n)}
    lib().orc_records_from_idx(_p(cloud), _p(nn_idx), _p(p), _p(r), C.c_int64(n), C.c_int(variant), _p(res["accepted"]),
                               _p(res["type"]), _p(res["rgb"]), _p(res["xyz"]), _p(res["abcd"]), _p(res["dist"]), _p(res["angle"]))
    return res


def match_proj(p, cand):
    p = _f64(p).ravel(); cand = _f64(cand, 6)
    xyz = np.empty(3); abcd = np.empty(4); d = C.c_double(); a = C.c_double()
    ok = lib().orc_match_proj(_p(p), _p(cand), C.c_int(len(cand)), _p(xyz), _p(abcd), C.byref(d), C.byref(a))
    return bool(ok), xyz, abcd, d.value, a.value


def resjac(p, abcd, w, jac=True, threads=1):
    p, abcd, w = _f64(p, 3), _f64(abcd, 4), _f64(w).ravel()
    n = len(w)
    r = np.empty(n); J = np.empty((n, 3)) if jac else None
    if threads == 1:
        lib().orc_resjac(_p(p), _p(abcd), _p(w), C.c_int64(n), _p(r), _p(J))
    else:
        lib().orc_resjac_mt(_p(p), _p(abcd), _p(w), C.c_int64(n), _p(r), _p(J), C.c_int(threads))
    return r, J


def filter_outliers(p, lidar_xyz, type_, max_proj, max_icp):
    p, l = _f64(p, 3), _f64(lidar_xyz, 3)
    t = np.ascontiguousarray(type_, dtype=np.uint8)
    e = np.empty(len(t), dtype=np.uint8)
    cnt = lib().orc_filter_outliers(_p(p), _p(l), _p(t), C.c_int64(len(t)), C.c_double(max_proj), C.c_double(max_icp), _p(e))
    return e, int(cnt)


class Submap:
    def __init__(self, cloud, opt):
        self.cloud = np.ascontiguousarray(cloud, dtype=np.float32).reshape(-1, 8)
        self.opt = opt
        self._h = C.c_void_p(lib().orc_submap_build(_p(self.cloud), C.c_int64(len(self.cloud)), C.byref(opt)))
        self.n_cells = int(lib().orc_submap_cells(self._h))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_submap_free(self._h)
            self._h = None

    def pyramid(self, qvec, tvec, cam, w, h):
        q = _f64(qvec).ravel(); t = _f64(tvec).ravel(); c = _f64(cam).ravel()
        planes = np.empty((5, 4), np.float32); verts = np.empty((5, 3), np.float32)
        lib().orc_pyramid(_p(q), _p(t), _p(c), C.c_int(w), C.c_int(h), C.byref(self.opt), _p(planes), _p(verts))
        return planes, verts

    def select(self, planes):
        sel = np.empty(self.n_cells, dtype=np.uint8)
        cnt = lib().orc_select_cells(self._h, _p(np.ascontiguousarray(planes, np.float32)), _p(sel))
        return sel, int(cnt)

    def project(self, qvec, tvec, cam, w, h, coef, feat_mask=None):
        """Sequential z-buffer for one image.  Returns (zbuf uint64 [img_h, img_w], counters)."""
        planes, _ = self.pyramid(qvec, tvec, cam, w, h)
        sel, _ = self.select(planes)
        s = self.opt.depth_image_scale
        ih, iw = int(h * s), int(w * s)
        z = np.full((ih, iw), np.iinfo(np.uint64).max, dtype=np.uint64)
        q = _f64(qvec).ravel(); t = _f64(tvec).ravel(); c = _f64(cam).ravel(); cf = _f64(coef).ravel()
        fm = None if feat_mask is None else np.ascontiguousarray(feat_mask, dtype=np.uint8)
        counters = np.zeros(2, dtype=np.int64)
        lib().orc_project(self._h, _p(sel), _p(q), _p(t), _p(c), C.c_int(w), C.c_int(h), _p(cf), _p(fm), _p(z), _p(counters))
        return z, counters, int(sel.sum())


def splat_coef(cam, opt):
    c = _f64(cam).ravel(); out = np.empty(4)
    lib().orc_splat_coef(_p(c), C.byref(opt), _p(out))
    return out


def feature_pixels(xy, scale, w, h):
    xy = _f64(xy, 2); px = np.empty(len(xy), dtype=np.int64)
    lib().orc_feature_pixels(_p(xy), C.c_int64(len(xy)), C.c_double(scale), C.c_int(w), C.c_int(h), _p(px))
    return px


def ray_plane(cloud, gi, cam, u, v):
    cloud = np.ascontiguousarray(cloud, dtype=np.float32).reshape(-1, 8)
    c = _f64(cam).ravel(); out = np.empty(3)
    lib().orc_ray_plane(_p(cloud), C.c_int32(gi), _p(c), C.c_double(u), C.c_double(v), _p(out))
    return out


def max_threads():
    return int(lib().orc_max_threads())


def project2image(track_image, image_id, pair_corrs, threshold, cached_ids=()):
    """Which track elements' images Project2Image hands to SetNewImage (src/optim/bundle_adjustment.cc:241-281)."""
    t = np.ascontiguousarray(track_image, dtype=np.uint32); pc = np.ascontiguousarray(pair_corrs, dtype=np.int64)
    c = np.ascontiguousarray(list(cached_ids), dtype=np.uint32); sel = np.zeros(len(t), dtype=np.uint8)
    lib().orc_project2image.restype = C.c_int64
    n = lib().orc_project2image(_p(t), C.c_int64(len(t)), C.c_uint32(image_id), _p(pc), C.c_int(threshold), _p(c) if len(c) else None,
                                C.c_int64(len(c)), _p(sel))
    return sel, int(n)
