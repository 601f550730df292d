"""ctypes binding over the C ABI in include/pcd_b200.h (test / benchmark harness side).

The product is libpcd_b200.so + the C++ shim under host/; this module only marshals numpy
arrays and raw device pointers into the C entry points.  It raises if the shared library is
missing -- there is no CPU fallback anywhere on the product path.
"""
import ctypes as C
import os
import weakref
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PCD_B200_LIB") or os.path.join(_HERE, "libpcd_b200.so")      # PCD_B200_LIB: another build of the same library (make dbg)

K_NN_KEYSORT, K_NN_SEARCH, K_ASSOC_RECORD, K_RESJAC, K_PROJECT, K_NN_FAR = 0, 1, 2, 3, 4, 5
TYPE_PROJ, TYPE_ICP, TYPE_ICP_GROUND, TYPE_NONE = 0, 1, 2, 255
ASSOC_LOCAL, ASSOC_GLOBAL, ASSOC_WHOLE_MAP = 0, 1, 2


class ProjOptions(C.Structure):
    """pcd_proj_options == PcdProjectionOptions (src/lidar/pcd_projection.h:31-47)."""
    _fields_ = [("depth_image_scale", C.c_double), ("max_proj_scale", C.c_int32), ("min_proj_scale", C.c_int32),
                ("min_proj_dist", C.c_double), ("submap_length", C.c_float), ("submap_width", C.c_float),
                ("submap_height", C.c_float), ("choose_meter", C.c_float), ("min_lidar_proj_dist", C.c_double)]

    @classmethod
    def default(cls, **kw):
        o = cls(0.2, 10, 2, 2.0, 1.0, 1.0, 1.0, 40.0, 0.5)
        for k, v in kw.items():
            setattr(o, k, v)
        return o


class RangeSweep(C.Structure):
    _fields_ = [("max_r", C.c_double), ("min_r", C.c_double), ("drop", C.c_double)]


class AssocOut(C.Structure):
    _fields_ = [("idx", C.c_void_p), ("accepted", C.c_void_p), ("type", C.c_void_p), ("rgb", C.c_void_p),
                ("xyz", C.c_void_p), ("abcd", C.c_void_p), ("dist", C.c_void_p), ("angle", C.c_void_p)]


class Image(C.Structure):
    _fields_ = [("qvec", C.c_double * 4), ("tvec", C.c_double * 3), ("cam", C.c_double * 8),
                ("width", C.c_int32), ("height", C.c_int32)]

    @classmethod
    def make(cls, qvec, tvec, cam, width, height):
        im = cls()
        im.qvec[:] = [float(x) for x in qvec]
        im.tvec[:] = [float(x) for x in tvec]
        im.cam[:] = [float(x) for x in cam]
        im.width, im.height = int(width), int(height)
        return im


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with __graft_entry__.build() "
                               "(there is no CPU fallback for the product path)")
        _lib = C.CDLL(LIB_PATH)
        _lib.pcd_last_error.restype = C.c_char_p
        _lib.pcd_last_error.argtypes = [C.c_void_p]
        _lib.pcd_cloud_size.restype = C.c_int64
        _lib.pcd_cloud_size.argtypes = [C.c_void_p]
        _lib.pcd_submap_cells.restype = C.c_int64
        _lib.pcd_submap_cells.argtypes = [C.c_void_p]
        _lib.pcd_ctx_launches.restype = C.c_int64
        _lib.pcd_ctx_launches.argtypes = [C.c_void_p]
        _lib.pcd_residual_jacobian_noalloc_max.restype = C.c_int64
        _lib.pcd_residual_jacobian_noalloc_max.argtypes = []
        _lib.pcd_cloud_device_ptr.restype = C.c_void_p
        _lib.pcd_cloud_device_ptr.argtypes = [C.c_void_p]
        for name in ("pcd_ctx_destroy", "pcd_cloud_destroy", "pcd_host_free", "pcd_free"):
            getattr(_lib, name).restype = None
            getattr(_lib, name).argtypes = [C.c_void_p]
    return _lib


class PcdError(RuntimeError):
    pass


def _ptr(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _f64(a, cols=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if cols is not None:
        a = a.reshape(-1, cols)
    return a


class Context:
    def __init__(self, device=0):
        self._h = C.c_void_p()
        rc = lib().pcd_ctx_create(C.c_int(device), C.byref(self._h))
        if rc:
            raise PcdError(f"pcd_ctx_create failed ({rc}): {lib().pcd_last_error(None).decode()}")
        self.device = device

    def check(self, rc, what):
        if rc:
            raise PcdError(f"{what} failed ({rc}): {lib().pcd_last_error(self._h).decode()}")

    @property
    def launches(self):
        return int(lib().pcd_ctx_launches(self._h))

    def close(self):
        if self._h:
            lib().pcd_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def profile(self, enable=True):
        self.check(lib().pcd_ctx_profile(self._h, C.c_int(int(enable))), "pcd_ctx_profile")

    def kernel_time(self, which):
        """(total milliseconds, launches) of kernel class `which` (K_*) since profile(True)."""
        ms = C.c_double(); n = C.c_int64()
        self.check(lib().pcd_ctx_kernel_time(self._h, C.c_int(which), C.byref(ms), C.byref(n)), "pcd_ctx_kernel_time")
        return ms.value, n.value

    def atomic_peak(self, words=1920 * 1080, sweeps=64):
        """Contention-free, warp-coalesced 64-bit atomicMin updates per second (the z-buffer's roofline denominator)."""
        v = C.c_double()
        self.check(lib().pcd_atomic_peak(self._h, C.c_int64(words), C.c_int(sweeps), C.byref(v)), "pcd_atomic_peak")
        return v.value

    def weights_from_type_dev(self, d_type, n, w_proj, w_icp, w_ground, d_w, stream=0):
        self.check(lib().pcd_weights_from_type_dev(self._h, C.c_void_p(d_type), C.c_int64(n), C.c_double(w_proj),
                                                   C.c_double(w_icp), C.c_double(w_ground), C.c_void_p(d_w),
                                                   C.c_void_p(stream)), "pcd_weights_from_type_dev")

    def pack_blocks_f32_dev(self, d_r, d_J, n, d_out, stream=0):
        """fp32 wire format {r, J0, J1, J2} of n residual blocks (device pointers), for the multi-GPU exchange."""
        self.check(lib().pcd_pack_blocks_f32_dev(self._h, C.c_void_p(d_r), C.c_void_p(d_J), C.c_int64(n), C.c_void_p(d_out),
                                                 C.c_void_p(stream)), "pcd_pack_blocks_f32_dev")

    # -- a19 --------------------------------------------------------------------------------
    def residual_jacobian(self, p, abcd, w, jac=True, want_sum=True, out=None):
        p, abcd, w = _f64(p, 3), _f64(abcd, 4), _f64(w).ravel()
        n = len(w)
        if out is not None:
            r, J = out
        else:
            r = np.empty(n); J = np.empty((n, 3)) if jac else None
        s = C.c_double(0.0)
        self.check(lib().pcd_residual_jacobian_batch(self._h, _ptr(p), _ptr(abcd), _ptr(w), C.c_int64(n), _ptr(r), _ptr(J),
                                                     C.byref(s) if want_sum else None), "pcd_residual_jacobian_batch")
        return r, J, s.value

    def residual_jacobian_dev(self, d_p, d_abcd, d_w, n, d_r, d_J, d_sum, stream=0):
        self.check(lib().pcd_residual_jacobian_batch_dev(self._h, C.c_void_p(d_p), C.c_void_p(d_abcd), C.c_void_p(d_w),
                                                         C.c_int64(n), C.c_void_p(d_r), C.c_void_p(d_J) if d_J else None,
                                                         C.c_void_p(d_sum) if d_sum else None, C.c_void_p(stream)),
                   "pcd_residual_jacobian_batch_dev")

    def residual_jacobian_dev_noalloc(self, d_p, d_abcd, d_w, n, d_r, d_J, d_sum, stream=0):
        """pcd_residual_jacobian_batch_dev_noalloc: the same evaluation with its scratch in a fixed device array."""
        self.check(lib().pcd_residual_jacobian_batch_dev_noalloc(self._h, C.c_void_p(d_p), C.c_void_p(d_abcd), C.c_void_p(d_w),
                                                                 C.c_int64(n), C.c_void_p(d_r), C.c_void_p(d_J) if d_J else None,
                                                                 C.c_void_p(d_sum) if d_sum else None, C.c_void_p(stream)),
                   "pcd_residual_jacobian_batch_dev_noalloc")

    def residual_pose_jacobian(self, p_cam, qvec, tvec, abcd, w):
        p_cam, qvec, tvec, abcd, w = _f64(p_cam, 3), _f64(qvec, 4), _f64(tvec, 3), _f64(abcd, 4), _f64(w).ravel()
        n = len(w)
        r = np.empty(n); Jp = np.empty((n, 3)); Jq = np.empty((n, 7))
        self.check(lib().pcd_residual_pose_jacobian_batch(self._h, _ptr(p_cam), _ptr(qvec), _ptr(tvec), _ptr(abcd), _ptr(w),
                                                          C.c_int64(n), _ptr(r), _ptr(Jp), _ptr(Jq)), "pcd_residual_pose_jacobian_batch")
        return r, Jp, Jq

    def weights_from_type(self, type_, w_proj, w_icp, w_ground, out=None):
        t = np.ascontiguousarray(type_, dtype=np.uint8)
        w = out if out is not None else np.empty(len(t))
        self.check(lib().pcd_weights_from_type(self._h, _ptr(t), C.c_int64(len(t)), C.c_double(w_proj), C.c_double(w_icp),
                                               C.c_double(w_ground), _ptr(w)), "pcd_weights_from_type")
        return w

    def filter_outliers(self, p, lidar_xyz, type_, max_proj, max_icp):
        p, l = _f64(p, 3), _f64(lidar_xyz, 3)
        t = np.ascontiguousarray(type_, dtype=np.uint8)
        e = np.empty(len(t), dtype=np.uint8); cnt = C.c_int64(0)
        self.check(lib().pcd_filter_outliers_batch(self._h, _ptr(p), _ptr(l), _ptr(t), C.c_int64(len(t)), C.c_double(max_proj),
                                                   C.c_double(max_icp), _ptr(e), C.byref(cnt)), "pcd_filter_outliers_batch")
        return e, cnt.value

    def lidar_point_ops(self, xyz, abcd, pt=None, normalize=True):
        """pcd_lidar_point_ops: (abcd_out, ComputeDist, ComputePointToPointDist, ComputeAngle) of n records."""
        xyz, abcd = _f64(xyz, 3), _f64(abcd, 4)
        n = len(xyz)
        p = None if pt is None else _f64(pt, 3)
        out = np.empty((n, 4)); dp = np.empty(n); dd = np.empty(n); an = np.empty(n)
        self.check(lib().pcd_lidar_point_ops(self._h, _ptr(xyz), _ptr(abcd), _ptr(p), C.c_int64(n), C.c_int(int(normalize)), _ptr(out),
                                             _ptr(dp), _ptr(dd), _ptr(an)), "pcd_lidar_point_ops")
        return out, dp, dd, an

    def match_proj(self, p, cand, cand_start):
        p, cand = _f64(p, 3), _f64(cand, 6)
        cs = np.ascontiguousarray(cand_start, dtype=np.int64)
        n = len(p)
        found = np.empty(n, dtype=np.uint8); xyz = np.empty((n, 3)); abcd = np.empty((n, 4)); dist = np.empty(n); ang = np.empty(n)
        self.check(lib().pcd_match_proj_batch(self._h, _ptr(p), _ptr(cand), _ptr(cs), C.c_int64(n), _ptr(found), _ptr(xyz),
                                              _ptr(abcd), _ptr(dist), _ptr(ang)), "pcd_match_proj_batch")
        return found, xyz, abcd, dist, ang


class Cloud:
    """Device-resident cloud + indices (PointCloudProcess' data: src/lidar/ply.h:22-40)."""

    def __init__(self, ctx, aos8=None, direction_trans=False, device_ptr=None, n=None):
        self.ctx = ctx
        self._h = C.c_void_p()
        if device_ptr is not None:
            ctx.check(lib().pcd_cloud_from_device(ctx._h, C.c_void_p(device_ptr), C.c_int64(n), C.c_int(int(direction_trans)),
                                                  C.byref(self._h)), "pcd_cloud_from_device")
        else:
            a = np.ascontiguousarray(aos8, dtype=np.float32).reshape(-1, 8)
            ctx.check(lib().pcd_cloud_from_host(ctx._h, _ptr(a), C.c_int64(len(a)), C.c_int(int(direction_trans)),
                                                C.byref(self._h)), "pcd_cloud_from_host")

    @classmethod
    def load_ply(cls, ctx, path, opt=None):
        self = cls.__new__(cls)
        self.ctx = ctx
        self._h = C.c_void_p()
        ctx.check(lib().pcd_cloud_load_ply(ctx._h, path.encode(), C.byref(opt) if opt is not None else None, C.byref(self._h)),
                  "pcd_cloud_load_ply")
        return self

    @classmethod
    def load_ply_cached(cls, ctx, path, opt=None):
        """pcd_cloud_load_ply_cached: (cloud, cache_hit)."""
        h = C.c_void_p(); hit = C.c_int()
        ctx.check(lib().pcd_cloud_load_ply_cached(ctx._h, path.encode(), C.byref(opt) if opt is not None else None, C.byref(h), C.byref(hit)),
                  "pcd_cloud_load_ply_cached")
        c = cls.__new__(cls)
        c.ctx, c._h, c.opt = ctx, h, opt
        return c, bool(hit.value)

    def close(self):
        if self._h:
            lib().pcd_cloud_destroy(self._h)
            self._h = C.c_void_p()

    def __len__(self):
        return int(lib().pcd_cloud_size(self._h))

    @property
    def device_ptr(self):
        return lib().pcd_cloud_device_ptr(self._h)

    def download(self):
        out = np.empty((len(self), 8), dtype=np.float32)
        self.ctx.check(lib().pcd_cloud_download(self._h, _ptr(out)), "pcd_cloud_download")
        return out

    def build_index(self, cell_size=0.0):
        self.ctx.check(lib().pcd_index_build(self._h, C.c_float(cell_size)), "pcd_index_build")
        return self.index_info()

    def index_info(self):
        cs = C.c_float(); dims = (C.c_int32 * 3)(); org = (C.c_float * 3)(); nc = C.c_int64()
        self.ctx.check(lib().pcd_index_info(self._h, C.byref(cs), dims, org, C.byref(nc)), "pcd_index_info")
        return {"cell": cs.value, "dims": list(dims), "origin": list(org), "n_cells": nc.value}

    def build_submap(self, opt):
        self.opt = opt
        self.ctx.check(lib().pcd_submap_build(self._h, C.byref(opt)), "pcd_submap_build")
        return int(lib().pcd_submap_cells(self._h))

    # -- a4/a5 ------------------------------------------------------------------------------
    def nn(self, q, radius=None):
        q = _f64(q, 3); n = len(q)
        r = None if radius is None else _f64(np.broadcast_to(radius, (n,))).ravel()
        idx = np.empty(n, dtype=np.int32); d2 = np.empty(n, dtype=np.float32)
        self.ctx.check(lib().pcd_nn_batch(self._h, _ptr(q), _ptr(r), C.c_int64(n), _ptr(idx), _ptr(d2)), "pcd_nn_batch")
        return idx, d2

    def nn_dev(self, d_q, d_radius, n, d_idx, d_d2, stream=0):
        self.ctx.check(lib().pcd_nn_batch_dev(self._h, C.c_void_p(d_q), C.c_void_p(d_radius) if d_radius else None, C.c_int64(n),
                                              C.c_void_p(d_idx), C.c_void_p(d_d2) if d_d2 else None, C.c_void_p(stream)),
                       "pcd_nn_batch_dev")

    def search_nn(self, q):
        q = _f64(q, 3); n = len(q)
        l = np.empty((n, 6)); ok = np.empty(n, dtype=np.uint8)
        self.ctx.check(lib().pcd_search_nn_batch(self._h, _ptr(q), C.c_int64(n), _ptr(l), _ptr(ok)), "pcd_search_nn_batch")
        return l, ok.astype(bool)

    # -- a6 ---------------------------------------------------------------------------------
    def associate(self, p, range_=None, opt_num=None, sweep=None, variant=ASSOC_LOCAL, fields=None, out=None):
        p = _f64(p, 3); n = len(p)
        allf = ("idx", "accepted", "type", "rgb", "xyz", "abcd", "dist", "angle")
        fields = allf if fields is None else fields
        shapes = {"idx": ((n,), np.int32), "accepted": ((n,), np.uint8), "type": ((n,), np.uint8), "rgb": ((n, 3), np.uint8),
                  "xyz": ((n, 3), np.float64), "abcd": ((n, 4), np.float64), "dist": ((n,), np.float64), "angle": ((n,), np.float64)}
        res = out if out is not None else {f: np.empty(*shapes[f]) for f in fields}
        out = AssocOut(*[res[f].ctypes.data if f in res else None for f in allf])
        r = None if range_ is None else _f64(np.broadcast_to(range_, (n,))).ravel()
        on = None if opt_num is None else np.ascontiguousarray(opt_num, dtype=np.int32)
        sw = None if sweep is None else RangeSweep(*sweep)
        self.ctx.check(lib().pcd_associate_batch(self._h, _ptr(p), _ptr(r), _ptr(on), C.byref(sw) if sw is not None else None,
                                                 C.c_int64(n), C.c_int(variant), C.byref(out)), "pcd_associate_batch")
        return res

    def associate_evaluate(self, p, weights, range_=None, opt_num=None, sweep=None, variant=ASSOC_LOCAL, fields=None, out=None,
                           out_rj=None, jac=True):
        """pcd_associate_evaluate_batch: association records + (r, J, sum r^2) of the accepted matches in one
        pipelined host-pointer pass.  weights = (w_proj, w_icp, w_ground).  Returns (records, r, J, sum_sq)."""
        p = _f64(p, 3); n = len(p)
        allf = ("idx", "accepted", "type", "rgb", "xyz", "abcd", "dist", "angle")
        fields = allf if fields is None else fields
        shapes = {"idx": ((n,), np.int32), "accepted": ((n,), np.uint8), "type": ((n,), np.uint8), "rgb": ((n, 3), np.uint8),
                  "xyz": ((n, 3), np.float64), "abcd": ((n, 4), np.float64), "dist": ((n,), np.float64), "angle": ((n,), np.float64)}
        res = out if out is not None else {f: np.empty(*shapes[f]) for f in fields}
        o = AssocOut(*[res[f].ctypes.data if f in res else None for f in allf])
        r = None if range_ is None else _f64(np.broadcast_to(range_, (n,))).ravel()
        on = None if opt_num is None else np.ascontiguousarray(opt_num, dtype=np.int32)
        sw = None if sweep is None else RangeSweep(*sweep)
        rr, J = out_rj if out_rj is not None else (np.empty(n), np.empty((n, 3)) if jac else None)
        s = C.c_double(0.0)
        self.ctx.check(lib().pcd_associate_evaluate_batch(self._h, _ptr(p), _ptr(r), _ptr(on), C.byref(sw) if sw is not None else None,
                                                          C.c_int64(n), C.c_int(variant), C.c_double(weights[0]), C.c_double(weights[1]),
                                                          C.c_double(weights[2]), C.byref(o), _ptr(rr), _ptr(J), C.byref(s)),
                       "pcd_associate_evaluate_batch")
        return res, rr, J, s.value

    def associate_evaluate_dev(self, d_p, d_range, d_opt_num, sweep, n, variant, weights, out_ptrs, d_r, d_J, d_sum, stream=0):
        """pcd_associate_evaluate_batch_dev: device pointers (ints); out_ptrs: dict field -> device pointer."""
        allf = ("idx", "accepted", "type", "rgb", "xyz", "abcd", "dist", "angle")
        out = AssocOut(*[out_ptrs.get(f) for f in allf])
        sw = None if sweep is None else RangeSweep(*sweep)
        self.ctx.check(lib().pcd_associate_evaluate_batch_dev(
            self._h, C.c_void_p(d_p), C.c_void_p(d_range) if d_range else None, C.c_void_p(d_opt_num) if d_opt_num else None,
            C.byref(sw) if sw is not None else None, C.c_int64(n), C.c_int(variant), C.c_double(weights[0]), C.c_double(weights[1]),
            C.c_double(weights[2]), C.byref(out), C.c_void_p(d_r), C.c_void_p(d_J) if d_J else None,
            C.c_void_p(d_sum) if d_sum else None, C.c_void_p(stream)), "pcd_associate_evaluate_batch_dev")

    def associate_dev(self, d_p, d_range, d_opt_num, sweep, n, variant, out_ptrs, stream=0):
        """out_ptrs: dict field -> device pointer (int)."""
        allf = ("idx", "accepted", "type", "rgb", "xyz", "abcd", "dist", "angle")
        out = AssocOut(*[out_ptrs.get(f) for f in allf])
        sw = None if sweep is None else RangeSweep(*sweep)
        self.ctx.check(lib().pcd_associate_batch_dev(self._h, C.c_void_p(d_p), C.c_void_p(d_range) if d_range else None,
                                                     C.c_void_p(d_opt_num) if d_opt_num else None,
                                                     C.byref(sw) if sw is not None else None, C.c_int64(n), C.c_int(variant),
                                                     C.byref(out), C.c_void_p(stream)), "pcd_associate_batch_dev")

    # -- a10-a15 ----------------------------------------------------------------------------
    def project_image(self, image, coef, feat_xy=None, dense=False):
        coef_c = (C.c_double * 4)(*coef)
        f = None if feat_xy is None else _f64(feat_xy, 2)
        nf = 0 if f is None else len(f)
        widx = np.empty(nf, dtype=np.int32); wr = np.empty(nf, dtype=np.float32)
        scale = self.opt.depth_image_scale
        w, h = int(image.width * scale), int(image.height * scale)
        z = np.empty((h, w), dtype=np.uint64) if dense else None
        stats = np.zeros(4, dtype=np.int64)
        self.ctx.check(lib().pcd_project_image(self._h, C.byref(image), coef_c, _ptr(f), C.c_int64(nf), _ptr(widx), _ptr(wr),
                                               _ptr(z), _ptr(stats)), "pcd_project_image")
        return widx, wr, z, stats

    def project_images(self, images, coef, feat_xy, feat_start):
        arr = (Image * len(images))(*images)
        coef_c = (C.c_double * 4)(*coef)
        f = _f64(feat_xy, 2); fs = np.ascontiguousarray(feat_start, dtype=np.int64)
        nf = len(f)
        widx = np.empty(nf, dtype=np.int32); wr = np.empty(nf, dtype=np.float32); stats = np.zeros(4, dtype=np.int64)
        self.ctx.check(lib().pcd_project_images(self._h, arr, C.c_int64(len(images)), coef_c, _ptr(f), _ptr(fs), _ptr(widx),
                                                _ptr(wr), _ptr(stats)), "pcd_project_images")
        return widx, wr, stats

    def ray_plane(self, image, feat_xy, winner_idx):
        f = _f64(feat_xy, 2); w = np.ascontiguousarray(winner_idx, dtype=np.int32); n = len(w)
        m = np.empty(n, dtype=np.uint8); xyz = np.empty((n, 3))
        self.ctx.check(lib().pcd_ray_plane_batch(self._h, C.byref(image), _ptr(f), _ptr(w), C.c_int64(n), _ptr(m), _ptr(xyz)),
                       "pcd_ray_plane_batch")
        return m.astype(bool), xyz


def project2image_select(track_image, image_id, pair_corrs, threshold, cached=None):
    """pcd_project2image_select: flags of the track elements whose images must be projected now, and their count."""
    t = np.ascontiguousarray(track_image, dtype=np.uint32); pc = np.ascontiguousarray(pair_corrs, dtype=np.int64)
    c = None if cached is None else np.ascontiguousarray(cached, dtype=np.uint8)
    sel = np.zeros(len(t), dtype=np.uint8)
    lib().pcd_project2image_select.restype = C.c_int64
    n = lib().pcd_project2image_select(_ptr(t), C.c_int64(len(t)), C.c_uint32(image_id), _ptr(pc), C.c_int(threshold), _ptr(c), _ptr(sel))
    if n < 0:
        raise PcdError("pcd_project2image_select: bad argument")
    return sel, int(n)


def file_key(path):
    k = C.c_uint64()
    if lib().pcd_file_key(path.encode(), C.byref(k)):
        raise PcdError(lib().pcd_last_error(None).decode())
    return k.value


def sidecar_write(path, key, aos8, direction_trans=True):
    a = np.ascontiguousarray(aos8, dtype=np.float32).reshape(-1, 8)
    if lib().pcd_sidecar_write(path.encode(), C.c_uint64(key), C.c_int(int(direction_trans)), _ptr(a), C.c_int64(len(a))):
        raise PcdError(lib().pcd_last_error(None).decode())


def sidecar_read(path, key, direction_trans=True):
    """The cloud of a sidecar, or None when it is missing / damaged / belongs to another source file."""
    p = C.POINTER(C.c_float)(); n = C.c_int64()
    if lib().pcd_sidecar_read(path.encode(), C.c_uint64(key), C.c_int(int(direction_trans)), C.byref(p), C.byref(n)):
        return None
    out = np.ctypeslib.as_array(p, shape=(n.value * 8,)).copy().reshape(-1, 8) if n.value else np.zeros((0, 8), np.float32)
    lib().pcd_free(p)
    return out


def host_bind_near(device):
    """Binds the calling thread (and what it allocates next) to the CPUs next to `device`; > 0: CPUs bound to, < 0: unchanged."""
    return int(lib().pcd_host_bind_near(C.c_int(device)))


def pinned_empty(shape, dtype):
    """numpy array over cudaMallocHost memory (pcd_host_alloc); freed when the array is collected."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = C.c_void_p()
    rc = lib().pcd_host_alloc(C.byref(p), C.c_int64(max(n, 1)))
    if rc:
        raise PcdError(f"pcd_host_alloc failed ({rc}): {lib().pcd_last_error(None).decode()}")
    buf = (C.c_char * max(n, 1)).from_address(p.value)
    a = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    weakref.finalize(buf, lambda q=p.value: lib().pcd_host_free(C.c_void_p(q)))
    return a



def splat_coef(cam, opt):
    cam_c = (C.c_double * 8)(*[float(x) for x in cam]); out = (C.c_double * 4)()
    rc = lib().pcd_splat_coef(cam_c, C.byref(opt), out)
    if rc:
        raise PcdError("pcd_splat_coef failed")
    return list(out)


def ply_read(path):
    p = C.POINTER(C.c_float)(); n = C.c_int64()
    rc = lib().pcd_ply_read(path.encode(), C.byref(p), C.byref(n))
    if rc:
        raise PcdError(f"pcd_ply_read failed ({rc}): {lib().pcd_last_error(None).decode()}")
    a = np.ctypeslib.as_array(p, shape=(n.value, 8)).copy() if n.value else np.zeros((0, 8), np.float32)
    lib().pcd_free(p)
    return a


def ply_write(path, aos8, binary=True):
    a = np.ascontiguousarray(aos8, dtype=np.float32).reshape(-1, 8)
    rc = lib().pcd_ply_write(path.encode(), _ptr(a), C.c_int64(len(a)), C.c_int(int(binary)))
    if rc:
        raise PcdError(f"pcd_ply_write failed ({rc}): {lib().pcd_last_error(None).decode()}")


def pose_ply_read(path):
    """pcd_pose_ply_read: returns (image_ids int64 [n], poses float64 [n, 7] = tx ty tz qw qx qy qz)."""
    n = C.c_int64(0)
    rc = lib().pcd_pose_ply_read(path.encode(), None, None, C.c_int64(0), C.byref(n))
    if rc:
        raise PcdError(lib().pcd_last_error(None).decode())
    ids = np.empty(n.value, dtype=np.int64); tq = np.empty((n.value, 7), dtype=np.float64)
    rc = lib().pcd_pose_ply_read(path.encode(), _ptr(ids), _ptr(tq), C.c_int64(n.value), C.byref(n))
    if rc:
        raise PcdError(lib().pcd_last_error(None).decode())
    return ids, tq


def pose_ply_write(path, xyzrpy):
    a = _f64(xyzrpy, 6)
    rc = lib().pcd_pose_ply_write(path.encode(), _ptr(a), C.c_int64(len(a)))
    if rc:
        raise PcdError(lib().pcd_last_error(None).decode())


class Group:
    """pcd_group_*: the path sharded over several devices of one node behind the C ABI (a device may repeat)."""

    def __init__(self, devices):
        self.devices = list(devices)
        arr = (C.c_int * len(self.devices))(*self.devices)
        self._h = C.c_void_p()
        rc = lib().pcd_group_create(arr, C.c_int(len(self.devices)), C.byref(self._h))
        if rc:
            raise PcdError(f"pcd_group_create: {lib().pcd_last_error(None).decode()}")
        lib().pcd_group_last_error.restype = C.c_char_p
        lib().pcd_group_last_error.argtypes = [C.c_void_p]
        lib().pcd_group_destroy.restype = None
        lib().pcd_group_cloud_destroy.restype = None
        self._gc = None
        self.opt = None

    def _check(self, rc, what):
        if rc:
            raise PcdError(f"{what}: {lib().pcd_group_last_error(self._h).decode()}")

    def cloud_from_host(self, aos8, direction_trans=False, cell_size=0.0, opt=None):
        a = np.ascontiguousarray(aos8, dtype=np.float32).reshape(-1, 8)
        gc = C.c_void_p()
        self._check(lib().pcd_group_cloud_from_host(self._h, _ptr(a), C.c_int64(len(a)), C.c_int(int(direction_trans)), C.c_float(cell_size),
                                                    C.byref(opt) if opt is not None else None, C.byref(gc)), "pcd_group_cloud_from_host")
        self._gc, self.opt = gc, opt

    def associate_evaluate(self, p, weights, range_=None, opt_num=None, sweep=None, variant=ASSOC_LOCAL):
        p = _f64(p, 3); n = len(p)
        rng = None if range_ is None else _f64(range_).ravel()
        on = None if opt_num is None else np.ascontiguousarray(opt_num, dtype=np.int32)
        sw = None if sweep is None else RangeSweep(*sweep)
        o = {"idx": np.empty(n, np.int32), "accepted": np.empty(n, np.uint8), "type": np.empty(n, np.uint8), "rgb": np.empty((n, 3), np.uint8),
             "xyz": np.empty((n, 3)), "abcd": np.empty((n, 4)), "dist": np.empty(n), "angle": np.empty(n)}
        out = AssocOut(*[_ptr(o[k]) for k in ("idx", "accepted", "type", "rgb", "xyz", "abcd", "dist", "angle")])
        r = np.empty(n); J = np.empty((n, 3)); s = C.c_double()
        self._check(lib().pcd_group_associate_evaluate_batch(self._gc, _ptr(p), _ptr(rng), _ptr(on), C.byref(sw) if sw is not None else None,
                                                             C.c_int64(n), C.c_int(variant), C.c_double(weights[0]), C.c_double(weights[1]),
                                                             C.c_double(weights[2]), C.byref(out), _ptr(r), _ptr(J), C.byref(s)),
                    "pcd_group_associate_evaluate_batch")
        return o, r, J, s.value

    def project_images(self, images, coef, feat_xy, feat_start):
        arr = (Image * len(images))(*images)
        coef_c = (C.c_double * 4)(*coef)
        f = _f64(feat_xy, 2); fs = np.ascontiguousarray(feat_start, dtype=np.int64)
        nf = len(f)
        widx = np.empty(nf, dtype=np.int32); wr = np.empty(nf, dtype=np.float32); stats = np.zeros(4, dtype=np.int64)
        self._check(lib().pcd_group_project_images(self._gc, arr, C.c_int64(len(images)), coef_c, _ptr(f), _ptr(fs), _ptr(widx), _ptr(wr),
                                                   _ptr(stats)), "pcd_group_project_images")
        return widx, wr, stats

    def close(self):
        if self._gc is not None:
            lib().pcd_group_cloud_destroy(self._gc); self._gc = None
        if self._h:
            lib().pcd_group_destroy(self._h); self._h = None
