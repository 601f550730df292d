#!/usr/bin/env python
"""Generates tests/golden/assoc_small.npz and tests/golden/zbuf_small.npz.

    python tests/golden/make_golden.py            (run in the build container; needs no GPU)

The reference (Wangshihu12/colmap-pcd) cannot be compiled or imported in this image (PCL, FLANN,
Eigen and Ceres are absent) and holds no golden vectors for src/lidar, so these fixtures are NOT
outputs of the reference: parity stays "unpinned".  What they give instead:

  assoc_small.npz  an INDEPENDENT numpy restatement, written from the reference's formulas and not
                   from oracle/pcd_oracle.c -- brute-force float32 argmin for the NN
                   (src/lidar/kdtree.cc:10-21, distance order of FLANN's L2_Simple), numpy float64
                   for the LidarPoint record (src/optim/bundle_adjustment.cc:358-410,
                   src/lidar/lidar_point.cc:21-50) and the Jet<double,3> rules for the cost
                   function (src/base/cost_functions.h:204-232).  Both the C oracle (CPU suite) and
                   the CUDA path (GPU suite) are checked against it.
  zbuf_small.npz   an INDEPENDENT numpy restatement of the depth projection too
                   (tests/golden/proj_numpy.py, written from src/lidar/pcd_projection.cc:258-594 and
                   pcd_projection.h:67-158, not from the oracle): frustum planes, cell selection,
                   float32 transform, fp64 distortion, rounding, depth-dependent splat incl. the b_y
                   quirk, min range with the first-visited / lowest-index tie rule.  Four images: two at
                   the reference's depth_image_scale 0.2 (4032x3024, distorted), two at 1.0
                   (1920x1080), one of them with an identity pose and hand-placed points at
                   min_proj_dist, one float above it, nearer than min_lidar_proj_dist and behind the
                   camera.  Dense packed z-buffers (range bits << 32 | winner) and feature winners.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

SEED = 20261104


def direction_trans(raw):
    """src/lidar/ply.cc:33-57: (x,y,z) <- (-y,-z,x), same for the normal, NaN rows dropped."""
    out = np.zeros_like(raw)
    out[:, 0], out[:, 1], out[:, 2] = -raw[:, 1], -raw[:, 2], raw[:, 0]
    out[:, 4], out[:, 5], out[:, 6] = -raw[:, 5], -raw[:, 6], raw[:, 4]
    keep = ~np.isnan(out[:, [0, 1, 2, 4, 5, 6]]).any(axis=1)
    return np.ascontiguousarray(out[keep])


def nn_brute_f32(cloud, q):
    """argmin of ((dx*dx + dy*dy) + dz*dz) in float32 on the float-cast query; first (= lowest) index on ties."""
    qf = q.astype(np.float32)
    idx = np.empty(len(q), np.int32)
    d2 = np.empty(len(q), np.float32)
    x, y, z = cloud[:, 0], cloud[:, 1], cloud[:, 2]
    for i in range(len(q)):
        dx, dy, dz = qf[i, 0] - x, qf[i, 1] - y, qf[i, 2] - z
        d = (dx * dx + dy * dy) + dz * dz
        assert d.dtype == np.float32
        idx[i] = int(np.argmin(d))
        d2[i] = d[idx[i]]
    return idx, d2


def records(cloud, idx, p, rng, variant):
    """LidarPoint record of the three twin loops (0 local, 1 global, 2 whole map)."""
    l = cloud[idx, 0:3].astype(np.float64)
    n = cloud[idx, 4:7].astype(np.float64)
    norm = np.sqrt(n[:, 0] ** 2 + n[:, 1] ** 2 + n[:, 2] ** 2)           # lidar_point.cc:43
    ok = ~(norm < 1e-6)                                                   # ply.cc:100-101 (NaN rows were dropped)
    with np.errstate(all="ignore"):
        abc = n / norm[:, None]
        d = 0.0 - abc[:, 0] * l[:, 0] - abc[:, 1] * l[:, 1] - abc[:, 2] * l[:, 2]   # :48, left to right
        ground = (np.abs(n[:, 1] / n[:, 0]) > 10) & (np.abs(n[:, 1] / n[:, 2]) > 10)  # bundle_adjustment.cc:381
        v = l - p
        dist = np.sqrt(v[:, 0] ** 2 + v[:, 1] ** 2 + v[:, 2] ** 2)       # lidar_point.cc:28
        pdist = np.abs(abc[:, 0] * p[:, 0] + abc[:, 1] * p[:, 1] + abc[:, 2] * p[:, 2] + d)
        w = p - l
        angle = np.abs(abc[:, 0] * w[:, 0] + abc[:, 1] * w[:, 1] + abc[:, 2] * w[:, 2]) / np.sqrt((w ** 2).sum(1))
    if variant == 2:      # controllers/bundle_adjustment.cc:141-183: fixed thresholds
        acc = ok & ~(pdist > 1.0) & ~(dist > 2.0)
    else:                 # optim/bundle_adjustment.cc:400 / sfm/incremental_mapper.cc:1458
        acc = ok & ~(dist > rng)
    typ = np.where(ground, 2, 1).astype(np.uint8)
    return dict(accepted=acc.astype(np.uint8), type=typ, xyz=l, abcd=np.concatenate([abc, d[:, None]], 1), dist=dist,
                angle=angle)


def resjac(p, abcd, w):
    """Jet<double,3> evaluation of cost_functions.h:226-229: u = 0 - (p0 a + p1 b + p2 c + d); r = w sqrt(u u)."""
    u = 0.0 - (p[:, 0] * abcd[:, 0] + p[:, 1] * abcd[:, 1] + p[:, 2] * abcd[:, 2] + abcd[:, 3])
    du = -abcd[:, 0:3]                                   # d u / d p
    uu = u * u
    duu = u[:, None] * du + du * u[:, None]              # product rule
    s = np.sqrt(uu)
    with np.errstate(all="ignore"):
        ds = duu * (1.0 / (2.0 * s))[:, None]            # sqrt rule
    return w * s, w[:, None] * ds


def main():
    from importlib import util

    def load(name, path):
        spec = util.spec_from_file_location(name, path)
        m = util.module_from_spec(spec)
        spec.loader.exec_module(m)
        return m

    synth = load("pcd_synth", os.path.join(ROOT, "colmap-pcd_b200", "synth.py"))
    raw, info = synth.make_cloud(6000, seed=SEED)
    # exercise the NaN drop, zero normals, exact ties and the ground / non-ground split
    raw[17, 4] = np.nan
    raw[99, 1] = np.nan
    raw[123, 4:7] = 0.0
    raw[301, 0:3] = raw[300, 0:3]                        # duplicate position: tie -> lowest index
    world = direction_trans(raw)
    q, opt_num = synth.make_queries(world, 400, seed=SEED + 1)
    q[0] = world[300, 0:3].astype(np.float64)            # exact hit on the duplicated point
    q[1] = (world[10, 0:3].astype(np.float64) + world[11, 0:3].astype(np.float64)) / 2
    q[2] = [1e4, -1e4, 1e4]                              # far outside the cloud
    idx, d2 = nn_brute_f32(world, q)
    rng = np.where(1.5 - 0.1 * opt_num <= 0.2, 0.2, 1.5 - 0.1 * opt_num)     # incremental_mapper.cc:1160-1163
    out = dict(raw=raw, world=world, q=q, opt_num=opt_num.astype(np.int32), range=rng, nn_idx=idx, nn_d2=d2)
    for v in (0, 1, 2):
        rec = records(world, idx, q, rng, v)
        for k, a in rec.items():
            out[f"v{v}_{k}"] = a
    acc = out["v1_accepted"].astype(bool)
    wts = np.where(out["v1_type"] == 2, 10000.0, 1000.0)
    r, J = resjac(q[acc], out["v1_abcd"][acc], wts[acc])
    out.update(res_w=wts[acc], res_r=r, res_J=J)
    np.savez_compressed(os.path.join(HERE, "assoc_small.npz"), **out)
    print("assoc_small.npz:", len(world), "points,", len(q), "queries,", int(acc.sum()), "accepted (global variant)")

    # ---- z-buffer: independent numpy restatement (tests/golden/proj_numpy.py) ------------------
    pn = load("proj_numpy", os.path.join(HERE, "proj_numpy.py"))
    raw2, info2 = synth.make_cloud(20_000, seed=SEED + 7)
    world2 = direction_trans(raw2)
    # hand-placed points for the identity-pose camera (camera frame = world frame): exactly at min_proj_dist (near
    # branch, :402-405), one float above it (linear branch), nearer than min_lidar_proj_dist (skipped, :400), behind
    extra = np.zeros((5, 8), np.float32)
    extra[:, 0:3] = [[0.1, 0.2, 2.0], [-0.2, 0.1, np.nextafter(np.float32(2.0), np.float32(3.0))], [0.0, 0.1, 0.3], [0.05, 0.0, -1.0],
                     [0.1, 0.05, 0.5]]
    extra[:, 4:7] = [0.0, 0.0, -1.0]
    world2 = np.concatenate([world2, extra]).astype(np.float32)
    # back to the lidar frame so that the fixture is fed through the loader like any cloud: inverse of (x,y,z) <- (-y,-z,x)
    raw2 = np.zeros_like(world2)
    raw2[:, 0], raw2[:, 1], raw2[:, 2] = world2[:, 2], -world2[:, 0], -world2[:, 1]
    raw2[:, 4], raw2[:, 5], raw2[:, 6] = world2[:, 6], -world2[:, 4], -world2[:, 5]
    assert np.array_equal(direction_trans(raw2).view(np.uint32), world2.view(np.uint32))
    cams = synth.make_cameras(3, info2)
    ident = dict(qvec=np.array([1.0, 0.0, 0.0, 0.0]), tvec=np.zeros(3), cam=list(synth.OPENCV_1080["cam"]), width=1920, height=1080)
    runs = [(cams[0], 0.2), (cams[1], 0.2), (ident, 1.0), (dict(cams[2], cam=list(synth.OPENCV_1080["cam"]), width=1920, height=1080), 1.0)]
    rng_f = np.random.Generator(np.random.PCG64(SEED + 9))
    z = {"raw": raw2, "n_img": np.asarray([len(runs)], np.int32)}
    first_cam = {}
    for k, (cam, scale) in enumerate(runs):
        opt = dict(depth_image_scale=scale, max_proj_scale=10, min_proj_scale=2, min_proj_dist=2.0, submap_length=1.0, submap_width=1.0,
                   submap_height=1.0, choose_meter=40.0, min_lidar_proj_dist=0.5)
        if scale not in first_cam:
            first_cam[scale] = cam["cam"]                 # the function-local statics are set by the first camera of a run
        coef = pn.splat_coef(first_cam[scale], opt)
        w, h = cam["width"], cam["height"]
        dense = pn.project(world2, cam["qvec"], cam["tvec"], cam["cam"], w, h, opt, coef)
        assert np.array_equal(dense["first"], dense["lowest"]), "tie between cells: first-visited != lowest index"
        feat = np.stack([rng_f.uniform(-5, w + 5, 400), rng_f.uniform(-5, h + 5, 400)], 1)
        fpix = pn.feature_pixels(feat, scale, w, h)
        mask = np.zeros(dense["range"].size, dtype=bool)
        mask[fpix[fpix >= 0]] = True
        fw = pn.project(world2, cam["qvec"], cam["tvec"], cam["cam"], w, h, opt, coef, feat_mask=mask)
        assert np.array_equal(fw["first"], fw["lowest"])
        win = np.where(fpix >= 0, fw["lowest"].reshape(-1)[np.maximum(fpix, 0)], -1)
        packed = dense["packed_lowest"].reshape(-1)
        pix = np.nonzero(packed != np.iinfo(np.uint64).max)[0].astype(np.int32)
        z[f"img{k}_qvec"] = np.asarray(cam["qvec"], np.float64)
        z[f"img{k}_tvec"] = np.asarray(cam["tvec"], np.float64)
        z[f"img{k}_cam"] = np.asarray(cam["cam"], np.float64)
        z[f"img{k}_wh"] = np.asarray([w, h], np.int32)
        z[f"img{k}_scale"] = np.asarray([scale], np.float64)
        z[f"img{k}_coef"] = coef
        z[f"img{k}_shape"] = np.asarray(dense["range"].shape, np.int32)
        z[f"img{k}_nsel"] = np.asarray([dense["n_sel"]], np.int64)
        z[f"img{k}_covered"] = np.asarray([len(pix)], np.int64)
        z[f"img{k}_sha256"] = np.frombuffer(hashlib.sha256(packed.tobytes()).digest(), np.uint8)   # whole packed buffer
        samp = pix[:: max(1, len(pix) // 4000)]
        z[f"img{k}_pix"] = samp
        z[f"img{k}_val"] = packed[samp]
        z[f"img{k}_feat"] = feat
        z[f"img{k}_feat_pix"] = fpix
        z[f"img{k}_feat_win"] = win.astype(np.int64)
        print(f"zbuf img{k}: scale {scale}, {dense['n_sel']} cells, {len(pix)} covered pixels, {int((win >= 0).sum())} feature winners")
    # the hand-placed points must have done what they were placed for (identity-pose image = run 2)
    n0 = len(world2) - 5
    d2 = pn.project(world2[n0:], ident["qvec"], ident["tvec"], ident["cam"], 1920, 1080,
                    dict(depth_image_scale=1.0, max_proj_scale=10, min_proj_scale=2, min_proj_dist=2.0, submap_length=1.0, submap_width=1.0,
                         submap_height=1.0, choose_meter=40.0, min_lidar_proj_dist=0.5), z["img2_coef"])
    won = set(np.unique(d2["lowest"][d2["lowest"] >= 0]).tolist())
    assert won == {0, 1, 4}, won                          # the too-near and the behind-the-camera points project nothing
    np.savez_compressed(os.path.join(HERE, "zbuf_small.npz"), **z)


if __name__ == "__main__":
    main()
