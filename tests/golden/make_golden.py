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
  zbuf_small.npz   a REGRESSION pin of the depth-projection z-buffer: produced by the C oracle
                   itself (sequential reading of src/lidar/pcd_projection.cc:258-559), so it only
                   guards against drift; the independent check of the z-buffer is
                   tests/test_cpu.py::test_zbuffer_sequential_properties.
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

    # ---- z-buffer regression pin (produced by the C oracle) -------------------------------------
    orc = load("orc", os.path.join(ROOT, "oracle", "orc.py"))
    raw2, info2 = synth.make_cloud(20_000, seed=SEED + 7)
    world2 = orc.direction_trans(raw2)
    opt = orc.ProjOptions.default()
    sub = orc.Submap(world2, opt)
    cams = synth.make_cameras(3, info2)
    coef = orc.splat_coef(cams[0]["cam"], opt)
    z = {}
    for k, cam in enumerate(cams):
        zb, _, _ = sub.project(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"], coef)
        z[f"img{k}_qvec"] = np.asarray(cam["qvec"], np.float64)
        z[f"img{k}_tvec"] = np.asarray(cam["tvec"], np.float64)
        z[f"img{k}_cam"] = np.asarray(cam["cam"], np.float64)
        z[f"img{k}_wh"] = np.asarray([cam["width"], cam["height"]], np.int32)
        flat = zb.reshape(-1)
        pix = np.nonzero(flat != np.iinfo(np.uint64).max)[0].astype(np.int32)
        z[f"img{k}_covered"] = np.asarray([len(pix)], np.int64)
        z[f"img{k}_sha256"] = np.frombuffer(hashlib.sha256(flat.tobytes()).digest(), np.uint8)   # whole buffer
        pix = pix[:: max(1, len(pix) // 4000)]               # plus a readable sample of packed winners
        z[f"img{k}_pix"] = pix
        z[f"img{k}_val"] = flat[pix]
        z[f"img{k}_shape"] = np.asarray(zb.shape, np.int32)
    z["raw"] = raw2
    z["coef"] = np.asarray(coef, np.float64)
    np.savez_compressed(os.path.join(HERE, "zbuf_small.npz"), **z)
    print("zbuf_small.npz:", len(world2), "points, 3 images,", sum(int(z[f"img{k}_covered"][0]) for k in range(3)), "covered pixels")


if __name__ == "__main__":
    main()
