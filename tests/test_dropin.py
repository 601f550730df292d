"""The C++ drop-in shim (colmap-pcd_b200/host/pcd_b200_dropin.h): the reference's own class and method
names over the C ABI, driven by host/dropin_check.cpp the way the reference's call sites drive them."""
import os
import subprocess
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "colmap-pcd_b200", "host", "dropin_check")
REL_TOL = 1e-5


def _close(a, b, tol=REL_TOL):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    both_nan = np.isnan(a) & np.isnan(b)             # e.g. the ray/plane point of a zero normal: 0 / 0 on both sides
    with np.errstate(invalid="ignore"):
        return bool((both_nan | (np.abs(a - b) <= tol * np.maximum(np.abs(a), np.abs(b)) + 1e-300)).all())


def test_shim_builds_and_fails_loudly_without_gpu(tmp_path):
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "colmap-pcd_b200"), "host/dropin_check"])
    assert os.path.exists(EXE)
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    r = subprocess.run([EXE, str(tmp_path / "a.ply"), str(tmp_path / "q"), str(tmp_path / "o")], capture_output=True, text=True)
    assert r.returncode != 0 and "no CPU fallback" in r.stderr      # Initialize() returned false, no host formula ran


@pytest.mark.gpu
def test_dropin_shim_matches_oracle(tmp_path, orc, synth, pcd):
    raw, info = synth.make_cloud(80_000, seed=synth.SEED0 + 11)
    world = orc.direction_trans(raw)
    ply = str(tmp_path / "cloud.ply")
    pcd.ply_write(ply, raw, binary=True)
    q, on = synth.make_queries(world, 3000, seed=synth.SEED0 + 12)
    np.concatenate([q, on[:, None].astype(np.float64)], 1).tofile(str(tmp_path / "q.f64"))
    cam = synth.make_cameras(2, info)[1]
    feats = synth.make_features(1500)
    np.concatenate([np.asarray(cam["qvec"], np.float64), np.asarray(cam["tvec"], np.float64), np.asarray(cam["cam"], np.float64),
                    [cam["width"], cam["height"], len(feats)], feats.ravel()]).tofile(str(tmp_path / "im.f64"))
    r = subprocess.run([EXE, ply, str(tmp_path / "q.f64"), str(tmp_path / "o.f64"), str(tmp_path / "im.f64")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    o = np.fromfile(str(tmp_path / "o.f64"))
    # the same program over a two-rank group (PCD_B200_DEVICES; one GPU: two shards on device 0): identical dump
    import torch
    devs = "0,1" if torch.cuda.device_count() >= 2 else "0,0"
    r2 = subprocess.run([EXE, ply, str(tmp_path / "q.f64"), str(tmp_path / "o2.f64"), str(tmp_path / "im.f64")],
                        capture_output=True, text=True, env=dict(os.environ, PCD_B200_DEVICES=devs))
    assert r2.returncode == 0, r2.stderr
    o2 = np.fromfile(str(tmp_path / "o2.f64"))
    assert np.array_equal(o, o2, equal_nan=True)
    pos = 0
    assert o[pos] == len(world); pos += 1
    tree = orc.KdTree(world)
    oi, _ = tree.nn(q)
    # (1) serial SearchNearestNeiborByKdtree and Kdtree::GetClosestPoint
    for i in range(64):
        ok, l, _ = tree.search_nn(q[i])
        assert bool(o[pos]) == ok
        if ok:
            assert np.array_equal(o[pos + 1:pos + 7], l)
        pos += 7
    for i in range(64):
        assert o[pos] == 1.0 and np.array_equal(o[pos + 1:pos + 4].astype(np.float32), world[oi[i], 0:3]); pos += 4
    # (2) batched association (global variant) + cost function through the evaluation callback
    rng = orc.search_range(on)
    want = tree.associate(q, rng, variant=1)
    acc = np.flatnonzero(want["accepted"])
    assert o[pos] == len(acc); pos += 1
    wts = np.where(want["type"] == 2, 10000.0, 1000.0)
    ro, Jo = orc.resjac(q[acc], want["abcd"][acc], wts[acc])
    rec = o[pos:pos + 14 * len(acc)].reshape(len(acc), 14); pos += 14 * len(acc)
    assert np.array_equal(rec[:, 0], acc + 1000)
    assert np.array_equal(rec[:, 1].astype(np.uint8), want["type"][acc])
    assert _close(rec[:, 2:5], want["xyz"][acc]) and _close(rec[:, 5:9], want["abcd"][acc]) and _close(rec[:, 9], want["dist"][acc])
    assert _close(rec[:, 10], ro) and _close(rec[:, 11:14], Jo)
    moved = q[acc[0]] + [0.5, 0, 0]
    rm, _ = orc.resjac([moved], want["abcd"][acc[:1]], wts[acc[:1]])
    assert _close(o[pos], rm[0]); pos += 1
    # after the probe the batch is still bound to the blocks: the first block moved by 0.25, prepared again
    rm2, _ = orc.resjac([q[acc[0]] + [0.25, 0, 0]], want["abcd"][acc[:1]], wts[acc[:1]])
    assert _close(o[pos], rm2[0]); pos += 1
    # (3) PcdProj::SetNewImage (map overload) == oracle z-buffer winners at the feature pixels
    opt = orc.ProjOptions.default()
    sub = orc.Submap(world, opt)
    coef = orc.splat_coef(cam["cam"], opt)
    px = orc.feature_pixels(feats, 0.2, cam["width"], cam["height"])
    ih, iw = int(cam["height"] * 0.2), int(cam["width"] * 0.2)
    mask = np.zeros(ih * iw, np.uint8); mask[px[px >= 0]] = 1
    zo, _, _ = sub.project(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"], coef, mask)
    flat = zo.reshape(-1)
    exp = {}
    for k, p in enumerate(px):
        if p >= 0 and flat[p] != np.iinfo(np.uint64).max:
            exp[k] = int(flat[p] & 0xffffffff)
    m = int(o[pos]); pos += 1
    assert m == len(exp) and m > 50
    got = o[pos:pos + 7 * m].reshape(m, 7)
    for row in got:
        gi = exp[int(row[0])]
        assert np.array_equal(row[1:4].astype(np.float32), world[gi, 0:3]) and np.array_equal(row[4:7].astype(np.float32), world[gi, 4:7])
    pos += 7 * m
    # (4) the initial-pair overload: matched flag and camera-frame point per keypoint (pcd_projection.cc:102-220)
    nf = int(o[pos]); pos += 1
    assert nf == len(feats)
    init = o[pos:pos + 4 * nf].reshape(nf, 4); pos += 4 * nf
    assert pos == len(o)
    for k in range(nf):
        assert bool(init[k, 0]) == (k in exp)
        if k in exp:
            assert _close(init[k, 1:4], orc.ray_plane(world, exp[k], cam["cam"], feats[k, 0], feats[k, 1]))
        else:
            assert not init[k, 1:4].any()
