"""CPU-side tests: the oracle against itself / closed forms / committed fixtures, and the C-ABI
library's exported surface.  No GPU needed (no compute entry point of the product is called)."""
import ctypes
import os
import re
import subprocess
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def small(orc, synth):
    raw, info = synth.make_cloud(120_000, seed=synth.SEED0)
    world = orc.direction_trans(raw)
    q, opt_num = synth.make_queries(world, 8_000)
    return dict(raw=raw, world=world, q=q, opt_num=opt_num, info=info, tree=orc.KdTree(world))


# ------------------------------------------------------------------ library surface ---------
def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "pcd_b200.h")).read()
    return sorted(set(re.findall(r"PCD_API[^;(]*?\b(pcd_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(pcd):
    names = _declared_symbols()
    assert len(names) >= 30
    lib = pcd.lib()
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/pcd_b200.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", pcd.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (pcd_[a-z0-9_]+)", out))
    assert set(names) <= exported
    # nothing of the oracle is linked into the product
    assert "orc_" not in out


def test_product_does_not_reference_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "colmap-pcd_b200")):
        for f in files:
            if f.endswith((".cu", ".cuh", ".cpp", ".h", ".hpp", ".py")) and f != "synth.py":
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liborc" not in txt and "pcd_oracle" not in txt and "import orc" not in txt, f


def test_no_gpu_means_loud_failure(pcd):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pcd.PcdError):
        pcd.Context(0)


def test_struct_layouts_match_header(pcd):
    assert ctypes.sizeof(pcd.ProjOptions) == 48
    assert ctypes.sizeof(pcd.Image) == 4 * 8 + 3 * 8 + 8 * 8 + 8
    assert ctypes.sizeof(pcd.AssocOut) == 8 * 8
    assert ctypes.sizeof(pcd.RangeSweep) == 24


def test_ply_io_host_only(pcd, small, tmp_path):
    raw = small["raw"][:3000]
    for binary in (True, False):
        path = str(tmp_path / f"a{int(binary)}.ply")
        pcd.ply_write(path, raw, binary=binary)
        back = pcd.ply_read(path)
        assert back.shape == raw.shape
        assert np.array_equal(np.isnan(back), np.isnan(raw))
        assert np.array_equal(np.nan_to_num(back[:, [0, 1, 2, 4, 5, 6]]), np.nan_to_num(raw[:, [0, 1, 2, 4, 5, 6]]))
    with pytest.raises(pcd.PcdError):
        pcd.ply_read(str(tmp_path / "nope.ply"))
    # COLMAP-style property names (nx ny nz) and extra properties are accepted (src/util/ply.cc:100-184)
    p = tmp_path / "named.ply"
    p.write_text("ply\nformat ascii 1.0\nelement vertex 2\nproperty float x\nproperty float y\nproperty float z\n"
                 "property float nx\nproperty float ny\nproperty float nz\nproperty uchar red\nend_header\n"
                 "1 2 3 0 0 1 7\n4 5 6 0 1 0 9\n")
    a = pcd.ply_read(str(p))
    assert a.shape == (2, 8) and np.array_equal(a[1, [0, 1, 2, 4, 5, 6]], [4, 5, 6, 0, 1, 0])


def test_ply_reader_threaded_paths(pcd, tmp_path):
    """The loader (a1, src/lidar/ply.cc:14) parses the vertex block on several host threads: byte ranges cut at line
    ends for ASCII, index ranges for binary.  Every range boundary, line shape and token form must give the value a
    single-threaded float(strtod(token)) gives; trailing elements are ignored; damage is an error, never a guess."""
    rng = np.random.default_rng(3)
    n = 300_000                                              # several MB of text: more than one byte range per thread pool
    a = (rng.normal(0, 30, (n, 6)) * 10.0 ** rng.integers(-6, 6, (n, 1))).astype(np.float32)
    a[::1000, 3] = np.nan; a[1::1000, 4] = np.inf; a[2::1000, 5] = -np.inf
    fmts = ["%.9g", "%+.8e", "%.10f", "%r"]                  # lines of very different lengths, '+' signs, exponents
    lines = []
    for i in range(n):
        f = fmts[i & 3]
        lines.append(("\t" if i % 7 == 0 else " ").join((f % float(v)) if f != "%r" else repr(float(v)) for v in a[i]) + ("\r\n" if i % 5 == 0 else "\n"))
    head = ("ply\nformat ascii 1.0\ncomment made by test\nelement vertex %d\nproperty float x\nproperty float y\nproperty float z\n"
            "property float normal_x\nproperty float normal_y\nproperty float normal_z\nelement face 1\n"
            "property list uchar int vertex_indices\nend_header\n" % n)
    path = tmp_path / "big_ascii.ply"
    path.write_text(head + "".join(lines) + "3 0 1 2", newline="")     # a face after the vertices, no trailing newline
    got = pcd.ply_read(str(path))
    want = np.zeros((n, 8), np.float32)                      # what float(strtod(token)) gives, token by token
    want[:, [0, 1, 2, 4, 5, 6]] = np.array([[float(t) for t in ln.split()] for ln in lines], dtype=np.float64).astype(np.float32)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.array_equal(np.nan_to_num(got, nan=12345.0), np.nan_to_num(want, nan=12345.0))
    assert np.array_equal(np.signbit(got), np.signbit(want))
    # last vertex line without a newline at the end of the file
    (tmp_path / "nonl.ply").write_text(head.replace("element face 1\nproperty list uchar int vertex_indices\n", "") + "".join(lines).rstrip("\r\n"), newline="")
    assert np.array_equal(np.nan_to_num(pcd.ply_read(str(tmp_path / "nonl.ply")), nan=12345.0), np.nan_to_num(want, nan=12345.0))
    # one vertex line short, and one damaged token deep inside the file
    (tmp_path / "short.ply").write_text(head.replace("element face 1\nproperty list uchar int vertex_indices\n", "") + "".join(lines[:-1]), newline="")
    with pytest.raises(pcd.PcdError, match="truncated"):
        pcd.ply_read(str(tmp_path / "short.ply"))
    broken = list(lines); broken[n - 12345] = "1 2 oops 4 5 6\n"
    (tmp_path / "broken.ply").write_text(head + "".join(broken), newline="")
    with pytest.raises(pcd.PcdError, match="bad vertex line"):
        pcd.ply_read(str(tmp_path / "broken.ply"))
    # binary little-endian with mixed property types and a trailing element; then the same file cut short
    m = 600_000
    rec = np.zeros(m, dtype=[("x", "<f8"), ("red", "u1"), ("y", "<f4"), ("z", "<f4"), ("nx", "<f4"), ("k", "<i2"), ("ny", "<f4"), ("nz", "<f8")])
    b = rng.normal(0, 50, (m, 6))
    for j, name in enumerate(("x", "y", "z", "nx", "ny", "nz")):
        rec[name] = b[:, j]
    rec["red"] = 200; rec["k"] = -3
    bhead = ("ply\nformat binary_little_endian 1.0\nelement vertex %d\nproperty double x\nproperty uchar red\nproperty float y\n"
             "property float z\nproperty float nx\nproperty short k\nproperty float ny\nproperty double nz\nelement face 0\n"
             "property list uchar int vertex_indices\nend_header\n" % m).encode()
    (tmp_path / "mixed.ply").write_bytes(bhead + rec.tobytes() + b"tail")
    got = pcd.ply_read(str(tmp_path / "mixed.ply"))
    wantb = np.zeros((m, 8), np.float32)
    for j, name in zip((0, 1, 2, 4, 5, 6), ("x", "y", "z", "nx", "ny", "nz")):
        wantb[:, j] = rec[name].astype(np.float32)
    assert np.array_equal(got, wantb)
    (tmp_path / "cut.ply").write_bytes(bhead + rec.tobytes()[:-5])
    with pytest.raises(pcd.PcdError, match="truncated"):
        pcd.ply_read(str(tmp_path / "cut.ply"))


# ------------------------------------------------------------------ oracle: loader / NN -----
def test_direction_trans(orc):
    a = np.zeros((4, 8), np.float32)
    a[:, :3] = [[1, 2, 3], [4, 5, 6], [7, 8, 9], [1, 1, 1]]
    a[:, 4:7] = [[0, 0, 1], [0, 1, 0], [np.nan, 0, 0], [1, 0, 0]]
    w = orc.direction_trans(a)
    assert len(w) == 3                                   # NaN row dropped, order kept
    assert np.array_equal(w[0, :3], [-2, -3, 1]) and np.array_equal(w[0, 4:7], [0, -1, 0])
    assert np.array_equal(w[2, :3], [-1, -1, 1])


def test_kdtree_equals_brute_force(orc, small):
    i, d = small["tree"].nn(small["q"])
    bi, bd = orc.nn_brute(small["world"], small["q"][:1200])
    assert np.array_equal(i[:1200], bi) and np.array_equal(d[:1200].view(np.uint32), bd.view(np.uint32))
    # threads do not change results
    i2, d2 = small["tree"].nn(small["q"], threads=1)
    assert np.array_equal(i, i2) and np.array_equal(d, d2)


def test_nn_against_scipy(orc, small):
    from scipy.spatial import cKDTree
    w = small["world"][:, :3].astype(np.float64)
    qf = small["q"].astype(np.float32).astype(np.float64)
    _, j = cKDTree(w).query(qf)
    i, d = small["tree"].nn(small["q"])
    diff = np.flatnonzero(i != j)
    # any disagreement must be a float32 near-tie
    for k in diff:
        da = np.sum((w[i[k]] - qf[k]) ** 2); db = np.sum((w[j[k]] - qf[k]) ** 2)
        assert abs(da - db) <= 1e-6 * max(da, db)
    assert len(diff) < 0.001 * len(i)


def test_ties_go_to_lowest_index(orc):
    c = np.zeros((6, 8), np.float32)
    c[:, :3] = [[1, 0, 0], [-1, 0, 0], [0, 1, 0], [1, 0, 0], [0, -1, 0], [0, 0, 1]]
    c[:, 5] = 1
    i, d = orc.nn_brute(c, np.array([[0.0, 0, 0], [1.0, 0, 0]]))
    assert list(i) == [0, 0] and d[0] == 1.0 and d[1] == 0.0
    i2, _ = orc.KdTree(c).nn(np.array([[0.0, 0, 0], [1.0, 0, 0]]))
    assert list(i2) == [0, 0]


def test_search_range_sweep(orc):
    r = orc.search_range(np.arange(0, 20))
    want = np.maximum(0.2, 1.5 - 0.1 * np.arange(0, 20))
    assert np.allclose(r, want, rtol=0, atol=1e-12) and len(np.unique(np.round(r, 9))) == 14
    assert r[13] == 0.2 or abs(r[13] - 0.2) < 1e-12


# ------------------------------------------------------------------ oracle: record ----------
def test_record_closed_forms(orc, small):
    rng = orc.search_range(small["opt_num"])
    rec = small["tree"].associate(small["q"], rng, variant=1)
    acc = rec["accepted"].astype(bool)
    assert 0.3 < acc.mean() < 0.999
    l = rec["xyz"][acc]; p = small["q"][acc]; pl = rec["abcd"][acc]
    assert np.allclose(np.linalg.norm(pl[:, :3], axis=1), 1.0, atol=1e-12)
    assert np.allclose((l * pl[:, :3]).sum(1) + pl[:, 3], 0.0, atol=1e-9)       # plane passes through l
    assert np.allclose(rec["dist"][acc], np.linalg.norm(l - p, axis=1), rtol=1e-12)
    assert (rec["dist"][acc] <= rng[acc]).all()
    n_raw = small["world"][rec["idx"][acc], 4:7].astype(np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        ground = (np.abs(n_raw[:, 1] / n_raw[:, 0]) > 10) & (np.abs(n_raw[:, 1] / n_raw[:, 2]) > 10)
    assert np.array_equal(rec["type"][acc] == 2, ground)
    assert np.array_equal(rec["rgb"][acc][ground], np.tile([255, 255, 0], (ground.sum(), 1)))
    assert np.array_equal(rec["rgb"][acc][~ground], np.tile([0, 0, 255], ((~ground).sum(), 1)))
    # rejected rows: zero normals (|n| < 1e-6) or dist > range
    rej = ~acc
    idx = rec["idx"][rej]
    nn_i, _ = small["tree"].nn(small["q"][rej])
    nrm = np.linalg.norm(small["world"][nn_i, 4:7].astype(np.float64), axis=1)
    d = np.linalg.norm(small["world"][nn_i, :3].astype(np.float64) - small["q"][rej], axis=1)
    assert ((nrm < 1e-6) | (d > rng[rej])).all()
    assert idx.shape == nn_i.shape


def test_whole_map_variant_thresholds(orc, small):
    rec = small["tree"].associate(small["q"], None, variant=2)
    acc = rec["accepted"].astype(bool)
    p = small["q"]; pl = rec["abcd"]; l = rec["xyz"]
    d2plane = np.abs((p * pl[:, :3]).sum(1) + pl[:, 3])
    dist = np.linalg.norm(l - p, axis=1)
    valid = np.linalg.norm(pl[:, :3], axis=1) > 0.5
    assert np.array_equal(acc[valid], ((d2plane <= 1) & (dist <= 2))[valid])


def test_records_from_idx_equals_associate(orc, small):
    rng = orc.search_range(small["opt_num"])
    a = small["tree"].associate(small["q"], rng, variant=0)
    b = orc.records_from_idx(small["world"], a["idx"], small["q"], rng, variant=0)
    for f in b:
        assert np.array_equal(a[f], b[f], equal_nan=True) if a[f].dtype.kind == "f" else np.array_equal(a[f], b[f])


# ------------------------------------------------------------------ oracle: cost function ---
def test_resjac_closed_form_and_fd(orc):
    rng = np.random.default_rng(0)
    n = 2000
    nn = rng.normal(0, 1, (n, 3)); nn /= np.linalg.norm(nn, axis=1, keepdims=True)
    abcd = np.concatenate([nn, rng.normal(0, 2, (n, 1))], 1)
    p = rng.normal(0, 5, (n, 3)); w = rng.choice([10.0, 1000.0, 10000.0], n)
    r, J = orc.resjac(p, abcd, w)
    s = (p * nn).sum(1) + abcd[:, 3]
    assert np.allclose(r, w * np.abs(s), rtol=1e-12)
    assert np.allclose(J, (w * np.sign(s))[:, None] * nn, rtol=1e-12)
    eps = 1e-6
    for k in range(3):
        dp = p.copy(); dp[:, k] += eps
        fd = (orc.resjac(dp, abcd, w)[0] - r) / eps
        ok = np.abs(s) > 1e-3
        assert np.allclose(fd[ok], J[ok, k], rtol=1e-4, atol=1e-4 * w[ok])
    r1, J1 = orc.resjac(p, abcd, w, threads=0)
    assert np.array_equal(r, r1) and np.array_equal(J, J1)


def test_resjac_nan_at_zero(orc):
    r, J = orc.resjac([[3.0, 1.6, 4.0]], [[0.0, -1.0, 0.0, 1.6]], [1000.0])
    assert r[0] == 0.0 and np.isnan(J).all()             # autodiff gives 0 * inf


def test_resjac_nan_plane_adds_no_block(orc):
    # AddLidarToProblem returns before Create when a plane coefficient is NaN (src/optim/bundle_adjustment.cc:1003-1007)
    r, J = orc.resjac([[1.0, 2.0, 3.0], [1.0, 2.0, 3.0]], [[np.nan, 0.0, 1.0, -1.0], [0.0, 0.0, 1.0, np.nan]], [10.0, 10.0])
    assert not r.any() and not J.any()


def test_known_answer_vectors(orc):
    # hand-computed known answers (the reference has none for this path)
    r, J = orc.resjac([[1.0, 2.0, 3.0]], [[0.0, 0.0, 1.0, -1.0]], [10.0])
    assert r[0] == 20.0 and np.array_equal(J[0], [0.0, 0.0, 10.0])
    r, J = orc.resjac([[1.0, 2.0, -3.0]], [[0.0, 0.0, 1.0, -1.0]], [10.0])
    assert r[0] == 40.0 and np.array_equal(J[0], [-0.0, -0.0, -10.0])
    ok, xyz, abcd, d, a = orc.match_proj([0.0, 0.0, 0.0], [[0.0, 0.0, 2.0, 0.0, 0.0, 2.0], [0.0, 3.0, 0.0, 1.0, 0.0, 0.0]])
    assert ok and np.array_equal(xyz, [0.0, 3.0, 0.0]) and np.array_equal(abcd, [1.0, 0.0, 0.0, -0.0]) and d == 0.0 and a == 0.0


# ------------------------------------------------------------------ oracle: z-buffer ---------
def test_zbuffer_sequential_properties(orc, synth, small):
    opt = orc.ProjOptions.default()
    sub = orc.Submap(small["world"], opt)
    cam = synth.make_cameras(3, small["info"])[1]
    coef = orc.splat_coef(cam["cam"], opt)
    z, counters, n_sel = sub.project(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"], coef)
    assert z.shape == (604, 806) and n_sel > 10 and counters[0] > 1000
    hit = z != np.iinfo(np.uint64).max
    assert 0.02 < hit.mean() <= 1.0
    idx = (z[hit] & np.uint64(0xffffffff)).astype(np.int64)
    rng_ = (z[hit] >> np.uint64(32)).astype(np.uint32).view(np.float32)
    assert idx.max() < len(small["world"]) and (rng_ >= 0.5).all() and (rng_ < 60).all()
    # the stored range is |R p + t| of the stored index
    from_q = _rot(cam["qvec"]).astype(np.float32)
    pc = small["world"][idx, :3] @ from_q.T + np.asarray(cam["tvec"], np.float32)
    assert np.allclose(np.linalg.norm(pc, axis=1), rng_, rtol=1e-5)
    # restricting to a feature mask gives the same winners at those pixels
    mask = np.zeros(z.size, np.uint8); sel = np.random.default_rng(1).integers(0, z.size, 4000); mask[sel] = 1
    zm, _, _ = sub.project(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"], coef, mask)
    assert np.array_equal(zm.ravel()[sel], z.ravel()[sel])
    assert (zm.ravel()[mask == 0] == np.iinfo(np.uint64).max).all()


def _rot(q):
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def test_splat_coef_quirk(orc):
    opt = orc.ProjOptions.default()
    c = orc.splat_coef([3039.0, 3039.0, 2016.0, 1512.0, 0, 0, 0, 0], opt)
    # a = (max - min) / (min_proj_dist - choose_meter) = 8 / -38 ; b_x = min - a*40 ; b_y uses raw min_proj_scale (F7)
    assert np.isclose(c[0], 8 / -38.0) and np.isclose(c[1], 2 + 8 / 38.0 * 40) and np.isclose(c[3], 2 + 8 / 38.0 * 40)
    c2 = orc.splat_coef([1447.0, 1500.0, 960.0, 540.0, 0, 0, 0, 0], orc.ProjOptions.default(depth_image_scale=1.0))
    min_y = 2 * (1500.0 / 3039.0) * 5.0
    assert not np.isclose(c2[3], min_y - c2[2] * 40.0) and np.isclose(c2[3], 2 - c2[2] * 40.0)


def test_feature_pixels_truncate(orc):
    px = orc.feature_pixels([[9.99, 4.99], [-0.1, 3.0], [-6.0, 3.0], [4031.9, 3023.9], [4035.0, 10.0]], 0.2, 4032, 3024)
    # (xy * s).cast<int>() truncates toward zero: -0.02 -> 0 is in bounds, -1.2 -> -1 is not
    assert list(px) == [0 * 806 + 1, 0 * 806 + 0, -1, 604 * 806 - 1 - 0, -1] or px[0] == 1


# ------------------------------------------------------------------ N4: pose.ply format ---------
def test_pose_ply_roundtrip_matches_reference_conversion(pcd, tmp_path):
    """LoadPose (src/controllers/incremental_mapper.cc:922-997) restated with scipy rotations."""
    from scipy.spatial.transform import Rotation as Rot
    rng = np.random.default_rng(11)
    poses = np.concatenate([rng.uniform(-30, 30, (40, 3)), rng.uniform(-np.pi, np.pi, (40, 3))], 1)
    poses[7] = np.nan                                           # skipped, but its image id is consumed
    path = str(tmp_path / "pose.ply")
    pcd.pose_ply_write(path, poses)
    ids, tq = pcd.pose_ply_read(path)
    keep = [i for i in range(40) if i != 7]
    assert np.array_equal(ids, np.asarray(keep) + 1)
    for row, i in zip(tq, keep):
        x, y, z, roll, pitch, yaw = poses[i]
        R_wc = (Rot.from_rotvec([0, -yaw, 0]) * Rot.from_rotvec([-pitch, 0, 0]) * Rot.from_rotvec([0, 0, roll])).as_matrix()
        R_cw = R_wc.T
        t_cw = -R_cw @ np.array([-y, -z, x])
        assert np.allclose(row[0:3], t_cw, rtol=1e-12, atol=1e-12)
        q = row[3:7]                                            # w x y z
        R_back = Rot.from_quat([q[1], q[2], q[3], q[0]]).as_matrix()
        assert np.allclose(R_back, R_cw, atol=1e-12) and abs(np.linalg.norm(q) - 1) < 1e-12
    with pytest.raises(pcd.PcdError):
        pcd.pose_ply_read(str(tmp_path / "missing.ply"))


def test_project2image_selection_logic(orc):
    """a16: which images of a point's track BundleAdjustmentConfig::Project2Image projects
    (src/optim/bundle_adjustment.cc:241-281) -- library (host logic, no GPU) vs the oracle's walk of the reference loop,
    plus hand-worked cases."""
    import __graft_entry__ as entry
    pcd = entry.package()
    # hand-worked: image 5 is the new image, threshold 30
    #   el 0: image 5 itself            -> projected (not cached)
    #   el 1: image 7, pair has 30      -> skipped (corrs <= threshold)
    #   el 2: image 8, pair has 31      -> projected
    #   el 3: image 9, no such pair     -> projected
    #   el 4: image 8 again             -> already cached by el 2
    #   el 5: image 6, pair has 500, cached before the call -> not projected
    track = [5, 7, 8, 9, 8, 6]; corrs = [-1, 30, 31, -1, 31, 500]; cached = [0, 0, 0, 0, 0, 1]
    sel, n = pcd.project2image_select(track, 5, corrs, 30, cached)
    assert sel.tolist() == [1, 0, 1, 1, 0, 0] and n == 3
    so, no = orc.project2image(track, 5, corrs, 30, cached_ids=[6])
    assert so.tolist() == sel.tolist() and no == n
    # a negative threshold converts to a huge size_t (the reference compares size_t <= int): every existing pair is skipped
    sel, n = pcd.project2image_select([1, 2, 3], 1, [10, 10, -1], -1)
    assert sel.tolist() == [1, 0, 1] and n == 2
    rng = np.random.default_rng(4)
    for _ in range(300):
        m = int(rng.integers(0, 40))
        track = rng.integers(0, 12, m).astype(np.uint32)
        image_id = int(rng.integers(0, 12))
        pair = {int(i): (int(rng.integers(0, 60)) if rng.random() < 0.7 else -1) for i in range(12)}
        corrs = np.array([pair[int(t)] for t in track], np.int64)
        cached_ids = [int(i) for i in range(12) if rng.random() < 0.3]
        flags = np.array([int(t) in cached_ids for t in track], np.uint8)
        thr = int(rng.integers(-1, 50))
        sel, n = pcd.project2image_select(track, image_id, corrs, thr, flags)
        so, no = orc.project2image(track, image_id, corrs, thr, cached_ids)
        assert np.array_equal(sel, so) and n == no
    assert pcd.project2image_select([], 0, [], 3)[1] == 0


def test_sidecar_roundtrip_and_rejection(tmp_path, synth):
    """N3: the sidecar of the parsed cloud -- exact round trip, and a stale / truncated / foreign file is refused."""
    import __graft_entry__ as entry
    pcd = entry.package()
    raw, _ = synth.make_cloud(5000, seed=synth.SEED0 + 21)
    src = str(tmp_path / "map.ply"); side = src + ".pcdb200"
    pcd.ply_write(src, raw, binary=True)
    key = pcd.file_key(src)
    assert key == pcd.file_key(src)
    assert pcd.sidecar_read(side, key) is None                       # missing
    pcd.sidecar_write(side, key, raw)
    got = pcd.sidecar_read(side, key)
    assert got is not None and np.array_equal(got.view(np.uint32), raw.view(np.uint32))     # NaN normals and all
    assert pcd.sidecar_read(side, key ^ 1) is None                    # another source file
    assert pcd.sidecar_read(side, key, direction_trans=False) is None
    data = open(side, "rb").read()
    open(side, "wb").write(data[:-7])                                 # truncated
    assert pcd.sidecar_read(side, key) is None
    open(side, "wb").write(b"NOTASIDE" + data[8:])                    # foreign magic
    assert pcd.sidecar_read(side, key) is None
    raw2 = raw.copy(); raw2[0, 0] += 1.0
    pcd.ply_write(src, raw2, binary=True)                             # the source changed: the key moves
    assert pcd.file_key(src) != key
    pcd.sidecar_write(side, key, np.zeros((0, 8), np.float32))
    assert len(pcd.sidecar_read(side, key)) == 0
