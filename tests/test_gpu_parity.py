"""Parity of the CUDA path (through the C ABI) against the CPU oracle.  All tests need a B200."""
import os
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL_TOL = 1e-5   # north_star: residuals / Jacobians / fp64 record fields within 1e-5 relative


def _close(a, b, tol=REL_TOL):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    both_nan = np.isnan(a) & np.isnan(b)
    ok = both_nan | (np.abs(a - b) <= tol * np.maximum(np.abs(a), np.abs(b)) + 1e-300)
    return bool(ok.all())


@pytest.fixture(scope="module")
def scene(orc, synth, pcd, ctx):
    raw, info = synth.make_cloud(300_000, seed=synth.SEED0)
    world = orc.direction_trans(raw)
    cloud = pcd.Cloud(ctx, raw, direction_trans=True)
    cloud.build_index()
    q, opt_num = synth.make_queries(world, 40_000)
    tree = orc.KdTree(world)
    yield dict(raw=raw, world=world, cloud=cloud, q=q, opt_num=opt_num, tree=tree, info=info)
    cloud.close()


# ---------------------------------------------------------------- loader (a1, a2) ----------
def test_direction_trans_and_nan_drop(scene):
    got = scene["cloud"].download()
    assert got.shape == scene["world"].shape
    assert np.array_equal(got.view(np.uint32), scene["world"].view(np.uint32))   # bit-exact incl. -0.0


def test_ply_roundtrip_and_load(tmp_path, scene, pcd, ctx, orc):
    raw = scene["raw"][:5000]
    for binary in (True, False):
        path = str(tmp_path / f"c{int(binary)}.ply")
        pcd.ply_write(path, raw, binary=binary)
        back = pcd.ply_read(path)
        assert np.array_equal(np.isnan(back), np.isnan(raw))
        assert np.array_equal(np.nan_to_num(back), np.nan_to_num(raw))
        cl = pcd.Cloud.load_ply(ctx, path, pcd.ProjOptions.default())
        assert np.array_equal(np.nan_to_num(cl.download()), np.nan_to_num(orc.direction_trans(raw)))
        cl.close()
    with pytest.raises(pcd.PcdError):
        pcd.Cloud.load_ply(ctx, str(tmp_path / "missing.ply"))


# ---------------------------------------------------------------- NN (a3-a5) ---------------
def test_nn_unbounded_bit_exact(scene, orc):
    idx, d2 = scene["cloud"].nn(scene["q"])
    oi, od = scene["tree"].nn(scene["q"])
    assert np.array_equal(idx, oi)
    assert np.array_equal(d2.view(np.uint32), od.view(np.uint32))
    # the kd-tree oracle itself against exhaustive search on a slice
    bi, bd = orc.nn_brute(scene["world"], scene["q"][:1500])
    assert np.array_equal(bi, oi[:1500]) and np.array_equal(bd, od[:1500])


def test_nn_radius_semantics(scene, orc):
    rng = orc.search_range(scene["opt_num"])          # 14 distinct radii 1.5 .. 0.2
    assert len(np.unique(rng)) == 14
    idx, d2 = scene["cloud"].nn(scene["q"], rng)
    oi, _ = scene["tree"].nn(scene["q"])
    l = scene["world"][oi, 0:3].astype(np.float64) - scene["q"]
    dist = np.sqrt(l[:, 0] ** 2 + (l[:, 1] ** 2 + l[:, 2] ** 2))
    want = np.where(dist > rng, -1, oi)
    assert np.array_equal(idx, want)
    assert (idx < 0).sum() > 100 and (idx >= 0).sum() > 100
    assert np.isinf(d2[idx < 0]).all()


@pytest.mark.parametrize("cell", [0.0625, 0.25, 1.0, 4.0])
def test_nn_any_cell_size(scene, pcd, ctx, cell):
    cl = pcd.Cloud(ctx, scene["world"], direction_trans=False)
    info = cl.build_index(cell)
    assert info["cell"] == cell
    q = scene["q"][:8000]
    idx, d2 = cl.nn(q)
    oi, od = scene["tree"].nn(q)
    assert np.array_equal(idx, oi) and np.array_equal(d2, od)
    cl.close()


def test_nn_edge_cases(scene, pcd, ctx, orc):
    world = scene["world"]
    cl = scene["cloud"]
    # empty batch
    idx, d2 = cl.nn(np.zeros((0, 3)))
    assert len(idx) == 0
    # NaN / inf queries -> no neighbour; far-away queries still find the exact NN when unbounded
    lo, hi = world[:, :3].min(0), world[:, :3].max(0)
    q = np.array([[np.nan, 0, 0], [np.inf, 0, 0], [hi[0] + 500.0, 0.3, 1.0], [lo[0] - 80.0, lo[1] - 90.0, lo[2] - 70.0],
                  [0.5 * (lo[0] + hi[0]), lo[1] - 30.0, 0.5 * (lo[2] + hi[2])]])
    idx, d2 = cl.nn(q)
    assert idx[0] == -1 and idx[1] == -1
    bi, bd = orc.nn_brute(world, q[2:])
    assert np.array_equal(idx[2:], bi) and np.array_equal(d2[2:], bd)
    # same far queries with a radius: nothing within 1.5 m
    idx, _ = cl.nn(q, 1.5)
    assert (idx == -1).all()
    # small batch path (no sort) and a batch in the sorted path give identical answers
    a, _ = cl.nn(scene["q"][:100]); b, _ = cl.nn(scene["q"][:5000])
    assert np.array_equal(a, b[:100])


def test_nn_radius_boundary_at_map_scale(pcd, ctx, orc):
    """The accept test is the reference's fp64 |l - p| <= R on the DOUBLE query; the search works on the float cast,
    which moves a query by up to ~7e-6 m at 100 m coordinates.  Radii placed one fp64 ulp either side of the true
    distance must still decide by that test (the float ball has to cover the cast displacement)."""
    rng = np.random.default_rng(11)
    off = np.array([412.7, -233.1, 167.9])
    pts = (rng.uniform(-4, 4, (30_000, 3)) + off).astype(np.float32)
    cloud = np.zeros((len(pts), 8), np.float32); cloud[:, :3] = pts; cloud[:, 5] = -1.0
    cl = pcd.Cloud(ctx, cloud); cl.build_index()
    q = rng.uniform(-3.5, 3.5, (6000, 3)) + off
    bi, _ = orc.nn_brute(cloud, q)
    l = pts[bi].astype(np.float64)
    v = l - q
    dist = np.sqrt(v[:, 0] * v[:, 0] + (v[:, 1] * v[:, 1] + v[:, 2] * v[:, 2]))      # Eigen's (l - p).norm()
    for radius, hit in ((dist, True), (np.nextafter(dist, np.inf), True), (np.nextafter(dist, -np.inf), False)):
        idx, _ = cl.nn(q, radius)
        assert np.array_equal(idx >= 0, np.full(len(q), hit)), (hit, int((idx >= 0).sum()))
        if hit:
            assert np.array_equal(idx, bi)
    cl.close()


def test_fused_pass_nan_plane_adds_no_block(pcd, ctx, orc):
    """An infinite normal survives the NaN / zero-normal test of ply.cc:100-101, normalises to NaN coefficients and is
    stored as a match, but AddLidarToProblem (bundle_adjustment.cc:1003-1007) adds no residual block for it."""
    rng = np.random.default_rng(3)
    pts = rng.uniform(-2, 2, (5000, 3)).astype(np.float32)
    cloud = np.zeros((len(pts), 8), np.float32); cloud[:, :3] = pts; cloud[:, 5] = -1.0
    cloud[::7, 4] = np.inf                                      # every 7th point: normal (inf, -1, 0)
    cl = pcd.Cloud(ctx, cloud); cl.build_index()
    q = pts[:3000].astype(np.float64) + rng.normal(0, 0.01, (3000, 3))
    W = (10.0, 1000.0, 10000.0)
    rec, r, J, s = cl.associate_evaluate(q, W, range_=np.full(len(q), 0.5), variant=pcd.ASSOC_GLOBAL)
    tree = orc.KdTree(cloud)
    want = tree.associate(q, np.full(len(q), 0.5), variant=1)
    acc = want["accepted"].astype(bool)
    assert np.array_equal(rec["accepted"].astype(bool), acc)
    bad = acc & np.isnan(want["abcd"]).any(axis=1)
    assert bad.sum() > 100 and np.array_equal(np.isnan(rec["abcd"]).any(axis=1)[acc], np.isnan(want["abcd"]).any(axis=1)[acc])
    assert not r[bad].any() and not J[bad].any() and np.isfinite(s)
    w = np.where(want["type"] == 2, W[2], W[1])
    ro, Jo = orc.resjac(q[acc], want["abcd"][acc], w[acc])
    assert _close(r[acc], ro) and _close(J[acc], Jo) and _close(s, float(np.sum(ro * ro)), 1e-9)
    r3, J3, s3 = ctx.residual_jacobian(q[acc], rec["abcd"][acc], w[acc])            # the standalone kernel: same rule
    assert np.array_equal(r3, r[acc]) and np.array_equal(J3, J[acc], equal_nan=True)
    cl.close()


def test_nn_ties_lowest_index(pcd, ctx, orc):
    # duplicated and mirrored points: exact float ties must resolve to the lowest index
    rng = np.random.default_rng(5)
    base = rng.uniform(-3, 3, (2000, 3)).astype(np.float32)
    pts = np.concatenate([base, base[::-1], base])            # every location three times
    cloud = np.zeros((len(pts), 8), np.float32); cloud[:, :3] = pts; cloud[:, 5] = -1.0
    cl = pcd.Cloud(ctx, cloud); cl.build_index()
    q = np.concatenate([base[:500].astype(np.float64), rng.uniform(-3, 3, (3000, 3))])
    idx, d2 = cl.nn(q)
    bi, bd = orc.nn_brute(cloud, q)
    assert np.array_equal(idx, bi) and np.array_equal(d2, bd)
    assert (idx[:500] == np.arange(500)).all()
    cl.close()


def test_nn_tiny_and_degenerate_clouds(pcd, ctx, orc):
    one = np.zeros((1, 8), np.float32); one[0, :3] = [1, 2, 3]; one[0, 4] = 1
    cl = pcd.Cloud(ctx, one); cl.build_index()
    idx, d2 = cl.nn(np.array([[0.0, 0, 0], [100.0, 5, 5]]))
    assert (idx == 0).all() and d2[0] == np.float32(14.0)
    cl.close()
    empty = pcd.Cloud(ctx, np.zeros((0, 8), np.float32)); empty.build_index()
    idx, _ = empty.nn(np.zeros((3, 3)))
    assert (idx == -1).all()
    empty.close()
    # all points on a line (two grid dimensions collapse to 1)
    line = np.zeros((5000, 8), np.float32); line[:, 0] = np.linspace(0, 50, 5000); line[:, 5] = 1
    cl = pcd.Cloud(ctx, line); cl.build_index()
    q = np.random.default_rng(1).uniform(-5, 55, (3000, 3)) * [1, 0.05, 0.05]
    idx, d2 = cl.nn(q)
    bi, bd = orc.nn_brute(line, q)
    assert np.array_equal(idx, bi) and np.array_equal(d2, bd)
    cl.close()


def test_search_nn_wrapper(scene, orc):
    # a5: widen + NaN / zero-normal rejection.  The synthetic cloud holds 0.1% zero normals.
    q = scene["q"][:6000]
    l, ok = scene["cloud"].search_nn(q)
    oi, _ = scene["tree"].nn(q)
    nrm = np.linalg.norm(scene["world"][oi, 4:7].astype(np.float64), axis=1)
    assert np.array_equal(ok, nrm >= 1e-6)
    assert np.array_equal(l[:, :3], scene["world"][oi, :3].astype(np.float64))
    assert np.array_equal(l[:, 3:], scene["world"][oi, 4:7].astype(np.float64))
    ok1, l1, i1 = scene["tree"].search_nn(q[0])
    assert ok1 == ok[0] and np.array_equal(l1, l[0])


# ---------------------------------------------------------------- association (a6-a8) ------
@pytest.mark.parametrize("variant", [0, 1, 2])
def test_associate_matches_reference_loops(scene, orc, variant):
    rng = orc.search_range(scene["opt_num"])
    got = scene["cloud"].associate(scene["q"], range_=None if variant == 2 else rng, variant=variant)
    want = scene["tree"].associate(scene["q"], None if variant == 2 else rng, variant=variant)
    assert np.array_equal(got["accepted"], want["accepted"])
    acc = want["accepted"].astype(bool)
    assert 0.3 < acc.mean() < 0.999
    assert np.array_equal(got["idx"][acc], want["idx"][acc])
    assert np.array_equal(got["type"][acc], want["type"][acc])
    assert np.array_equal(got["rgb"][acc], want["rgb"][acc])
    assert set(np.unique(want["type"][acc])) == {1, 2}
    for f in ("xyz", "abcd", "dist", "angle"):
        assert _close(got[f][acc], want[f][acc]), f
    # in practice the fp64 fields are bit-identical (same operation order, no FMA)
    assert np.array_equal(got["abcd"][acc], want["abcd"][acc])


def test_associate_idx_is_minus_one_beyond_range(scene, orc):
    """include/pcd_b200.h: idx = -1 when there is no neighbour within the search range (the reference finds the
    farther point and drops it at bundle_adjustment.cc:400); the unbounded NN of those queries is still >= 0."""
    q = scene["q"]
    rng = np.full(len(q), 0.05)
    got = scene["cloud"].associate(q, range_=rng, variant=0)
    oi, _ = scene["tree"].nn(q)
    l = scene["world"][oi, 0:3].astype(np.float64)
    v = l - q
    dist = np.sqrt(v[:, 0] ** 2 + (v[:, 1] ** 2 + v[:, 2] ** 2))
    beyond = dist > rng
    assert beyond.sum() > 1000 and (~beyond).sum() > 1000
    assert (got["idx"][beyond] == -1).all() and not got["accepted"][beyond].any()
    assert np.array_equal(got["idx"][~beyond], oi[~beyond])


def test_associate_opt_num_sweep_equals_explicit_range(scene, orc):
    a = scene["cloud"].associate(scene["q"][:20000], opt_num=scene["opt_num"][:20000], sweep=(1.5, 0.2, 0.1), variant=1)
    b = scene["cloud"].associate(scene["q"][:20000], range_=orc.search_range(scene["opt_num"][:20000]), variant=1)
    for f in a:
        assert np.array_equal(a[f], b[f], equal_nan=True) if a[f].dtype.kind == "f" else np.array_equal(a[f], b[f])


def test_associate_multi_chunk_pipeline(scene, orc, pcd, ctx):
    # > 2^20 queries exercises the double-buffered H2D / compute / D2H pipeline
    rng = np.random.default_rng(3)
    sel = rng.integers(0, len(scene["q"]), 2_300_000)
    q = scene["q"][sel] + rng.normal(0, 0.02, (len(sel), 3))
    got = scene["cloud"].associate(q, range_=0.6, variant=0, fields=("idx", "accepted", "abcd"))
    want = scene["tree"].associate(q, 0.6, variant=0)
    assert np.array_equal(got["accepted"], want["accepted"])
    acc = want["accepted"].astype(bool)
    assert np.array_equal(got["idx"][acc], want["idx"][acc])
    assert np.array_equal(got["abcd"][acc], want["abcd"][acc])


def test_associate_evaluate_fused_pass(scene, orc, pcd, ctx):
    """pcd_associate_evaluate_batch == association + weights + cost function on the accepted matches; multi-chunk."""
    rng_ = np.random.default_rng(5)
    reps = 28                                   # 1.12 M queries: crosses the 1 Mi chunk boundary
    q = np.tile(scene["q"], (reps, 1)) + rng_.normal(0, 1e-3, (reps * len(scene["q"]), 3))
    on = np.tile(scene["opt_num"], reps)
    W = (10.0, 1000.0, 10000.0)
    rec, r, J, s = scene["cloud"].associate_evaluate(q, W, opt_num=on, sweep=(1.5, 0.2, 0.1), variant=pcd.ASSOC_GLOBAL)
    want = scene["tree"].associate(q, orc.search_range(on), variant=1)
    acc = want["accepted"].astype(bool)
    assert np.array_equal(rec["accepted"].astype(bool), acc) and np.array_equal(rec["idx"][acc], want["idx"][acc])
    assert _close(rec["abcd"][acc], want["abcd"][acc]) and _close(rec["dist"][acc], want["dist"][acc])
    w = np.where(want["type"] == 2, W[2], W[1])
    ro, Jo = orc.resjac(q[acc], want["abcd"][acc], w[acc])
    assert _close(r[acc], ro) and _close(J[acc], Jo)
    assert not r[~acc].any() and not J[~acc].any()          # no residual block for a rejected query
    assert _close(s, float(np.sum(ro * ro)), 1e-9)
    # and it equals the three separate calls
    w3 = ctx.weights_from_type(rec["type"], *W) * rec["accepted"]
    r3, J3, _ = ctx.residual_jacobian(q[acc], rec["abcd"][acc], w3[acc])
    assert np.array_equal(r3, r[acc]) and np.array_equal(J3, J[acc], equal_nan=True)


def test_associate_evaluate_device_pointers(scene, orc, pcd, ctx):
    """pcd_associate_evaluate_batch_dev (what bench.py times) == the host-pointer fused pass == oracle."""
    import torch
    q, on = scene["q"], scene["opt_num"]
    n = len(q)
    W = (10.0, 1000.0, 10000.0)
    d_q = torch.from_numpy(q).cuda(); d_on = torch.from_numpy(on.astype(np.int32)).cuda()
    o = {"idx": torch.empty(n, dtype=torch.int32, device="cuda"), "accepted": torch.empty(n, dtype=torch.uint8, device="cuda"),
         "type": torch.empty(n, dtype=torch.uint8, device="cuda"), "abcd": torch.empty((n, 4), dtype=torch.float64, device="cuda")}
    d_r = torch.empty(n, dtype=torch.float64, device="cuda"); d_J = torch.empty((n, 3), dtype=torch.float64, device="cuda")
    d_s = torch.zeros(1, dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    scene["cloud"].associate_evaluate_dev(d_q.data_ptr(), 0, d_on.data_ptr(), (1.5, 0.2, 0.1), n, pcd.ASSOC_GLOBAL, W,
                                          {k: v.data_ptr() for k, v in o.items()}, d_r.data_ptr(), d_J.data_ptr(), d_s.data_ptr(), st)
    torch.cuda.synchronize()
    rec, r, J, s = scene["cloud"].associate_evaluate(q, W, opt_num=on, sweep=(1.5, 0.2, 0.1), variant=pcd.ASSOC_GLOBAL)
    assert np.array_equal(o["accepted"].cpu().numpy(), rec["accepted"]) and np.array_equal(o["idx"].cpu().numpy(), rec["idx"])
    assert np.array_equal(d_r.cpu().numpy(), r) and np.array_equal(d_J.cpu().numpy(), J)
    assert d_s.item() == s
    want = scene["tree"].associate(q, orc.search_range(on), variant=1)
    acc = want["accepted"].astype(bool)
    w = np.where(want["type"] == 2, W[2], W[1])
    ro, Jo = orc.resjac(q[acc], want["abcd"][acc], w[acc])
    assert _close(r[acc], ro) and _close(J[acc], Jo) and _close(s, float(np.sum(ro * ro)), 1e-9)


def test_pack_blocks_f32_wire_format(ctx):
    """pcd_pack_blocks_f32_dev: {r, J0, J1, J2} rounded to nearest float (the multi-GPU exchange's fp32 wire format);
    bit-exact against numpy's float64 -> float32 conversion, NaN / inf / denormal-range values included."""
    import torch
    rng = np.random.default_rng(5)
    n = 100_003
    r = rng.normal(0, 50, n); J = rng.normal(0, 1e3, (n, 3))
    r[:6] = [0.0, -0.0, np.nan, np.inf, 1e-46, 3.5e38]; J[:3, 1] = [np.nan, -np.inf, 1e-40]
    d_r = torch.from_numpy(r).cuda(); d_J = torch.from_numpy(J).cuda()
    d_o = torch.full((n, 4), -1.0, dtype=torch.float32, device="cuda")
    ctx.pack_blocks_f32_dev(d_r.data_ptr(), d_J.data_ptr(), n, d_o.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    with np.errstate(over="ignore"):
        want = np.concatenate([r[:, None], J], 1).astype(np.float32)
    got = d_o.cpu().numpy()
    assert np.array_equal(got, want, equal_nan=True)                      # NaN payloads may differ between cvt.rn and the host
    assert np.array_equal(np.signbit(got[~np.isnan(want)]), np.signbit(want[~np.isnan(want)]))
    ok = np.isfinite(want) & (np.abs(want) > 1e-30)
    assert np.all(np.abs(want[ok] - np.concatenate([r[:, None], J], 1)[ok]) <= 2.0 ** -24 * np.abs(want[ok]) * 1.0000001)


def test_match_proj(ctx, orc):
    rng = np.random.default_rng(11)
    n = 3000
    counts = rng.integers(0, 5, n); cs = np.concatenate([[0], np.cumsum(counts)])
    p = rng.normal(0, 3, (n, 3))
    cand = np.concatenate([rng.normal(0, 3, (cs[-1], 3)), rng.normal(0, 1, (cs[-1], 3))], 1)
    found, xyz, abcd, dist, ang = ctx.match_proj(p, cand, cs)
    for i in range(n):
        ok, x, a, d, g = orc.match_proj(p[i], cand[cs[i]:cs[i + 1]])
        assert ok == bool(found[i])
        if ok:
            assert _close(xyz[i], x) and _close(abcd[i], a) and _close(dist[i], d) and _close(ang[i], g)


# ---------------------------------------------------------------- cost function (a18-a19) --
def test_residual_jacobian(scene, ctx, orc):
    rec = scene["tree"].associate(scene["q"], 1.5, variant=0)
    acc = rec["accepted"].astype(bool)
    p, abcd = scene["q"][acc], rec["abcd"][acc]
    w = ctx.weights_from_type(rec["type"][acc], 10.0, 1000.0, 10000.0)
    assert set(np.unique(w)) == {1000.0, 10000.0}
    r, J, s = ctx.residual_jacobian(p, abcd, w)
    ro, Jo = orc.resjac(p, abcd, w)
    assert _close(r, ro) and _close(J, Jo)
    assert np.array_equal(r, ro) and np.array_equal(J, Jo)       # bit-identical in practice
    assert abs(s - float(np.sum(ro * ro))) <= 1e-9 * s
    # closed form: J = w * sign(n.p + d) * n
    sgn = np.sign((p * abcd[:, :3]).sum(1) + abcd[:, 3])
    assert np.allclose(J, (w * sgn)[:, None] * abcd[:, :3], rtol=1e-12, atol=0)
    # run-to-run determinism of the reduction
    assert ctx.residual_jacobian(p, abcd, w)[2] == s


def test_residual_jacobian_nan_at_zero_and_fd(ctx, orc):
    abcd = np.array([[0.0, -1.0, 0.0, 1.6], [0.6, 0.0, 0.8, -2.0]])
    p = np.array([[3.0, 1.6, 4.0], [1.0, 5.0, 1.0]])     # first point lies exactly on its plane
    w = np.array([1000.0, 10.0])
    r, J, _ = ctx.residual_jacobian(p, abcd, w)
    ro, Jo = orc.resjac(p, abcd, w)
    assert r[0] == 0.0 and np.isnan(J[0]).all() and np.isnan(Jo[0]).all()     # F12: 0 * inf
    assert _close(r, ro) and _close(J, Jo)
    eps = 1e-6
    for k in range(3):
        dp = p.copy(); dp[1, k] += eps
        fd = (ctx.residual_jacobian(dp, abcd, w)[0][1] - r[1]) / eps
        assert abs(fd - J[1, k]) < 1e-4 * max(1.0, abs(J[1, k]))


def test_pose_jacobian_extension_fd(ctx):
    # north_star extension, no reference counterpart: checked against finite differences only
    rng = np.random.default_rng(2)
    n = 50
    q = rng.normal(0, 1, (n, 4)); q /= np.linalg.norm(q, axis=1, keepdims=True)
    t = rng.normal(0, 2, (n, 3)); pc = rng.normal(0, 5, (n, 3))
    nn = rng.normal(0, 1, (n, 3)); nn /= np.linalg.norm(nn, axis=1, keepdims=True)
    abcd = np.concatenate([nn, rng.normal(0, 1, (n, 1))], 1); w = np.full(n, 100.0)
    r, Jp, Jq = ctx.residual_pose_jacobian(pc, q, t, abcd, w)
    eps = 1e-7
    for k in range(7):
        q2, t2 = q.copy(), t.copy()
        if k < 4: q2[:, k] += eps
        else: t2[:, k - 4] += eps
        r2, _, _ = ctx.residual_pose_jacobian(pc, q2, t2, abcd, w)
        fd = (r2 - r) / eps
        assert np.allclose(fd, Jq[:, k], rtol=2e-4, atol=2e-3), k


def test_filter_outliers(scene, ctx, orc):
    rec = scene["tree"].associate(scene["q"][:10000], 1.5, variant=0)
    typ = rec["type"].copy(); typ[::7] = 0
    e, cnt = ctx.filter_outliers(scene["q"][:10000], rec["xyz"], typ, 0.05, 0.12)
    eo, co = orc.filter_outliers(scene["q"][:10000], rec["xyz"], typ, 0.05, 0.12)
    assert np.array_equal(e, eo) and cnt == co and 0 < cnt < 10000


# ---------------------------------------------------------------- depth projection (a9-a15) -
@pytest.fixture(scope="module")
def proj_scene(scene, orc, synth, pcd):
    opt_g, opt_o = pcd.ProjOptions.default(), orc.ProjOptions.default()
    n_cells = scene["cloud"].build_submap(opt_g)
    sub = orc.Submap(scene["world"], opt_o)
    assert n_cells == sub.n_cells
    cams = synth.make_cameras(6, scene["info"])
    coef = orc.splat_coef(cams[0]["cam"], opt_o)
    assert np.array_equal(coef, pcd.splat_coef(cams[0]["cam"], opt_g))
    return dict(sub=sub, cams=cams, coef=coef, opt_g=opt_g, opt_o=opt_o)


def _winners(z, px):
    idx = np.full(len(px), -1, np.int32); rng = np.full(len(px), np.inf, np.float32)
    ok = px >= 0
    v = z.ravel()[px[ok]]
    hit = v != np.iinfo(np.uint64).max
    ii = np.where(hit, (v & np.uint64(0xffffffff)).astype(np.int64), -1).astype(np.int32)
    rr = np.where(hit, (v >> np.uint64(32)).astype(np.uint32).view(np.float32), np.inf).astype(np.float32)
    idx[ok] = ii; rng[ok] = rr
    return idx, rng


def test_zbuffer_dense_bit_exact(scene, proj_scene, pcd):
    ps = proj_scene
    for cam in ps["cams"][:3]:
        im = pcd.Image.make(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"])
        _, _, z, stats = scene["cloud"].project_image(im, ps["coef"], None, dense=True)
        zo, counters, n_sel = ps["sub"].project(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"], ps["coef"])
        assert stats[0] == n_sel and stats[1] == counters[0] and stats[2] == counters[1]
        assert counters[0] > 1000
        assert np.array_equal(z, zo)
        assert (z != np.iinfo(np.uint64).max).mean() > 0.05


def test_zbuffer_feature_winners_bit_exact(scene, proj_scene, pcd, orc, synth):
    ps = proj_scene
    feats = synth.make_features(3000)
    for cam in ps["cams"]:
        im = pcd.Image.make(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"])
        widx, wr, _, stats = scene["cloud"].project_image(im, ps["coef"], feats)
        px = orc.feature_pixels(feats, 0.2, cam["width"], cam["height"])
        s = 0.2; ih, iw = int(cam["height"] * s), int(cam["width"] * s)
        mask = np.zeros(ih * iw, np.uint8); mask[px[px >= 0]] = 1
        zo, _, _ = ps["sub"].project(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"], ps["coef"], mask)
        oi, orng = _winners(zo, px)
        assert np.array_equal(widx, oi)
        assert np.array_equal(wr.view(np.uint32), orng.view(np.uint32))
        assert (widx >= 0).sum() > 200 and stats[3] == (widx >= 0).sum()
        # a15 ray/plane on the winners
        m, xyz = scene["cloud"].ray_plane(im, feats, widx)
        assert np.array_equal(m, widx >= 0)
        for k in np.flatnonzero(m)[:50]:
            assert _close(xyz[k], orc.ray_plane(scene["world"], int(widx[k]), cam["cam"], feats[k, 0], feats[k, 1]))


def test_project_images_batch_equals_single(scene, proj_scene, pcd, synth):
    ps = proj_scene
    feats = [synth.make_features(500 + 100 * k, seed=100 + k) for k in range(len(ps["cams"]))]
    ims = [pcd.Image.make(c["qvec"], c["tvec"], c["cam"], c["width"], c["height"]) for c in ps["cams"]]
    fs = np.concatenate([[0], np.cumsum([len(f) for f in feats])])
    widx, wr, stats = scene["cloud"].project_images(ims, ps["coef"], np.concatenate(feats), fs)
    for k, im in enumerate(ims):
        a, b, _, _ = scene["cloud"].project_image(im, ps["coef"], feats[k])
        assert np.array_equal(widx[fs[k]:fs[k + 1]], a) and np.array_equal(wr[fs[k]:fs[k + 1]], b)


def test_zbuffer_invariant_to_segment_length(scene, proj_scene, pcd, ctx, monkeypatch):
    """The projection's work list cuts every sub-map cell into segments (common.cuh SUBMAP_SEG); winners, the dense map
    and the counters (a selected cell counts once, however many segments it has) must not depend on the cut."""
    ps = proj_scene
    n = len(scene["world"])
    assert n / ps["sub"].n_cells > 32                  # cells hold many points: short segments really split them
    cam = ps["cams"][0]
    im = pcd.Image.make(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"])
    zo, counters, n_sel = ps["sub"].project(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"], ps["coef"])
    for seg in ("32", "65536"):
        monkeypatch.setenv("PCD_B200_SUBMAP_SEG", seg)
        cl = pcd.Cloud(ctx, scene["world"]); cl.build_submap(ps["opt_g"])
        _, _, z, stats = cl.project_image(im, ps["coef"], None, dense=True)
        assert np.array_equal(z, zo) and stats[0] == n_sel and stats[1] == counters[0] and stats[2] == counters[1]


def test_zbuffer_full_res_and_empty_frustum(scene, orc, synth, pcd, ctx):
    # depth_image_scale = 1 at 1920x1080 (config 4 shape), plus a camera looking away from everything
    opt_g = pcd.ProjOptions.default(depth_image_scale=1.0); opt_o = orc.ProjOptions.default(depth_image_scale=1.0)
    cl = pcd.Cloud(ctx, scene["world"]); cl.build_submap(opt_g)
    sub = orc.Submap(scene["world"], opt_o)
    cams = synth.make_cameras(2, scene["info"], cam=synth.OPENCV_1080)
    coef = orc.splat_coef(cams[0]["cam"], opt_o)
    cam = cams[1]
    im = pcd.Image.make(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"])
    _, _, z, stats = cl.project_image(im, coef, None, dense=True)
    zo, counters, n_sel = sub.project(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"], coef)
    assert z.shape == (1080, 1920) and np.array_equal(z, zo) and stats[2] == counters[1]
    far = pcd.Image.make([1, 0, 0, 0], [0, 0, 5000.0], cam["cam"], cam["width"], cam["height"])
    widx, _, z, stats = cl.project_image(far, coef, synth.make_features(100, cam=synth.OPENCV_1080), dense=True)
    assert stats[0] == 0 and (widx == -1).all() and (z == np.iinfo(np.uint64).max).all()
    cl.close()


# ---------------------------------------------------------------- several devices behind the C ABI (pcd_group_*)
def _group_devices():
    import torch
    n = torch.cuda.device_count()
    return [0, 1] if n >= 2 else [0, 0]            # one GPU: two shards on the same device (two contexts)


def test_group_association_equals_single_device(scene, orc, pcd, ctx):
    """pcd_group_associate_evaluate_batch (cloud + index replicated by device-to-device copies, queries in contiguous
    shards, one host thread per device) == pcd_associate_evaluate_batch on one device, bit for bit; sum r^2 is the
    shards' sums added in rank order."""
    rng_ = np.random.default_rng(9)
    reps = 6
    q = np.tile(scene["q"], (reps, 1)) + rng_.normal(0, 1e-3, (reps * len(scene["q"]), 3))
    on = np.tile(scene["opt_num"], reps)
    W = (10.0, 1000.0, 10000.0)
    grp = pcd.Group(_group_devices())
    try:
        grp.cloud_from_host(scene["raw"], direction_trans=True)
        o, r, J, s = grp.associate_evaluate(q, W, opt_num=on, sweep=(1.5, 0.2, 0.1), variant=pcd.ASSOC_GLOBAL)
        rec, r1, J1, s1 = scene["cloud"].associate_evaluate(q, W, opt_num=on, sweep=(1.5, 0.2, 0.1), variant=pcd.ASSOC_GLOBAL)
        for k in ("idx", "accepted", "type", "xyz", "abcd", "dist"):
            assert np.array_equal(o[k], rec[k], equal_nan=(o[k].dtype.kind == "f")), k
        assert np.array_equal(r, r1) and np.array_equal(J, J1, equal_nan=True)
        half = len(q) // 2
        _, _, _, sa = scene["cloud"].associate_evaluate(q[:half], W, opt_num=on[:half], sweep=(1.5, 0.2, 0.1), variant=pcd.ASSOC_GLOBAL)
        _, _, _, sb = scene["cloud"].associate_evaluate(q[half:], W, opt_num=on[half:], sweep=(1.5, 0.2, 0.1), variant=pcd.ASSOC_GLOBAL)
        assert s == sa + sb and _close(s, s1, 1e-12)
        want = scene["tree"].associate(q, orc.search_range(on), variant=1)
        assert np.array_equal(o["accepted"], want["accepted"])
    finally:
        grp.close()


def test_group_projection_equals_single_device(scene, proj_scene, pcd, synth):
    ps = proj_scene
    feats = [synth.make_features(400 + 50 * k, seed=200 + k) for k in range(len(ps["cams"]))]
    ims = [pcd.Image.make(c["qvec"], c["tvec"], c["cam"], c["width"], c["height"]) for c in ps["cams"]]
    fs = np.concatenate([[0], np.cumsum([len(f) for f in feats])])
    want = scene["cloud"].project_images(ims, ps["coef"], np.concatenate(feats), fs)
    grp = pcd.Group(_group_devices())
    try:
        grp.cloud_from_host(scene["raw"], direction_trans=True, opt=pcd.ProjOptions.default())
        got = grp.project_images(ims, ps["coef"], np.concatenate(feats), fs)
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1].view(np.uint32), want[1].view(np.uint32))
        assert np.array_equal(got[2], want[2])
    finally:
        grp.close()


def test_cached_ply_load(tmp_path, scene, pcd, ctx, orc):
    """N3: pcd_cloud_load_ply_cached -- first load parses the PLY and writes the sidecar, the second uses it; both give
    the cloud, the index and the sub-map of the plain loader; a changed source file is parsed again."""
    raw = scene["raw"][:60_000]
    src = str(tmp_path / "map.ply")
    pcd.ply_write(src, raw, binary=False)                              # ASCII: the expensive parse
    opt = pcd.ProjOptions.default()
    c0 = pcd.Cloud.load_ply(ctx, src, opt)
    c1, hit1 = pcd.Cloud.load_ply_cached(ctx, src, opt)
    c2, hit2 = pcd.Cloud.load_ply_cached(ctx, src, opt)
    assert (hit1, hit2) == (False, True) and os.path.exists(src + ".pcdb200")
    w0 = c0.download()
    assert np.array_equal(c1.download().view(np.uint32), w0.view(np.uint32)) and np.array_equal(c2.download().view(np.uint32), w0.view(np.uint32))
    q = scene["q"][:5000]
    i0, d0 = c0.nn(q); i2, d2 = c2.nn(q)
    assert np.array_equal(i0, i2) and np.array_equal(d0, d2)
    bi, _ = orc.nn_brute(w0, q[:300])
    assert np.array_equal(i2[:300], bi)
    pcd.ply_write(src, raw[:50_000], binary=False)
    c3, hit3 = pcd.Cloud.load_ply_cached(ctx, src, opt)
    assert not hit3 and len(c3) < len(c0)
    for c in (c0, c1, c2, c3):
        c.close()
