"""Full-size checks (BASELINE.json configs[2]: 10 M-point cloud, 1 M SfM points) through size-independent
properties, since the CPU oracle cannot answer these sizes in seconds:
  * exhaustive search on the GPU in plain torch float32 (same operation order) for a sample of the queries:
    indices and float distances bit-exact;
  * self-queries: every cloud point finds a point at distance 0 whose index is <= its own (ties -> lowest);
  * the reported d2 is the float distance to the reported index; radius-bounded answers are a subset of the
    unbounded ones and agree where present; the 14-radius drop sweep only ever shrinks the accepted set;
  * record / cost-function identities: |n| = 1, plane through the lidar point, r = w |n.p + d|, J = +-w n;
  * z-buffer: feature winners == the dense z-buffer read at the feature pixels; every winner lies in range.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
N_CLOUD, N_Q = 10_000_000, 1_000_000


@pytest.fixture(scope="module")
def big(pcd, ctx, synth):
    import torch
    raw, info = synth.make_cloud(N_CLOUD, seed=synth.SEED0 + 2)
    cloud = pcd.Cloud(ctx, raw, direction_trans=True)
    del raw
    cloud.build_index()
    world = cloud.download()
    q, on = synth.make_queries(world, N_Q, seed=synth.SEED0 + 102)
    idx, d2 = cloud.nn(q)
    yield dict(cloud=cloud, world=world, q=q, on=on, idx=idx, d2=d2, info=info, torch=torch)
    cloud.close()


def test_full_size_nn_equals_exhaustive_search_on_sample(big):
    torch = big["torch"]
    w = torch.from_numpy(big["world"][:, 0:3].copy()).cuda()
    x, y, z = w[:, 0].contiguous(), w[:, 1].contiguous(), w[:, 2].contiguous()
    rng = np.random.default_rng(3)
    sample = rng.choice(N_Q, 600, replace=False)
    qf = big["q"][sample].astype(np.float32)
    for k, i in enumerate(sample):
        dx, dy, dz = qf[k, 0] - x, qf[k, 1] - y, qf[k, 2] - z
        d = (dx * dx + dy * dy) + dz * dz                     # separate float32 kernels: no contraction
        dmin = d.min()
        j = int(torch.nonzero(d == dmin)[0, 0])               # lowest index among ties
        assert big["idx"][i] == j, (i, big["idx"][i], j)
        assert np.float32(dmin.item()).view(np.uint32) == big["d2"][i].view(np.uint32)


def test_full_size_self_queries_and_distance_identity(big):
    world = big["world"]
    sel = np.arange(0, len(world), 7)[:1_200_000]
    idx, d2 = big["cloud"].nn(world[sel, 0:3].astype(np.float64))
    assert (d2 == 0).all() and (idx <= sel).all()
    assert np.array_equal(world[idx, 0:3], world[sel, 0:3])   # same coordinates (a duplicate with a lower index)
    # d2 is the float distance to the reported point, in the reference's operation order
    qf = big["q"].astype(np.float32); p = world[big["idx"], 0:3]
    dx, dy, dz = qf[:, 0] - p[:, 0], qf[:, 1] - p[:, 1], qf[:, 2] - p[:, 2]
    assert np.array_equal(((dx * dx + dy * dy) + dz * dz).view(np.uint32), big["d2"].view(np.uint32))
    assert (big["idx"] >= 0).all()


def test_full_size_radius_and_sweep_properties(big, pcd):
    cloud, q, on = big["cloud"], big["q"], big["on"]
    prev = None
    for k in (0, 5, 13):                                       # passes of the drop sweep: opt_num += k
        rng_k = np.where(1.5 - 0.1 * (on + k) <= 0.2, 0.2, 1.5 - 0.1 * (on + k))
        ib, db = cloud.nn(q, radius=rng_k)
        hit = ib >= 0
        assert np.array_equal(ib[hit], big["idx"][hit]) and np.array_equal(db[hit], big["d2"][hit])
        l = big["world"][big["idx"], 0:3].astype(np.float64)
        v = l - q
        dist = np.sqrt(v[:, 0] ** 2 + (v[:, 1] ** 2 + v[:, 2] ** 2))
        assert np.array_equal(hit, ~(dist > rng_k))            # exactly the reference's accept test
        if prev is not None:
            assert not (hit & ~prev).any()                     # a smaller range never adds matches
        prev = hit
    rec = cloud.associate(q, opt_num=on, sweep=(1.5, 0.2, 0.1), variant=pcd.ASSOC_GLOBAL)
    acc = rec["accepted"].astype(bool)
    assert np.array_equal(rec["idx"][acc], big["idx"][acc])
    n = rec["abcd"][acc, 0:3]
    assert np.allclose(np.linalg.norm(n, axis=1), 1.0, rtol=0, atol=1e-12)
    assert np.allclose((n * rec["xyz"][acc]).sum(1) + rec["abcd"][acc, 3], 0.0, atol=1e-9)


def test_full_size_cost_function_identities(big, ctx, pcd):
    q, on = big["q"], big["on"]
    W = (10.0, 1000.0, 10000.0)
    rec, r, J, s = big["cloud"].associate_evaluate(q, W, opt_num=on, sweep=(1.5, 0.2, 0.1), variant=pcd.ASSOC_GLOBAL)
    acc = rec["accepted"].astype(bool)
    w = np.where(rec["type"] == 2, W[2], W[1])[acc]
    u = -((q[acc] * rec["abcd"][acc, 0:3]).sum(1) + rec["abcd"][acc, 3])
    assert np.allclose(r[acc], w * np.abs(u), rtol=1e-9, atol=0)
    ok = u != 0
    assert np.allclose(J[acc][ok], (-np.sign(u) * w)[ok, None] * rec["abcd"][acc, 0:3][ok], rtol=1e-9, atol=0)
    assert not r[~acc].any() and abs(s - float((r * r).sum())) <= 1e-9 * s


def test_full_size_zbuffer_consistency(big, pcd, synth):
    cloud = big["cloud"]
    opt = pcd.ProjOptions.default()
    cloud.build_submap(opt)
    cams = synth.make_cameras(3, big["info"])
    coef = pcd.splat_coef(cams[0]["cam"], opt)
    feats = synth.make_features(3000)
    for cam in cams:
        im = pcd.Image.make(cam["qvec"], cam["tvec"], cam["cam"], cam["width"], cam["height"])
        widx, wr, _, stats = cloud.project_image(im, coef, feats)
        _, _, z, dstats = cloud.project_image(im, coef, None, dense=True)
        assert dstats[1] == stats[1] and dstats[2] == stats[2]                 # same points, same splats
        u = (feats[:, 0] * 0.2).astype(np.int64); v = (feats[:, 1] * 0.2).astype(np.int64)   # truncation
        inb = (feats[:, 0] * 0.2 > -1) & (feats[:, 1] * 0.2 > -1) & (u >= 0) & (u < z.shape[1]) & (v >= 0) & (v < z.shape[0])
        inb &= (feats[:, 0] >= 0) & (feats[:, 1] >= 0)
        zz = z[v[inb], u[inb]]
        exp_idx = np.where(zz == np.iinfo(np.uint64).max, -1, (zz & 0xffffffff).astype(np.int64)).astype(np.int32)
        assert np.array_equal(widx[inb], exp_idx)
        hit = widx >= 0
        assert hit.sum() > 100 and (wr[hit] >= 0.5).all() and np.isfinite(wr[hit]).all()
        assert np.array_equal(wr[inb][exp_idx >= 0].view(np.uint32), (zz[exp_idx >= 0] >> np.uint64(32)).astype(np.uint32))
