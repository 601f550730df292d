"""pcd_residual_jacobian_batch_dev_noalloc against pcd_residual_jacobian_batch_dev, through the C ABI.

The allocation-free entry point keeps its scratch (the block partial sums) in a fixed device array of the library; it
has no counterpart in the reference.  Collected last on purpose (file name).
Inputs and outputs are torch device tensors made by the TEST; the entry point itself allocates nothing.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL_TOL = 1e-5   # the tolerance of tests/test_gpu_parity.py; the two entry points run the same kernels, so they agree exactly


def _inputs(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    p = rng.uniform(-20, 20, (n, 3))
    nrm = rng.normal(0, 1, (n, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    d = rng.uniform(-5, 5, (n, 1))
    abcd = np.concatenate([nrm, d], axis=1)
    w = rng.choice([10.0, 1000.0, 10000.0], n)
    if n > 8:
        abcd[3, 1] = np.nan            # a NaN plane adds no residual block: r = 0, J = 0
        p[5] = 0.0; abcd[5] = [0.0, 1.0, 0.0, 0.0]   # u == 0: J is NaN, as autodiff gives
    return p, abcd, w


def _run(fn, torch, p, abcd, w, want_sum=True):
    n = len(w)
    d_p = torch.from_numpy(p).cuda(); d_a = torch.from_numpy(abcd).cuda(); d_w = torch.from_numpy(w).cuda()
    d_r = torch.full((n,), -7.0, dtype=torch.float64, device="cuda")
    d_J = torch.full((n, 3), -7.0, dtype=torch.float64, device="cuda")
    d_s = torch.full((1,), -7.0, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    st = torch.cuda.current_stream().cuda_stream
    fn(d_p.data_ptr(), d_a.data_ptr(), d_w.data_ptr(), n, d_r.data_ptr(), d_J.data_ptr(), d_s.data_ptr() if want_sum else 0, st)
    torch.cuda.synchronize()
    return d_r.cpu().numpy(), d_J.cpu().numpy(), float(d_s.cpu().numpy()[0])


def _same(a, b):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    both_nan = np.isnan(a) & np.isnan(b)
    return bool((both_nan | (np.abs(a - b) <= REL_TOL * np.maximum(np.abs(a), np.abs(b)) + 1e-300)).all())


def test_noalloc_residual_jacobian_matches_existing_entry_point(pcd, ctx):
    import torch
    assert pcd.lib().pcd_residual_jacobian_noalloc_max() == 65536 * 256
    # two calls in a row on the shared scratch: a large batch first, then a smaller one with other data -- the second
    # must not see the first one's partial sums
    for n, seed in ((1_000_003, 11), (70_001, 12), (1, 13)):
        p, abcd, w = _inputs(n, seed)
        r0, J0, s0 = _run(ctx.residual_jacobian_dev, torch, p, abcd, w)
        r1, J1, s1 = _run(ctx.residual_jacobian_dev_noalloc, torch, p, abcd, w)
        assert _same(r1, r0) and _same(J1, J0), n
        assert abs(s1 - s0) <= REL_TOL * abs(s0), (n, s0, s1)
        assert s1 > 0 and np.isfinite(s1)
        if n > 8:
            assert r1[3] == 0.0 and (J1[3] == 0.0).all() and np.isnan(J1[5]).all()
    # the same call twice gives the same answer (the array keeps the last call's contents; nothing needs clearing)
    p, abcd, w = _inputs(300_000, 14)
    a = _run(ctx.residual_jacobian_dev_noalloc, torch, p, abcd, w)
    b = _run(ctx.residual_jacobian_dev_noalloc, torch, p, abcd, w)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1], equal_nan=True) and a[2] == b[2]
    # without the sum no scratch is touched
    r2, J2, s2 = _run(ctx.residual_jacobian_dev_noalloc, torch, p, abcd, w, want_sum=False)
    assert np.array_equal(r2, a[0]) and s2 == -7.0


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch(pcd, ctx):
    import torch
    d = torch.zeros(64, dtype=torch.float64, device="cuda")      # never read: the size check comes first
    too_many = pcd.lib().pcd_residual_jacobian_noalloc_max() + 1
    with pytest.raises(pcd.PcdError, match="exceed the fixed scratch"):
        ctx.residual_jacobian_dev_noalloc(d.data_ptr(), d.data_ptr(), d.data_ptr(), too_many, d.data_ptr(), d.data_ptr(), d.data_ptr(), 0)
    # n == 0 is a no-op that zeroes the sum
    s = torch.full((1,), 3.0, dtype=torch.float64, device="cuda")
    ctx.residual_jacobian_dev_noalloc(0, 0, 0, 0, 0, 0, s.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert float(s.cpu()[0]) == 0.0
