"""world_size-2 gloo test of the multi-GPU host logic (colmap-pcd_b200/shard.py): queries sharded,
cloud replicated, one exchange step.  The compute on each rank is the CPU oracle (tests may use it);
the sharded + gathered result must equal the single-process result bit for bit."""
import os
import socket
import sys
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, n_q, out_dir):
    sys.path.insert(0, ROOT)
    import __graft_entry__ as entry
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    shard = entry._load("pcd_shard", os.path.join(entry.PKG, "shard.py"))
    synth = entry._load("pcd_synth", os.path.join(entry.PKG, "synth.py"))
    orc = entry._load("orc", os.path.join(ROOT, "oracle", "orc.py"))
    raw, _ = synth.make_cloud(30_000, seed=synth.SEED0)          # replicated cloud
    world_cloud = orc.direction_trans(raw)
    q, on = synth.make_queries(world_cloud, n_q, seed=synth.SEED0 + 5)
    lo, hi = shard.shard_bounds(n_q, world, rank)                 # this rank's contiguous shard
    tree = orc.KdTree(world_cloud)
    rec = tree.associate(q[lo:hi], orc.search_range(on[lo:hi]), variant=1)
    w = np.where(rec["type"] == 2, 10000.0, 1000.0) * rec["accepted"]
    r, J = orc.resjac(q[lo:hi], rec["abcd"], w)
    r = np.nan_to_num(r); J = np.nan_to_num(J)
    block = torch.from_numpy(np.concatenate([r[:, None], J], 1))
    local = torch.tensor([float((r * r).sum()), float(rec["accepted"].sum()), float((rec["type"][rec["accepted"] > 0] == 2).sum())],
                         dtype=torch.float64)
    if n_q % world == 0:                  # equal shards: the bench's exchange step
        got = {}
        for root in (None, 0):            # all-gather, and gather-to-root (what bench.py uses)
            ex = shard.Exchange(hi - lo, "cpu", root=root)
            ex.start(torch.from_numpy(r), torch.from_numpy(J), local)
            r_all, J_all, stats = ex.wait()
            got[root] = torch.cat([r_all[:, None], J_all], 1)
            assert len(r_all) == (n_q if (root is None or rank == 0) else 0)
        if rank == 0:
            assert torch.equal(got[None], got[0])
        allb = got[None]
        # fp32 wire format: the same exchange on rounded float4 blocks (the CUDA pack kernel's arithmetic, restated)
        ex = shard.Exchange(hi - lo, "cpu", root=None, wire="f32",
                            pack=lambda r_, J_, out, st: out.copy_(torch.cat([r_[:, None], J_], 1).to(torch.float32)))
        ex.start(torch.from_numpy(r), torch.from_numpy(J), local, 1)
        r32, J32, stats32 = ex.wait()
        assert r32.dtype == torch.float32 and torch.equal(torch.cat([r32[:, None], J32], 1), allb.to(torch.float32))
        assert torch.equal(stats32, stats)
        stats = stats.clone()
    else:                                 # ragged shards
        allb = shard.gather_ragged(block)
        stats = shard.reduce_stats(float(local[0]), int(local[1]), int(local[2]))
    idx = shard.gather_ragged(torch.from_numpy(rec["idx"]))
    if rank == 0:
        np.savez(os.path.join(out_dir, "sharded.npz"), block=allb.numpy(), idx=idx.numpy(), stats=stats.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_q", [4000, 4001])
def test_two_rank_shard_equals_single_process(tmp_path, orc, synth, n_q):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, n_q, str(tmp_path)), nprocs=2, join=True)
    got = np.load(str(tmp_path / "sharded.npz"))
    raw, _ = synth.make_cloud(30_000, seed=synth.SEED0)
    world_cloud = orc.direction_trans(raw)
    q, on = synth.make_queries(world_cloud, n_q, seed=synth.SEED0 + 5)
    rec = orc.KdTree(world_cloud).associate(q, orc.search_range(on), variant=1)
    w = np.where(rec["type"] == 2, 10000.0, 1000.0) * rec["accepted"]
    r, J = orc.resjac(q, rec["abcd"], w)
    r = np.nan_to_num(r); J = np.nan_to_num(J)
    assert np.array_equal(got["idx"], rec["idx"])
    assert np.array_equal(got["block"], np.concatenate([r[:, None], J], 1))
    assert got["stats"][1] == rec["accepted"].sum()
    assert got["stats"][2] == (rec["type"][rec["accepted"] > 0] == 2).sum()
    assert abs(got["stats"][0] - (r * r).sum()) <= 1e-12 * (r * r).sum()


def test_shard_bounds_cover_and_balance(pcd):
    import importlib.util
    spec = importlib.util.spec_from_file_location("pcd_shard", os.path.join(ROOT, "colmap-pcd_b200", "shard.py"))
    shard = importlib.util.module_from_spec(spec); spec.loader.exec_module(shard)
    for n in (0, 1, 7, 450, 1_000_003):
        for world in (1, 2, 3, 8):
            b = [shard.shard_bounds(n, world, r) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n and all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
    imgs = [shard.shard_images(450, 8, r) for r in range(8)]
    assert sorted(sum(imgs, [])) == list(range(450)) and {len(x) for x in imgs} == {56, 57}
