"""GPU exploration: k_nn_search time vs cell size and lanes-per-query (run under gpurun)."""
import os
import sys
import time
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry

pcd = entry.package()
synth = entry._load("pcd_synth", os.path.join(entry.PKG, "synth.py"))
n_cloud = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
Q = int(sys.argv[2]) if len(sys.argv) > 2 else 8_000_000
cells = [float(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0.0625, 0.125, 0.25, 0.5]
groups = [int(x) for x in sys.argv[4].split(",")] if len(sys.argv) > 4 else [1, 2, 4, 8, 16]
ctx = pcd.Context(0)
raw, _ = synth.make_cloud(n_cloud, seed=synth.SEED0 + 4)
cloud = pcd.Cloud(ctx, raw, direction_trans=True)
del raw
cloud.build_index()
world = cloud.download()
q, on = synth.make_queries(world, Q, seed=synth.SEED0 + 100)
d_q = torch.from_numpy(q).cuda()
d_idx = torch.empty(Q, dtype=torch.int32, device="cuda")
d_d2 = torch.empty(Q, dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
ref = None
for cell in cells:
    info = cloud.build_index(cell)
    for g in groups:
        os.environ["PCD_B200_NN_GROUP"] = str(g)
        for _ in range(2):
            cloud.nn_dev(d_q.data_ptr(), 0, Q, d_idx.data_ptr(), d_d2.data_ptr(), st)
        torch.cuda.synchronize()
        ctx.profile(True)
        for _ in range(3):
            cloud.nn_dev(d_q.data_ptr(), 0, Q, d_idx.data_ptr(), d_d2.data_ptr(), st)
        torch.cuda.synchronize()
        ms, n = ctx.kernel_time(pcd.K_NN_SEARCH)
        ks, _ = ctx.kernel_time(pcd.K_NN_KEYSORT)
        ctx.profile(False)
        res = d_idx.cpu().numpy()
        if ref is None:
            ref = res
        same = bool(np.array_equal(ref, res))
        print(f"cell {cell:7.4f} cells {info['n_cells']:>11d} G {g:2d}: search {ms / n:8.3f} ms  keysort {ks / n:6.3f} ms  "
              f"{Q / (ms / n) / 1e6:8.1f} Mq/s  same={same}", flush=True)
