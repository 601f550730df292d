"""GPU exploration: k_nn_search time vs cell size, x sub-division and lanes-per-query (run under gpurun).

usage: explore_nn.py n_cloud n_queries cells xsubs groups [radius|unbounded]
(EXPLORE_ENVS="A=1;A=2" repeats the sweep once per ;-separated set of environment knobs)
"""
import os
import sys
import zlib
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry

pcd = entry.package()
synth = entry._load("pcd_synth", os.path.join(entry.PKG, "synth.py"))
arg = sys.argv[1:] + [""] * 8
n_cloud = int(arg[0] or 20_000_000)
Q = int(arg[1] or 8_000_000)
cells = [float(x) for x in (arg[2] or "0.125,0.25,0.5").split(",")]
xsubs = [int(x) for x in (arg[3] or "1,4").split(",")]
groups = [int(x) for x in (arg[4] or "1,2").split(",")]
mode = arg[5] or "radius"
ctx = pcd.Context(0)
raw, _ = synth.make_cloud(n_cloud, seed=synth.SEED0 + 4)
cloud = pcd.Cloud(ctx, raw, direction_trans=True)
del raw
cloud.build_index()
world = cloud.download()
q, on = synth.make_queries(world, Q, seed=synth.SEED0 + 100, outlier_frac=float(os.environ.get('EXPLORE_OUTLIERS', '0.05')))
d_q = torch.from_numpy(q).cuda()
rad = np.maximum(0.2, 1.5 - 0.1 * on.astype(np.float64))
d_rad = torch.from_numpy(rad).cuda() if mode == "radius" else None
d_idx = torch.empty(Q, dtype=torch.int32, device="cuda")
d_d2 = torch.empty(Q, dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
ref = None
envs = [e for e in os.environ.get("EXPLORE_ENVS", "").split(";")] or [""]     # e.g. "PCD_B200_SORT_SKIP=0;PCD_B200_SORT_SKIP=4"
for env in envs:
  for kv in filter(None, env.split(",")):
    os.environ[kv.split("=")[0]] = kv.split("=")[1]
  print("env:", env or "(default)", flush=True)
  for cell in cells:
      for xs in xsubs:
          os.environ["PCD_B200_XSUB"] = str(xs)
          try:
              info = cloud.build_index(cell)
          except pcd.PcdError as e:
              print(f"cell {cell} xs {xs}: {e}")
              continue
          for g in groups:
              if True:
                  os.environ["PCD_B200_NN_K1"] = str(g)
                  rp = d_rad.data_ptr() if d_rad is not None else 0
                  for _ in range(2):
                      cloud.nn_dev(d_q.data_ptr(), rp, Q, d_idx.data_ptr(), d_d2.data_ptr(), st)
                  torch.cuda.synchronize()
                  ctx.profile(True)
                  for _ in range(3):
                      cloud.nn_dev(d_q.data_ptr(), rp, Q, d_idx.data_ptr(), d_d2.data_ptr(), st)
                  torch.cuda.synchronize()
                  ms, n = ctx.kernel_time(pcd.K_NN_SEARCH)
                  ks, _ = ctx.kernel_time(pcd.K_NN_KEYSORT)
                  fs, _ = ctx.kernel_time(pcd.K_NN_FAR)
                  ctx.profile(False)
                  res = d_idx.cpu().numpy()
                  if ref is None:
                      ref = res
                  same = bool(np.array_equal(ref, res))
                  print(f"cell {cell:7.4f} xs {xs:2d} cells {info['n_cells']:>11d} K1 {g:2d}: search {ms / n:8.3f} ms  "
                        f"far {fs / n:6.3f} keysort {ks / n:6.3f} ms  total {(ms + fs + ks) / n:7.3f} ms  "
                        f"{Q / ((ms + fs) / n) / 1e3:8.1f} Mq/s (search+far)  hits {(res >= 0).mean():.3f} same={same} "
                        f"crc {zlib.crc32(res.tobytes()):08x}/{zlib.crc32(d_d2.cpu().numpy().tobytes()):08x}", flush=True)
