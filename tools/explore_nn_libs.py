"""GPU exploration: the NN kernels of several BUILDS of the library on one scene, in one process (run under gpurun).

usage: explore_nn_libs.py n_cloud n_queries lib[,lib...] [radius|unbounded]
Each lib is a path relative to the repo root (e.g. colmap-pcd_b200/build/var/libpcd_b200_x2.so; see
tools/build_nn_variants.sh).  The scene and the device query buffers are made once; every build gets its own context,
cloud upload and index.  Prints the CUDA-event times of the key sort, the search and the far kernel and a CRC of the
index and distance arrays, which must agree between builds.
"""
import os
import sys
import zlib
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry

synth = entry._load("pcd_synth", os.path.join(entry.PKG, "synth.py"))
arg = sys.argv[1:] + [""] * 4
n_cloud = int(arg[0] or 20_000_000)
Q = int(arg[1] or 8_000_000)
libs = [x for x in (arg[2] or "colmap-pcd_b200/libpcd_b200.so").split(",") if x]
mode = arg[3] or "radius"
reps = int(os.environ.get("EXPLORE_REPS", "5"))

raw, _ = synth.make_cloud(n_cloud, seed=synth.SEED0 + 4)
d_q = d_rad = None
d_idx = torch.empty(Q, dtype=torch.int32, device="cuda")
d_d2 = torch.empty(Q, dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
first = None
for k, rel in enumerate(libs):
    os.environ["PCD_B200_LIB"] = os.path.join(ROOT, rel)
    pcd = entry._load(f"pcdb200_build{k}", os.path.join(entry.PKG, "pcdb200.py"))
    ctx = pcd.Context(0)
    cloud = pcd.Cloud(ctx, raw, direction_trans=True)
    os.environ["PCD_B200_XSUB"] = "4"
    info = cloud.build_index(0.25)
    if d_q is None:
        world = cloud.download()
        q, on = synth.make_queries(world, Q, seed=synth.SEED0 + 100, outlier_frac=0.05)
        d_q = torch.from_numpy(q).cuda()
        rad = np.maximum(0.2, 1.5 - 0.1 * on.astype(np.float64))
        d_rad = torch.from_numpy(rad).cuda() if mode == "radius" else None
        del world
    rp = d_rad.data_ptr() if d_rad is not None else 0
    d_idx.fill_(-7)
    for _ in range(2):
        cloud.nn_dev(d_q.data_ptr(), rp, Q, d_idx.data_ptr(), d_d2.data_ptr(), st)
    torch.cuda.synchronize()
    ctx.profile(True)
    for _ in range(reps):
        cloud.nn_dev(d_q.data_ptr(), rp, Q, d_idx.data_ptr(), d_d2.data_ptr(), st)
    torch.cuda.synchronize()
    ms, n = ctx.kernel_time(pcd.K_NN_SEARCH)
    ks, _ = ctx.kernel_time(pcd.K_NN_KEYSORT)
    fs, _ = ctx.kernel_time(pcd.K_NN_FAR)
    ctx.profile(False)
    crc = (zlib.crc32(d_idx.cpu().numpy().tobytes()), zlib.crc32(d_d2.cpu().numpy().tobytes()))
    if first is None:
        first = crc
    print(f"{rel}: search {ms / n:7.3f} ms  far {fs / n:6.3f}  keysort {ks / n:6.3f}  total {(ms + fs + ks) / n:7.3f} ms  "
          f"{Q / ((ms + fs) / n) / 1e3:8.1f} Mq/s (search+far)  crc {crc[0]:08x}/{crc[1]:08x} same={crc == first}", flush=True)
    del cloud, ctx
