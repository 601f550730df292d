"""GPU exploration: k_project time vs the sub-map segment length (run under gpurun).

usage: explore_proj.py n_cloud n_images seg_lens [scale]
"""
import os
import sys
import time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry

pcd = entry.package()
if os.environ.get("EXPLORE_LIB"):          # compare against another build of the library
    pcd.LIB_PATH = os.path.abspath(os.environ["EXPLORE_LIB"])
synth = entry._load("pcd_synth", os.path.join(entry.PKG, "synth.py"))
arg = sys.argv[1:] + [""] * 4
n_cloud = int(arg[0] or 20_000_000)
n_img = int(arg[1] or 12)
segs = [int(x) for x in (arg[2] or "32,64,128,256,1000000").split(",")]
scale = float(arg[3] or 1.0)
ctx = pcd.Context(0)
raw, _ = synth.make_cloud(n_cloud, seed=synth.SEED0 + 4)
cloud = pcd.Cloud(ctx, raw, direction_trans=True)
del raw
opt = pcd.ProjOptions.default(depth_image_scale=scale)
cams = synth.make_cameras(n_img, synth.scene_extent(n_cloud), cam=synth.OPENCV_1080)
imgs = [pcd.Image.make(c["qvec"], c["tvec"], c["cam"], c["width"], c["height"]) for c in cams]
coef = pcd.splat_coef(cams[0]["cam"], opt)
feats = synth.make_features(3000, cam=synth.OPENCV_1080)
fxy = np.concatenate([feats] * n_img); fstart = np.arange(n_img + 1, dtype=np.int64) * len(feats)
ref = None
peak = ctx.atomic_peak(1920 * 1080, 64)
print(f"contention-free coalesced 64-bit atomicMin on one 1080p z-buffer: {peak / 1e9:.1f} G/s", flush=True)
for seg in segs:
  if True:
    os.environ["PCD_B200_SUBMAP_SEG"] = str(min(seg, 65536))
    n_cells = cloud.build_submap(opt)
    cloud.project_images(imgs[:1], coef, feats, fstart[:2])
    ctx.profile(True)
    t0 = time.perf_counter()
    widx, wr, stats = cloud.project_images(imgs, coef, fxy, fstart)
    dt = time.perf_counter() - t0
    k_ms, k_n = ctx.kernel_time(pcd.K_PROJECT)
    ctx.profile(False)
    ctx.profile(True)
    t0 = time.perf_counter()
    _, _, z, dstats = cloud.project_image(imgs[0], coef, None, dense=True)
    dd = time.perf_counter() - t0
    dk_ms, _ = ctx.kernel_time(pcd.K_PROJECT)
    ctx.profile(False)
    if ref is None:
        ref = (widx.copy(), wr.copy(), z.copy(), stats.copy())
    same = (bool(np.array_equal(ref[0], widx)) and bool(np.array_equal(ref[1].view(np.uint32), wr.view(np.uint32)))
            and bool(np.array_equal(ref[2], z)) and bool(np.array_equal(ref[3], stats)))
    print(f"seg {seg:8d}: k_project {k_ms / max(k_n, 1):7.3f} ms/image  {n_img / dt:8.1f} images/s  dense {1e3 * dd:7.2f} ms, kernel {dk_ms:6.3f} ms "
          f"({dstats[2] / (1e-3 * dk_ms) / 1e9:6.1f} G atomics/s in the kernel = {dstats[2] / (1e-3 * dk_ms) / peak:.2f} of the sweep peak)  cells {n_cells}  same={same}", flush=True)
