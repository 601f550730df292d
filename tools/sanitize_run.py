"""Small pass over every kernel family of the path, for compute-sanitizer (memcheck / racecheck) under gpurun:
smoke(), one multi-chunk pcd_associate_evaluate_batch (double-buffered H2D / kernels / D2H pipeline, far queue with its
atomic draw), one dense and one feature-mask depth projection, the group path on two contexts."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry

pcd = entry.package()
print("library:", pcd.LIB_PATH, flush=True)
entry.smoke()
synth = entry._load("pcd_synth", os.path.join(entry.PKG, "synth.py"))
ctx = pcd.Context(0)
raw, info = synth.make_cloud(150_000, seed=synth.SEED0 + 31)
cloud = pcd.Cloud(ctx, raw, direction_trans=True)
cloud.build_index()
world = cloud.download()
q, on = synth.make_queries(world, 1_150_000, seed=synth.SEED0 + 32)          # two chunks of the pipeline
rec, r, J, s = cloud.associate_evaluate(q, (10.0, 1000.0, 10000.0), opt_num=on, sweep=(1.5, 0.2, 0.1), variant=pcd.ASSOC_GLOBAL)
print("associate_evaluate:", int(rec["accepted"].sum()), "accepted, sum r^2", s)
idx, d2 = cloud.nn(q[:200_000], 0.7)
print("nn with radius:", int((idx >= 0).sum()), "hits")
opt = pcd.ProjOptions.default(depth_image_scale=1.0)
cloud.build_submap(opt)
cams = synth.make_cameras(2, info, cam=synth.OPENCV_1080)
coef = pcd.splat_coef(cams[0]["cam"], opt)
feats = synth.make_features(2000, cam=synth.OPENCV_1080)
im = pcd.Image.make(cams[1]["qvec"], cams[1]["tvec"], cams[1]["cam"], 1920, 1080)
w, wr, z, st = cloud.project_image(im, coef, feats, dense=True)
print("dense projection:", int((z != np.iinfo(np.uint64).max).sum()), "pixels;", int((w >= 0).sum()), "feature winners")
w2, _, _, _ = cloud.project_image(im, coef, feats, dense=False)
assert np.array_equal(w, w2)
print("atomic sweep:", ctx.atomic_peak(1920 * 1080, 4) / 1e9, "G/s")
grp = pcd.Group([0, 0])
grp.cloud_from_host(raw, direction_trans=True, opt=opt)
o, r2, J2, s2 = grp.associate_evaluate(q[:300_000], (10.0, 1000.0, 10000.0), opt_num=on[:300_000], sweep=(1.5, 0.2, 0.1), variant=pcd.ASSOC_GLOBAL)
assert np.array_equal(r2, r[:300_000])
grp.close()
cloud.close(); ctx.close()
print("sanitize_run ok")
