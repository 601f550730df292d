import sys, os, numpy as np
sys.path.insert(0, "/root/repo")
import __graft_entry__ as e
synth = e._load("pcd_synth", os.path.join(e.PKG, "synth.py"))
N = int(sys.argv[1]); Q = int(sys.argv[2])
raw, info = synth.make_cloud(N, seed=synth.SEED0 + 4, lidar_frame=False, bad_frac=0)
world = raw
q, on = synth.make_queries(world, Q, seed=synth.SEED0 + 100)
world[:, :3].astype(np.float32).tofile("pts.f32")
q.astype(np.float32).tofile("q.f32")
rad = np.maximum(0.2, 1.5 - 0.1 * on.astype(np.float64)).astype(np.float32)
rad.tofile("rad.f32")
print(info, world[:, :3].min(0), world[:, :3].max(0))
