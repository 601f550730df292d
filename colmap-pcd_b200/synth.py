"""Deterministic synthetic scenes for tests and bench: a "Smith-Hall-shaped"
street scene in the camera-aligned world frame (x right, y DOWN, z forward; ground plane at
y = +1.6 m with normal (0,-1,0)), building boxes with vertical facades, thin cylinders.
Surface density is kept near `density` points/m^2 (5 cm pitch) and the scene area grows with
the requested point count.  No reference data is involved; seeds are fixed per config.
"""
import numpy as np

SEED0 = 20261104
OPENCV_4032 = dict(width=4032, height=3024, cam=[3039.0, 3039.0, 2016.0, 1512.0, 0.05, -0.02, 1e-3, 1e-3])
OPENCV_1080 = dict(width=1920, height=1080, cam=[1447.0, 1447.0, 960.0, 540.0, 0.05, -0.02, 1e-3, 1e-3])


def _unit(v):
    return v / np.linalg.norm(v, axis=1, keepdims=True)


def make_cloud(n, seed=SEED0, density=400.0, bad_frac=1e-3, lidar_frame=True, chunk=4_000_000):
    """Returns AoS8 float32 [n, 8] (x y z 0 nx ny nz 0).  With lidar_frame=True the cloud is in the
    lidar frame (x front, y left, z up) so that PointCloudDirectionTrans maps it to the world
    frame above; bad_frac of the rows get a NaN normal component and bad_frac a zero normal."""
    rng = np.random.Generator(np.random.PCG64(seed))
    total_area = n / density
    L = float(np.sqrt(total_area * 0.5 / 0.75)); W = 0.75 * L           # ground takes ~half the area
    ground_area = L * W
    n_build = max(2, int(round((total_area * 0.45) / 900.0)))
    bx = rng.uniform(0.1 * L, 0.9 * L, n_build); bz = rng.uniform(0.1 * W, 0.9 * W, n_build)
    bw = rng.uniform(10, 30, n_build); bd = rng.uniform(10, 30, n_build); bh = rng.uniform(8, 15, n_build)
    n_cyl = max(4, int(round(total_area * 0.05 / 7.5)))
    cxs = rng.uniform(0, L, n_cyl); czs = rng.uniform(0, W, n_cyl)
    # surface list: (kind, area, params)
    areas = [ground_area]
    for i in range(n_build):
        areas += [bw[i] * bh[i], bw[i] * bh[i], bd[i] * bh[i], bd[i] * bh[i], bw[i] * bd[i]]
    areas += [2 * np.pi * 0.3 * 4.0] * n_cyl
    areas = np.asarray(areas); prob = areas / areas.sum()
    out = np.zeros((n, 8), dtype=np.float32)
    for lo in range(0, n, chunk):
        m = min(chunk, n - lo)
        surf = rng.choice(len(areas), size=m, p=prob)
        a = rng.random(m); b = rng.random(m)
        P = np.empty((m, 3)); N = np.empty((m, 3))
        g = surf == 0
        P[g] = np.stack([a[g] * L, np.full(g.sum(), 1.6), b[g] * W], 1); N[g] = [0, -1, 0]
        s1 = surf - 1
        isb = (surf >= 1) & (s1 < 5 * n_build)
        bi = np.clip(s1 // 5, 0, n_build - 1); face = s1 % 5
        x0 = bx[bi] - bw[bi] / 2; z0 = bz[bi] - bd[bi] / 2
        for f in range(5):
            k = isb & (face == f)
            if not k.any():
                continue
            i = bi[k]
            if f == 0:    # front wall z = z0, normal -z
                P[k] = np.stack([x0[k] + a[k] * bw[i], 1.6 - b[k] * bh[i], z0[k]], 1); N[k] = [0, 0, -1]
            elif f == 1:  # back wall
                P[k] = np.stack([x0[k] + a[k] * bw[i], 1.6 - b[k] * bh[i], z0[k] + bd[i]], 1); N[k] = [0, 0, 1]
            elif f == 2:  # left wall x = x0
                P[k] = np.stack([x0[k], 1.6 - b[k] * bh[i], z0[k] + a[k] * bd[i]], 1); N[k] = [-1, 0, 0]
            elif f == 3:
                P[k] = np.stack([x0[k] + bw[i], 1.6 - b[k] * bh[i], z0[k] + a[k] * bd[i]], 1); N[k] = [1, 0, 0]
            else:         # roof
                P[k] = np.stack([x0[k] + a[k] * bw[i], 1.6 - bh[i], z0[k] + b[k] * bd[i]], 1); N[k] = [0, -1, 0]
        c = surf >= 1 + 5 * n_build
        ci = np.clip(surf - 1 - 5 * n_build, 0, n_cyl - 1)[c]
        th = a[c] * 2 * np.pi
        N[c] = np.stack([np.cos(th), np.zeros(c.sum()), np.sin(th)], 1)
        P[c] = np.stack([cxs[ci] + 0.3 * np.cos(th), 1.6 - b[c] * 4.0, czs[ci] + 0.3 * np.sin(th)], 1)
        P += N * rng.normal(0, 0.005, (m, 1))                       # 5 mm along the normal
        Nn = _unit(N + rng.normal(0, np.deg2rad(2.0), (m, 3)))      # ~2 degrees of normal noise
        out[lo:lo + m, 0:3] = P.astype(np.float32)
        nf = Nn.astype(np.float32)
        nf /= np.linalg.norm(nf, axis=1, keepdims=True)
        out[lo:lo + m, 4:7] = nf
    if bad_frac > 0 and n >= 1000:
        k = max(1, int(n * bad_frac))
        bad = rng.choice(n, size=2 * k, replace=False)
        out[bad[:k], 4 + rng.integers(0, 3, k)] = np.nan
        out[bad[k:], 4:7] = 0.0
    if lidar_frame:   # inverse of (x,y,z) <- (-y,-z,x)
        w = out.copy()
        out[:, 0] = w[:, 2]; out[:, 1] = -w[:, 0]; out[:, 2] = -w[:, 1]
        out[:, 4] = w[:, 6]; out[:, 5] = -w[:, 4]; out[:, 6] = -w[:, 5]
    return out, dict(L=L, W=W, n_build=n_build)


def scene_extent(n, density=400.0):
    """L, W of the ground rectangle make_cloud(n) lays out (enough for make_cameras on ranks that did not generate)."""
    total_area = n / density
    L = float(np.sqrt(total_area * 0.5 / 0.75))
    return dict(L=L, W=0.75 * L)


def make_queries(world_cloud, q, seed=SEED0 + 1, sigma=0.10, outlier_frac=0.05):
    """SfM points: cloud points displaced by N(0, sigma^2 I); outlier_frac of them anywhere in a
    3 m ball instead.  Returns (p float64 [q,3], opt_num int32 [q] ~ U{0..15}), in hash order."""
    rng = np.random.Generator(np.random.PCG64(seed))
    n = len(world_cloud)
    pick = rng.integers(0, n, q)
    p = world_cloud[pick, 0:3].astype(np.float64) + rng.normal(0, sigma, (q, 3))
    k = int(q * outlier_frac)
    if k:
        sel = rng.choice(q, size=k, replace=False)
        d = _unit(rng.normal(0, 1, (k, 3))) * rng.uniform(0, 3.0, (k, 1))
        p[sel] = world_cloud[pick[sel], 0:3].astype(np.float64) + d
    opt_num = rng.integers(0, 16, q).astype(np.int32)
    return p, opt_num


def make_cameras(n_img, info, cam=OPENCV_4032, seed=SEED0 + 2):
    """Poses on a walking path 1.6 m above the ground, 1 m apart, yaw along the path +- 20 deg.
    Returns list of dicts(qvec, tvec, cam, width, height) with world->camera pose."""
    rng = np.random.Generator(np.random.PCG64(seed))
    L, W = info["L"], info["W"]
    out = []
    for i in range(n_img):
        if n_img <= 60:
            cx = 0.2 * L + 1.0 * i; cz = 0.3 * W; heading = np.pi / 2      # walking along +x
        else:
            t = 2 * np.pi * i / n_img
            cx = L / 2 + 0.3 * L * np.cos(t); cz = W / 2 + 0.3 * W * np.sin(t); heading = t + np.pi / 2
        yaw = heading + np.deg2rad(rng.uniform(-20, 20))
        # camera z axis points along (sin yaw, 0, cos yaw): R_wc = rot_y(yaw), R_cw = rot_y(-yaw)
        half = -yaw / 2
        q = np.array([np.cos(half), 0.0, np.sin(half), 0.0])
        c, s = np.cos(-yaw), np.sin(-yaw)
        R = np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])
        t = -R @ np.array([cx, 0.0, cz])
        out.append(dict(qvec=q, tvec=t, cam=list(cam["cam"]), width=cam["width"], height=cam["height"]))
    return out


def make_features(n_feat, cam=OPENCV_4032, seed=SEED0 + 3):
    rng = np.random.Generator(np.random.PCG64(seed))
    xy = np.stack([rng.uniform(-5, cam["width"] + 5, n_feat), rng.uniform(-5, cam["height"] + 5, n_feat)], 1)
    return xy
