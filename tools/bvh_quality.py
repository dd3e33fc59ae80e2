"""Tree quality on the CPU (tests/hostsim = the device headers compiled for the host): nodes / triangles per ray of the
wide BVH on the C3 scene for camera rays and shadow segments, per builder setting.  usage: python tools/bvh_quality.py [n_tris] [radius,...] [leaf,...]"""
import sys, time
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import numpy as np
import bench
from hostsim_py import HostSim

n_tris = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
radii = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [2]
leafs = [int(x) for x in sys.argv[3].split(",")] if len(sys.argv) > 3 else [3]
bench.C3["tris"] = n_tris
sd = bench.c3_scene()
rng = np.random.default_rng(5)
cam = sd.camera
eye = np.array(cam["eye"], np.float32); look = np.array(cam["look_at"], np.float32)
n_cam = 40000
fwd = look - eye; fwd /= np.linalg.norm(fwd)
d = fwd[None, :] + rng.uniform(-0.6, 0.6, (n_cam, 3)).astype(np.float32)      # camera rays: a cone around the view direction
d /= np.linalg.norm(d, axis=1, keepdims=True)
o = np.repeat(eye[None, :], n_cam, 0)
for radius in radii:
    for leaf in leafs:
        t0 = time.time()
        hs = HostSim(sd, max_leaf=leaf | (radius << 8))
        tb = time.time() - t0
        ids, t, cnt = hs.intersect(o, d.astype(np.float32), 0)
        hit = ids >= 0
        P = (o + d * t[:, None])[hit] - d[hit] * 1e-3
        k = 100000                                                            # shadow segments between visible surface points / lights
        a = P[rng.integers(0, len(P), k)]; b = P[rng.integers(0, len(P), k)]
        L = np.array(sd.lights["pos"], np.float32)
        b[: k // 3] = L[rng.integers(0, len(L), k // 3)]
        vis, _, cnt2 = hs.intersect(a.astype(np.float32), b.astype(np.float32), 1)
        info = hs.bvh_info()
        print(f"radius {radius} leaf {leaf}: build {tb:.1f}s nodes {info['n_nodes']} depth {info['depth']} | camera: {cnt[0]/n_cam:.2f} nodes {cnt[1]/n_cam:.2f} tris per ray | "
              f"shadow: {cnt2[0]/k:.2f} nodes {cnt2[1]/k:.2f} tris per ray, visible {vis.mean():.2f}", flush=True)
        hs.close()
