"""Golden data at the BASELINE.json configurations, from the UNMODIFIED reference (oracle/_ref/libref_bdpt.so = cpu_bdpt.cpp,
object.cpp compiled where they lie).  Run here, where /root/reference exists (about half an hour on 8 cores):

    python tests/golden/make_baseline_golden.py [c1] [c3] [c5]

  c1 -> film_c1.npz   two films of input.txt at C1 = 512 x 512, 16 spp, spl 8, eye / light depth 8 / 8, all host threads
                      (dynamic OpenMP scheduling makes the eye streams of the two runs independent), stored as float16
  c3 -> film_c3.npz   two films of the generated C3 OBJ scene (bench.c3_scene(): gen.write_obj_scene seed 1, ~1 M triangles,
                      parsed through tinyobjloader) at 120 x 68, 8 spp, spl 8, depth 16 / 16 (the reduced film BASELINE.md
                      section 3 proposes; cpu_find_closest_hit stays tractable through the adapter's spatial `G` groups)
  c5 -> hits_c5.npz   closest hits (obj id, t) and visibility of 10 000 random rays / segments (gen.random_rays seed 7) against
                      the 10 M-triangle soup of C5 (gen.soup_scene seed 12345) by the reference's brute-force intersector
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.join(HERE, "..", "..")
sys.path.insert(0, ROOT)
from oracle import oracle_py as O                      # noqa: E402
from path_tracing_b200 import gen                      # noqa: E402
from path_tracing_b200.scene import reference_scene    # noqa: E402

what = set(sys.argv[1:]) or {"c1", "c3", "c5"}
O.build()
R = O.Ref()
threads = len(os.sched_getaffinity(0))


def desc(sd):
    f = sd.to_flat()
    return O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])


def quiet_render(h, d, depth, W, H, spp, spl):
    devnull = os.open(os.devnull, os.O_WRONLY); saved = os.dup(1)
    sys.stdout.flush(); os.dup2(devnull, 1)             # the reference prints per-thread timings
    try:
        return R.render(h, d.cam, depth, depth, W, H, spp, spl, nthreads=threads)
    finally:
        os.dup2(saved, 1); os.close(devnull); os.close(saved)


if "c1" in what:
    d = desc(reference_scene("cornell")); h = R.scene(d)
    cfg = dict(W=512, H=512, spp=16, spl=8, depth=8)
    films = []
    for k in range(2):
        t0 = time.time()
        films.append(quiet_render(h, d, cfg["depth"], cfg["W"], cfg["H"], cfg["spp"], cfg["spl"]))
        print("c1 film", k, "mean", films[-1].mean(axis=(0, 1)), f"{time.time() - t0:.0f} s", flush=True)
    R.destroy(h)
    np.savez_compressed(os.path.join(HERE, "film_c1.npz"), a=films[0].astype(np.float16), b=films[1].astype(np.float16),
                        cfg=np.array([cfg["W"], cfg["H"], cfg["spp"], cfg["spl"], cfg["depth"], cfg["depth"]], np.int32))

if "c3" in what:
    import bench
    sd = bench.c3_scene()
    d = desc(sd); h = R.scene(d)
    cfg = dict(W=120, H=68, spp=8, spl=8, depth=16)
    films = []
    for k in range(2):
        t0 = time.time()
        films.append(quiet_render(h, d, cfg["depth"], cfg["W"], cfg["H"], cfg["spp"], cfg["spl"]))
        print("c3 film", k, "mean", films[-1].mean(axis=(0, 1)), f"{time.time() - t0:.0f} s", flush=True)
    R.destroy(h)
    np.savez_compressed(os.path.join(HERE, "film_c3.npz"), a=films[0], b=films[1], n_tris=np.int64(len(sd.tris)),
                        cfg=np.array([cfg["W"], cfg["H"], cfg["spp"], cfg["spl"], cfg["depth"], cfg["depth"]], np.int32))

if "c5" in what:
    sd = gen.soup_scene(10_000_000)
    d = desc(sd); h = R.scene(d)
    n = 10_000
    org, dr, p2 = gen.random_rays(n, 0.0, 1.0, seed=7)
    ids = np.zeros(n, np.int32); t = np.zeros(n, np.float32); vis = np.zeros(n, np.float32)
    t0 = time.time()
    # chunks over a thread pool: the shim's closest_hit loops over rays on one thread, ctypes releases the GIL
    from concurrent.futures import ThreadPoolExecutor
    chunks = [slice(i, min(i + 50, n)) for i in range(0, n, 50)]

    def work(s):
        r = R.closest_hit(h, org[s], dr[s])
        ids[s] = r["id"]; t[s] = r["t"]
        vis[s] = R.visibility(h, org[s], p2[s])
        return s.start

    with ThreadPoolExecutor(threads) as ex:
        for k, s0 in enumerate(ex.map(work, chunks)):
            if k % 20 == 0:
                print("c5 rays", s0, f"{time.time() - t0:.0f} s", flush=True)
    R.destroy(h)
    np.savez_compressed(os.path.join(HERE, "hits_c5.npz"), id=ids, t=t, vis=vis.astype(np.int8), n_tris=np.int64(len(sd.tris)))
    print("c5: hit fraction", float((ids >= 0).mean()), "visible fraction", float(vis.mean()), f"{time.time() - t0:.0f} s")

if "c5s" in what:       # short segments (length 0.004): about half of them are visible, unlike box-to-box segments through 10 M triangles
    sd = gen.soup_scene(10_000_000)
    d = desc(sd); h = R.scene(d)
    n = 10_000
    org, dr, _ = gen.random_rays(n, 0.0, 1.0, seed=7)
    p2 = (org + dr * np.float32(0.004)).astype(np.float32)
    vis = np.zeros(n, np.float32)
    from concurrent.futures import ThreadPoolExecutor
    chunks = [slice(i, min(i + 50, n)) for i in range(0, n, 50)]

    def work_s(s):
        vis[s] = R.visibility(h, org[s], p2[s])

    t0 = time.time()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(work_s, chunks))
    R.destroy(h)
    g = dict(np.load(os.path.join(HERE, "hits_c5.npz")))
    g["vis_short"] = vis.astype(np.int8)
    np.savez_compressed(os.path.join(HERE, "hits_c5.npz"), **g)
    print("c5s: visible fraction", float(vis.mean()), f"{time.time() - t0:.0f} s")
