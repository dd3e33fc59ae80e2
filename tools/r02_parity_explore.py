"""Numbers behind the thresholds of tests/test_gpu_baseline_configs.py (run on the GPU box): python tools/r02_parity_explore.py [c2] [fast] [tie] [c3] [c5] [c1]"""
import os, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from oracle import oracle_py as O
from path_tracing_b200 import api, gen
from path_tracing_b200.scene import reference_scene, SceneData

what = set(sys.argv[1:]) or {"c2", "fast", "tie", "c3", "c5", "c1"}
GOLD = "tests/golden"


def desc(sd):
    f = sd.to_flat()
    return O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])


def rel_rmse(x, y):
    return np.sqrt(((x - y) ** 2).mean(axis=(0, 1))) / np.maximum((0.5 * (x + y)).mean(axis=(0, 1)), 1e-12)


def box(x, k):
    H, W, _ = x.shape
    return x[: H // k * k, : W // k * k].reshape(H // k, k, W // k, k, 3).mean(axis=(1, 3))


def pixel_stats(fg, fo, tag):
    diff = np.abs(fg - fo); scale = np.abs(fo) * 1e-3 + 1e-4
    print(tag, "mean gpu/oracle", fg.mean(), fo.mean(), "frac within 1e-3", float((diff <= scale).mean()),
          "p99.9 |d|", float(np.quantile(diff, 0.999)), "max |d|", float(diff.max()), "max film", float(fo.max()), flush=True)


if "c2" in what:
    for name in ("mis", "mis_variant"):
        sd = reference_scene(name); P = O.Port(); d = desc(sd); h = P.scene(d)
        t0 = time.time()
        fo, so = P.render(h, d.cam, 8, 8, 1024, 1024, 64, 8, P.opts(rng_mode=P.PHILOX, seed=1, max_delta=256), want_stats=True)
        t1 = time.time()
        sc = api.Scene(sd)
        fg, sg = sc.render_bdpt(width=1024, height=1024, spp=64, spl=8, eye_depth=8, light_depth=8, seed=1)
        print("C2", name, f"cpu {t1 - t0:.1f} s", {k: (sg[k], getattr(so, k)) for k in ("eye_samples", "closest_rays", "shadow_rays", "path_samples")})
        pixel_stats(fg, fo, "C2 " + name)
        P.destroy(h); sc.close()

if "fast" in what:
    for name in ("cornell", "mis"):
        g = np.load(os.path.join(GOLD, f"hits_{name}.npz"))
        sc = api.Scene(reference_scene(name))
        a = sc.visibility(g["p1"], g["p2"]); b = sc.visibility(g["p1"], g["p2"], fast_shadow=True)
        print("fast vs exact on golden segments", name, len(a), "disagree", int((a != b).sum()), "vs golden", int((a != g["vis"].astype(np.int32)).sum()), flush=True)
        sc.close()
    # grazing: segments through points ON triangle edges / vertices of a random soup
    rng = np.random.default_rng(4)
    n = 3000
    c = rng.uniform(-1, 1, (n, 1, 3)); e = rng.normal(size=(n, 3, 3)) * rng.uniform(0.02, 0.2, (n, 1, 1))
    tri = (c + e).astype(np.float32)
    kind = np.ones(n, np.int32); group = rng.integers(0, 4, n).astype(np.int32)
    mtl = np.tile(np.array([0.7, 0.7, 0.7, 0.5, 0.0, 0.0], np.float32), (n, 1))
    L = np.zeros((1, 12), np.float32); L[0, 3:6] = [0, -1, 0]; L[0, 6:9] = 1; L[0, 9] = 1; L[0, 11] = 0.05; L[0, 0:3] = 5
    sd = SceneData.from_flat(kind, group, tri.reshape(n, 9), mtl, L, np.array([0, 0, -3, 0, 0, 0, 0, 1, 0, 50], np.float32), 32, 32)
    sc = api.Scene(sd)
    m = 200000
    ti = rng.integers(0, n, m); w = rng.uniform(0, 1, (m, 1)).astype(np.float32)
    k0 = rng.integers(0, 3, m); k1 = (k0 + 1) % 3
    on_edge = tri[ti, k0] * w + tri[ti, k1] * (1 - w)
    on_edge[: m // 4] = tri[ti[: m // 4], k0[: m // 4]]                       # a quarter of them through a vertex
    dirs = rng.normal(size=(m, 3)).astype(np.float32); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    p1 = (on_edge - dirs * rng.uniform(0.2, 1.0, (m, 1))).astype(np.float32); p2 = (on_edge + dirs * rng.uniform(0.2, 1.0, (m, 1))).astype(np.float32)
    a = sc.visibility(p1, p2); b = sc.visibility(p1, p2, fast_shadow=True); bf = sc.visibility(p1, p2, brute_force=True)
    print("grazing segments", m, "exact bvh != brute", int((a != bf).sum()), "fast != exact", int((a != b).sum()), "visible frac", float(a.mean()), flush=True)
    p1r = rng.uniform(-1.2, 1.2, (m, 3)).astype(np.float32); p2r = rng.uniform(-1.2, 1.2, (m, 3)).astype(np.float32)
    a = sc.visibility(p1r, p2r); b = sc.visibility(p1r, p2r, fast_shadow=True)
    print("random segments", m, "fast != exact", int((a != b).sum()), flush=True)
    sc.close()

if "tie" in what:
    # a ray along +z hits triangle A of group 0 (plane z = 1) and triangle B of group 1 (also z = 1) at exactly t = 1:
    #   v1  group 1 = {B, C in z = 2}: its box starts at z = 1 = best.t -> the reference culls it (tmax <= tmin)     -> A (id 0)
    #   v2  as v1, group 0 also holds D in z = 3 (so A sits on ITS box's near face too; the first group is never culled) -> A (id 0)
    #   v3  group 1 = {B, C' in z = 0.5 off to the side}: box [0.5, 1], entered before best.t -> B replaces A (t <= best) -> B
    def T(z, dx=0.0):
        return [-1 + dx, -1, z, 1 + dx, -1, z, dx, 1, z]
    variants = {"v1": ([T(1.0), T(1.0), T(2.0)], [0, 1, 1]), "v2": ([T(1.0), T(3.0), T(1.0), T(2.0)], [0, 0, 1, 1]), "v3": ([T(1.0), T(1.0), T(0.5, 7.0)], [0, 1, 1])}
    for name, (tris, grp) in variants.items():
        n = len(tris)
        geom = np.array(tris, np.float32); group = np.array(grp, np.int32); kind = np.ones(n, np.int32)
        mtl = np.tile(np.array([0.7, 0.7, 0.7, 0.5, 0.0, 0.0], np.float32), (n, 1))
        L = np.zeros((1, 12), np.float32); L[0, 0:3] = 9; L[0, 3:6] = [0, -1, 0]; L[0, 6:9] = 1; L[0, 9] = 1; L[0, 11] = 0.05
        sd = SceneData.from_flat(kind, group, geom, mtl, L, np.array([0, 0, -3, 0, 0, 0, 0, 1, 0, 50], np.float32), 32, 32)
        org = np.array([[0, 0, 0], [0.1, -0.2, 0], [0, 0, 0.25]], np.float32); dr = np.array([[0, 0, 1]] * 3, np.float32)
        P = O.Port(); h = P.scene(desc(sd)); r = P.closest_hit(h, org, dr)
        out = {"port": (r["id"].tolist(), r["t"].tolist())}
        if O.have_ref():
            Rf = O.Ref(); hr = Rf.scene(desc(sd)); rr = Rf.closest_hit(hr, org, dr); out["ref"] = (rr["id"].tolist(), rr["t"].tolist())
        sc = api.Scene(sd)
        for brute in (False, True):
            ids, t = sc.closest_hit(org, dr, brute_force=brute); out["gpu_brute" if brute else "gpu_bvh"] = (ids.tolist(), t.tolist())
        print("tie", name, out, flush=True)
        sc.close()

if "c3" in what:
    import bench
    sd = bench.c3_scene(); P = O.Port(); d = desc(sd); h = P.scene(d)
    sc = api.Scene(sd)
    W, H, spp, spl, depth = 48, 27, 1, 8, 16
    t0 = time.time()
    fo, so = P.render(h, d.cam, depth, depth, W, H, spp, spl, P.opts(rng_mode=P.PHILOX, seed=1, max_delta=256), want_stats=True)
    print("C3 port", f"{time.time() - t0:.1f} s")
    fg, sg = sc.render_bdpt(width=W, height=H, spp=spp, spl=spl, eye_depth=depth, light_depth=depth, seed=1)
    print("C3 identical streams", {k: (sg[k], getattr(so, k)) for k in ("eye_samples", "closest_rays", "shadow_rays", "path_samples")})
    pixel_stats(fg, fo, "C3")
    gp = os.path.join(GOLD, "film_c3.npz")
    if os.path.exists(gp):
        g = np.load(gp); A = [g["a"], g["b"]]; W, H, spp, spl, de, dl = [int(x) for x in g["cfg"]]
        G = [sc.render_bdpt(width=W, height=H, spp=spp, spl=spl, eye_depth=de, light_depth=dl, seed=k)[0] for k in (101, 102, 103, 104)]
        print("C3 golden means", [a.mean(axis=(0, 1)) for a in A], "gpu means", [x.mean(axis=(0, 1)) for x in G])
        for k in (1, 4):
            print(" box", k, "A-A", rel_rmse(box(A[0], k), box(A[1], k)), "G-G", [rel_rmse(box(G[i], k), box(G[j], k)) for i, j in ((0, 1), (2, 3))],
                  "G-A", [rel_rmse(box(x, k), box(a, k)) for x in G[:2] for a in A])
    P.destroy(h); sc.close()

if "c1" in what and os.path.exists(os.path.join(GOLD, "film_c1.npz")):
    g = np.load(os.path.join(GOLD, "film_c1.npz")); A = [g["a"].astype(np.float32), g["b"].astype(np.float32)]
    W, H, spp, spl, de, dl = [int(x) for x in g["cfg"]]
    sc = api.Scene(reference_scene("cornell"))
    G = [sc.render_bdpt(width=W, height=H, spp=spp, spl=spl, eye_depth=de, light_depth=dl, seed=k)[0] for k in range(101, 109)]
    print("C1 golden means", [a.mean(axis=(0, 1)) for a in A]); print("C1 gpu means", np.array([x.mean(axis=(0, 1)) for x in G]))
    for k in (1, 8):
        print(" box", k, "A-A", rel_rmse(box(A[0], k), box(A[1], k)), "G-G max", np.max([rel_rmse(box(G[i], k), box(G[j], k)) for i in range(8) for j in range(i)], axis=0),
              "G-A max", np.max([rel_rmse(box(x, k), box(a, k)) for x in G for a in A], axis=0))
    sc.close()

if "c5" in what:
    from hostsim_py import HostSim
    t0 = time.time(); sd = gen.soup_scene(10_000_000); print("soup", f"{time.time() - t0:.1f} s", flush=True)
    sc = api.Scene(sd)
    org, dr, p2 = gen.random_rays(1_000_000, 0.0, 1.0, seed=7)
    ids, t = sc.closest_hit(org, dr); vis = sc.visibility(org, p2)
    gp = os.path.join(GOLD, "hits_c5.npz")
    if os.path.exists(gp):
        g = np.load(gp); n = len(g["id"])
        og, dg, pg = gen.random_rays(n, 0.0, 1.0, seed=7)          # the golden file's own draw (a draw of another size is another stream)
        ig, tg = sc.closest_hit(og, dg); vg = sc.visibility(og, pg); vs = sc.visibility(og, (og + dg * np.float32(0.004)).astype(np.float32))
        print("C5 vs reference golden", n, "id mismatch", int((ig != g["id"]).sum()), "t bit mismatch", int((tg.view(np.uint32) != g["t"].view(np.uint32)).sum()),
              "vis mismatch", int((vg != g["vis"].astype(np.int32)).sum()), "short-segment vis mismatch", int((vs != g["vis_short"].astype(np.int32)).sum()),
              "hit fraction", float((g["id"] != -1).mean()), "short visible fraction", float(g["vis_short"].mean()), flush=True)
    t0 = time.time(); hs = HostSim(sd); print("hostsim build", f"{time.time() - t0:.1f} s", flush=True)
    t0 = time.time(); hi, ht, _ = hs.intersect(org, dr, 0); hv, _, _ = hs.intersect(org, p2, 1); print("hostsim 1e6 rays x2", f"{time.time() - t0:.1f} s", flush=True)
    print("C5 vs host BVH restatement 1e6: id mismatch", int((ids != hi).sum()), "t bit mismatch", int((t.view(np.uint32) != ht.view(np.uint32)).sum()), "vis mismatch", int((vis != hv).sum()), flush=True)
    hs.close(); sc.close()

if "small" in what:
    for name, depth, spl in (("cornell", 4, 8), ("cornell", 8, 8), ("mis", 8, 8), ("mis_variant", 8, 8)):
        sd = reference_scene(name); sc = api.Scene(sd); P = O.Port(); d = desc(sd); h = P.scene(d)
        opts = P.opts(rng_mode=P.PHILOX, seed=5, max_delta=256)
        pool_o = P.light_pool(h, d.n_lights, depth, spl, opts); pool_g = sc.light_pool(spl=spl, light_depth=depth, seed=5)
        ok = np.isclose(pool_g, pool_o, rtol=2e-4, atol=1e-6).all(axis=1)
        bad = ~ok
        print("small", name, depth, "pool bad", int(bad.sum()), "of", len(ok), "max rel dev of bad", float((np.abs(pool_g - pool_o) / (np.abs(pool_o) + 1e-3))[bad].max()) if bad.any() else 0.0)
        fo, st = P.render(h, d.cam, depth, depth, 48, 48, 4, spl, opts, want_stats=True)
        fg, sg = sc.render_bdpt(width=48, height=48, spp=4, spl=spl, eye_depth=depth, light_depth=depth, seed=5, max_batch_paths=3000)
        pixel_stats(fg, fo, f"small {name} {depth}")
        P.destroy(h); sc.close()
