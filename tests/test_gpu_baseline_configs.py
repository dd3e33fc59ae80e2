"""Parity gates AT the BASELINE.json configurations, through the C ABI:

  C1  input.txt, 512 x 512, 16 spp, spl 8, depth 8/8: the GPU film against two films of the UNMODIFIED reference
      (tests/golden/film_c1.npz, made by tests/golden/make_baseline_golden.py), SURVEY.md 8(d) gates 1-3
  C2  mis_test.txt (and its documented variant), 1024 x 1024, 64 spp, depth 8/8: EVERY pixel of the bench-size film against a
      live oracle render on identical Philox streams (the CPU needs a few seconds)
  C3  the generated 1 M-triangle OBJ scene: a reduced film against the unmodified reference (film_c3.npz) and, on identical
      streams, against the oracle port pixel by pixel
  C5  10 M-triangle soup: 10 000 rays / segments bit-exact against the reference's brute-force intersector (hits_c5.npz) and
      1 000 000 against the host compilation of the BVH code (tests/hostsim)
plus the two arithmetic points the round-1 review asked to pin: where the render's fast shadow-ray triangle test may differ
from the bit-exact one (and how often), and the group-box tie of cpu_bdpt.cpp:43-44.
Tolerances are stated where they are applied; the measured values behind them are in profiles/r02_parity.md."""
import os
import sys

import numpy as np
import pytest

from oracle import oracle_py as O
from path_tracing_b200 import api, gen
from path_tracing_b200.scene import SceneData, reference_scene

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")


def desc(sd):
    f = sd.to_flat()
    return O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def rel_rmse(x, y):
    return np.sqrt(((x - y) ** 2).mean(axis=(0, 1))) / np.maximum((0.5 * (x + y)).mean(axis=(0, 1)), 1e-12)


def box(x, k):
    H, W, _ = x.shape
    return x[: H // k * k, : W // k * k].reshape(H // k, k, W // k, k, 3).mean(axis=(1, 3))


def assert_same_render(fg, sg, fo, so):
    """GPU film / counters vs an oracle render on IDENTICAL Philox streams.  The arithmetic differs in the last bits (CUDA
    libm, FMA contraction, approximate reciprocals in shading), so a path may take another branch at a discontinuity:
    counters within 1e-4 relative; >= 99.9 % of the pixel channels within 1e-3 relative (+ 1e-4 absolute); and no pixel
    further off than 2 % of the brightest pixel (a bound on the worst case, not only on the fraction)."""
    for k in ("closest_rays", "shadow_rays", "path_samples"):
        assert abs(sg[k] - getattr(so, k)) <= 1e-4 * getattr(so, k) + 4, (k, sg[k], getattr(so, k))
    assert sg["eye_samples"] == so.eye_samples
    d = np.abs(fg - fo)
    assert (d <= 1e-3 * np.abs(fo) + 1e-4).mean() >= 0.999
    assert d.max() <= 0.02 * max(float(fo.max()), 1e-3), (float(d.max()), float(fo.max()))
    assert abs(float(fg.mean()) - float(fo.mean())) <= 1e-4 * float(fo.mean()) + 1e-7


@pytest.mark.parametrize("name", ["mis", "mis_variant"])
def test_c2_full_size_every_pixel_against_live_oracle(name):
    sd = reference_scene(name)
    P = O.Port(); d = desc(sd); h = P.scene(d)
    kw = dict(width=1024, height=1024, spp=64, spl=8, eye_depth=8, light_depth=8, seed=1)
    fo, so = P.render(h, d.cam, 8, 8, 1024, 1024, 64, 8, P.opts(rng_mode=P.PHILOX, seed=1, max_delta=256), want_stats=True)
    sc = api.Scene(sd)
    fg, sg = sc.render_bdpt(**kw)
    assert sg["eye_samples"] == 1024 * 1024 * 64
    assert_same_render(fg, sg, fo, so)
    P.destroy(h); sc.close()


@pytest.fixture(scope="module")
def c3():
    sys.path.insert(0, os.path.join(HERE, ".."))
    import bench
    sd = bench.c3_scene()                       # gen.write_obj_scene (seed 1) -> OBJ + MTL + PPM -> tinyobjloader -> adapter
    sc = api.Scene(sd)
    yield sd, sc
    sc.close()


def test_c3_obj_scene_on_identical_streams(c3):
    sd, sc = c3
    assert len(sd.tris) > 1_000_000
    P = O.Port(); d = desc(sd); h = P.scene(d)
    W, H, spp, spl, depth = 48, 27, 1, 8, 16
    fo, so = P.render(h, d.cam, depth, depth, W, H, spp, spl, P.opts(rng_mode=P.PHILOX, seed=1, max_delta=256), want_stats=True)
    fg, sg = sc.render_bdpt(width=W, height=H, spp=spp, spl=spl, eye_depth=depth, light_depth=depth, seed=1)
    assert so.shadow_rays > 100 * so.eye_samples            # the connection loop is what runs here
    assert_same_render(fg, sg, fo, so)
    P.destroy(h)


# ---- the render's fast shadow-ray test (csrc/b2_isect.cuh tri_test_fast + b2_bvh.cuh group_pass_fast) vs the bit-exact one ----
def test_fast_shadow_test_differs_only_on_grazing_segments():
    """k_connect's shadow rays use contracted FMAs and approximate reciprocals (B2_FAST_SHADOW through b2_intersect).  Accepted:
    NO difference on generic segments; <= 0.5 % on the reference golden segment sets (which contain edge / vertex / axis-aligned
    adversarial rays on purpose: measured 0.26 % and 0.29 %); <= 2 % on segments aimed exactly at triangle edges and vertices
    (measured 1.0 %).  The exact test itself stays equal to the golden answers."""
    for name in ("cornell", "mis"):
        g = np.load(os.path.join(GOLD, f"hits_{name}.npz"))
        sc = api.Scene(reference_scene(name))
        exact = sc.visibility(g["p1"], g["p2"]); fast = sc.visibility(g["p1"], g["p2"], fast_shadow=True)
        assert np.array_equal(exact, g["vis"].astype(np.int32))
        assert (exact != fast).mean() <= 0.005, (name, int((exact != fast).sum()))
        sc.close()
    rng = np.random.default_rng(4)
    n, m = 3000, 200_000
    c = rng.uniform(-1, 1, (n, 1, 3)); e = rng.normal(size=(n, 3, 3)) * rng.uniform(0.02, 0.2, (n, 1, 1))
    tri = (c + e).astype(np.float32)
    mtl = np.tile(np.array([0.7, 0.7, 0.7, 0.5, 0.0, 0.0], np.float32), (n, 1))
    L = np.zeros((1, 12), np.float32); L[0, 0:3] = 5; L[0, 3:6] = [0, -1, 0]; L[0, 6:9] = 1; L[0, 9] = 1; L[0, 11] = 0.05
    sd = SceneData.from_flat(np.ones(n, np.int32), rng.integers(0, 4, n).astype(np.int32), tri.reshape(n, 9), mtl, L,
                             np.array([0, 0, -3, 0, 0, 0, 0, 1, 0, 50], np.float32), 32, 32)
    sc = api.Scene(sd)
    ti = rng.integers(0, n, m); w = rng.uniform(0, 1, (m, 1)).astype(np.float32)
    k0 = rng.integers(0, 3, m); k1 = (k0 + 1) % 3
    on_edge = tri[ti, k0] * w + tri[ti, k1] * (1 - w)
    on_edge[: m // 4] = tri[ti[: m // 4], k0[: m // 4]]                       # a quarter of them through a vertex
    dirs = rng.normal(size=(m, 3)).astype(np.float32); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    p1 = (on_edge - dirs * rng.uniform(0.2, 1.0, (m, 1))).astype(np.float32); p2 = (on_edge + dirs * rng.uniform(0.2, 1.0, (m, 1))).astype(np.float32)
    exact = sc.visibility(p1, p2)
    assert np.array_equal(exact, sc.visibility(p1, p2, brute_force=True))     # the exact path does not care about grazing
    assert (exact != sc.visibility(p1, p2, fast_shadow=True)).mean() <= 0.02
    p1r = rng.uniform(-1.2, 1.2, (m, 3)).astype(np.float32); p2r = rng.uniform(-1.2, 1.2, (m, 3)).astype(np.float32)
    assert np.array_equal(sc.visibility(p1r, p2r), sc.visibility(p1r, p2r, fast_shadow=True))
    sc.close()


# ---- the group-box cull on exact ties (cpu_bdpt.cpp:43-44: intersectAABB(ray, 1e-4, best.t), object.cpp:118 tmax <= tmin) ----
def _tie_scene(tris, groups):
    n = len(tris)
    mtl = np.tile(np.array([0.7, 0.7, 0.7, 0.5, 0.0, 0.0], np.float32), (n, 1))
    L = np.zeros((1, 12), np.float32); L[0, 0:3] = 9; L[0, 3:6] = [0, -1, 0]; L[0, 6:9] = 1; L[0, 9] = 1; L[0, 11] = 0.05
    return SceneData.from_flat(np.ones(n, np.int32), np.array(groups, np.int32), np.array(tris, np.float32), mtl, L,
                               np.array([0, 0, -3, 0, 0, 0, 0, 1, 0, 50], np.float32), 32, 32)


def _T(z, dx=0.0):
    return [-1 + dx, -1, z, 1 + dx, -1, z, dx, 1, z]


@pytest.mark.parametrize("tris,groups,winner", [
    # a ray along +z meets A (group 0, plane z = 1) and B (group 1, plane z = 1) at exactly t = 1
    ([_T(1.0), _T(1.0), _T(2.0)], [0, 1, 1], 0),              # group 1's box STARTS at t = 1 = best.t: the reference culls it -> A
    ([_T(1.0), _T(3.0), _T(1.0), _T(2.0)], [0, 0, 1, 1], 0),  # A on the near face of its own box as well: the first group is never culled -> A
    ([_T(1.0), _T(1.0), _T(0.5, 7.0)], [0, 1, 1], 1),         # group 1's box is entered at 0.5: B is tested, t <= best.t accepted -> B
])
def test_group_box_cull_on_exact_ties(tris, groups, winner):
    sd = _tie_scene(tris, groups)
    org = np.array([[0, 0, 0], [0.1, -0.2, 0], [0, 0, 0.25]], np.float32); dr = np.array([[0, 0, 1]] * 3, np.float32)
    P = O.Port(); h = P.scene(desc(sd)); ref = P.closest_hit(h, org, dr)
    if O.have_ref():
        R = O.Ref(); hr = R.scene(desc(sd)); rr = R.closest_hit(hr, org, dr)
        assert np.array_equal(rr["id"], ref["id"]) and np.array_equal(bits(rr["t"]), bits(ref["t"]))
    assert ref["id"].tolist() == [winner] * 3
    sc = api.Scene(sd)
    for brute in (False, True):
        ids, t = sc.closest_hit(org, dr, brute_force=brute)
        assert ids.tolist() == [winner] * 3 and np.array_equal(bits(t), bits(ref["t"])), (brute, ids, t)
    sc.close()
    P.destroy(h)


def test_c5_hits_bit_exact_against_the_reference_and_the_host_bvh():
    """BASELINE config 5: hit ids bit-exact vs the CPU intersector on the 10 M-triangle soup.  10 000 rays + 10 000 segments against
    the UNMODIFIED reference's brute-force cpu_find_closest_hit / cpu_check_visibility (golden: 36 ms per ray on 8 cores), and
    1 000 000 + 1 000 000 against the host compilation of the same BVH build + traversal code (tests/hostsim), which is itself
    equal to brute force on the subsets of tests/test_hostsim.py."""
    sys.path.insert(0, HERE)
    from hostsim_py import HostSim
    g = np.load(os.path.join(GOLD, "hits_c5.npz"))
    sd = gen.soup_scene(10_000_000)
    assert len(sd.tris) == int(g["n_tris"])
    sc = api.Scene(sd)
    n = len(g["id"])
    assert n >= 10_000
    org, dr, p2 = gen.random_rays(n, 0.0, 1.0, seed=7)
    ids, t = sc.closest_hit(org, dr)
    assert np.array_equal(ids, g["id"]), int((ids != g["id"]).sum())
    hit = g["id"] != -1
    assert 0.2 < hit.mean() < 1.0
    assert np.all(np.abs(t[hit] - g["t"][hit]) <= 1e-5 * np.abs(g["t"][hit])) and np.array_equal(bits(t), bits(g["t"]))
    assert np.array_equal(sc.visibility(org, p2), g["vis"].astype(np.int32))
    p2s = (org + dr * np.float32(0.004)).astype(np.float32)          # short segments: about half are visible (box-to-box ones never are)
    vs = sc.visibility(org, p2s)
    assert np.array_equal(vs, g["vis_short"].astype(np.int32)) and 0.2 < vs.mean() < 0.8
    org, dr, p2 = gen.random_rays(1_000_000, 0.0, 1.0, seed=8)
    ids, t = sc.closest_hit(org, dr); vis = sc.visibility(org, p2)
    hs = HostSim(sd)
    hi, ht, _ = hs.intersect(org, dr, 0); hv, _, _ = hs.intersect(org, p2, 1)
    assert np.array_equal(ids, hi) and np.array_equal(bits(t), bits(ht)) and np.array_equal(vis, hv)
    hs.close(); sc.close()


def _family_gates(G, A, k_box, tag):
    """SURVEY.md 8(d) gates 1-3 for a family of GPU films G (different Philox seeds = different light pools) against films A of the
    unmodified reference at the same configuration.  Both families are samples of one distribution if the GPU path is right, so
    cross distances must look like within-family distances:
      gate 1  raw HDR:               max relRMSE_c(G_k, A_i) <= 1.25 * floor,  floor = max over the pairs inside A and inside G
      gate 2  k_box x k_box filter:  the same, <= max(1.25 * floor, 0.05); and again on films clamped to [0, 1] (the display range:
              fireflies of up to 15 per connection dominate the raw statistic, cpu_bdpt.cpp:437)
      gate 3  every reference film's channel means within 3 sigma (of the GPU family, whose seeds span the pool noise that one
              reference run cannot show: its pool comes from seed 1337 + thread id) of the GPU family's mean"""
    import itertools
    def floors(f):
        inside = [f(x, y) for fam in (A, G) for x, y in itertools.combinations(fam, 2)]
        cross = [f(x, a) for x in G for a in A]
        return np.max(inside, axis=0), np.max(cross, axis=0)
    fl, cr = floors(rel_rmse)
    assert np.all(cr <= 1.25 * fl), (tag, "gate 1", cr, fl)
    fl, cr = floors(lambda x, y: rel_rmse(box(x, k_box), box(y, k_box)))
    assert np.all(cr <= np.maximum(1.25 * fl, 0.05)), (tag, "gate 2", cr, fl)
    fl, cr = floors(lambda x, y: rel_rmse(box(np.clip(x, 0, 1), k_box), box(np.clip(y, 0, 1), k_box)))
    assert np.all(cr <= np.maximum(1.25 * fl, 0.05)), (tag, "gate 2 clamped", cr, fl)
    gm = np.array([x.mean(axis=(0, 1)) for x in G]); mu, sigma = gm.mean(axis=0), gm.std(axis=0, ddof=1)
    for a in A:
        assert np.all(np.abs(a.mean(axis=(0, 1)) - mu) <= 3.0 * sigma * np.sqrt(1.0 + 1.0 / len(G)) + 1e-3 * mu), (tag, "gate 3", a.mean(axis=(0, 1)), mu, sigma)
    for x in G:
        assert np.isfinite(x).all() and x.min() >= 0.0


def test_c1_full_size_against_films_of_the_unmodified_reference():
    """BASELINE config 1 (the PR1 reference render): input.txt, 512 x 512, 16 spp, spl 8, depth 8/8 -- ~450 shadow rays per eye
    sample, the loop-C workload.  Golden: two renders of oracle/_ref (cpu_bdpt.cpp unmodified, all host threads)."""
    g = np.load(os.path.join(GOLD, "film_c1.npz"))
    A = [g["a"].astype(np.float32), g["b"].astype(np.float32)]
    W, H, spp, spl, de, dl = [int(x) for x in g["cfg"]]
    assert (W, H, spp, spl, de, dl) == (512, 512, 16, 8, 8, 8)
    sc = api.Scene(reference_scene("cornell"))
    G = []
    for seed in range(101, 109):
        film, st = sc.render_bdpt(width=W, height=H, spp=spp, spl=spl, eye_depth=de, light_depth=dl, seed=seed)
        G.append(film)
    assert st["eye_samples"] == W * H * spp and st["shadow_rays"] > 300 * st["eye_samples"] and st["delta_cap_hits"] == 0
    _family_gates(G, A, 8, "C1")
    sc.close()


def test_c3_reduced_film_against_films_of_the_unmodified_reference(c3):
    """BASELINE config 3 at the reduced film BASELINE.md proposes: the generated 1 M-triangle OBJ scene, 120 x 68, 8 spp, spl 8,
    depth 16/16, rendered by the unmodified reference (its own group-AABB cull over the adapter's spatial G groups keeps
    cpu_find_closest_hit tractable) -- golden film_c3.npz."""
    sd, sc = c3
    g = np.load(os.path.join(GOLD, "film_c3.npz"))
    assert len(sd.tris) == int(g["n_tris"])                 # the generator is deterministic: same mesh here and where the golden was made
    A = [g["a"], g["b"]]
    W, H, spp, spl, de, dl = [int(x) for x in g["cfg"]]
    G = [sc.render_bdpt(width=W, height=H, spp=spp, spl=spl, eye_depth=de, light_depth=dl, seed=seed)[0] for seed in range(101, 109)]
    _family_gates(G, A, 4, "C3")
