"""Device headers (BVH build, traversal, BSDF, RNG, wavefront BDPT stage bodies) compiled for the HOST by tests/hostsim
and checked against the reference golden vectors and the oracle.  This is how the kernel logic is validated on the
CPU-only build box; the same bodies run inside the CUDA kernels (tests/test_gpu_parity.py checks those on the GPU)."""
import os
import numpy as np
import pytest
from hostsim_py import HostSim
from oracle import oracle_py as O
from path_tracing_b200.scene import SceneData, reference_scene

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def desc(sd):
    f = sd.to_flat()
    return O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])


@pytest.mark.parametrize("name", ["cornell", "mis"])
def test_traversal_is_bit_exact_against_reference_golden(name):
    g = np.load(os.path.join(GOLD, f"hits_{name}.npz"))
    hs = HostSim(reference_scene(name))
    for mode in (0, 16):      # wide BVH, exhaustive
        ids, t, _ = hs.intersect(g["org"], g["dir"], mode)
        assert np.array_equal(ids, g["id"])
        assert np.array_equal(t.view(np.uint32), g["t"].view(np.uint32))
    for mode in (1, 17):
        vis, _, _ = hs.intersect(g["p1"], g["p2"], mode)
        assert np.array_equal(vis, g["vis"].astype(np.int32))
    hs.close()


def random_soup(n, seed, lights=2):
    rng = np.random.default_rng(seed)
    c = rng.uniform(-1, 1, (n, 1, 3)); e = rng.normal(size=(n, 3, 3)) * rng.uniform(0.01, 0.2, (n, 1, 1))
    geom = (c + e).reshape(n, 9).astype(np.float32)
    kind = np.ones(n, np.int32); group = rng.integers(0, 3, n).astype(np.int32)
    mtl = np.tile(np.array([0.7, 0.7, 0.7, 0.5, 0.0, 0.0], np.float32), (n, 1))
    mtl[rng.uniform(size=n) < 0.2, 5] = 1.5     # transparent to shadow rays
    L = np.zeros((lights, 12), np.float32)
    L[:, 0:3] = rng.uniform(-1, 1, (lights, 3)); L[:, 3:6] = [0, -1, 0]; L[:, 6:9] = 1; L[:, 9] = 1.0; L[:, 11] = 0.05
    cam = np.array([0, 0, -3, 0, 0, 0, 0, 1, 0, 50], np.float32)
    return SceneData.from_flat(kind, group, geom, mtl, L, cam, 32, 32)


@pytest.mark.parametrize("n,max_leaf", [(1, 3), (2, 1), (3, 3), (500, 3), (4000, 2)])
def test_bvh_equals_brute_force_and_oracle_on_triangle_soup(n, max_leaf):
    sd = random_soup(n, 100 + n)
    hs = HostSim(sd, max_leaf=max_leaf)
    rng = np.random.default_rng(n)
    m = 3000
    org = rng.uniform(-1.2, 1.2, (m, 3)).astype(np.float32)
    dr = rng.normal(size=(m, 3)).astype(np.float32); dr /= np.linalg.norm(dr, axis=1, keepdims=True)
    dr[:50, 1:] = 0; dr[:50, 0] = 1                     # axis-aligned directions (zero components)
    a_id, a_t, cnt = hs.intersect(org, dr, 0)
    b_id, b_t, _ = hs.intersect(org, dr, 16)
    assert np.array_equal(a_id, b_id) and np.array_equal(a_t.view(np.uint32), b_t.view(np.uint32))
    P = O.Port(); h = P.scene(desc(sd))
    r = P.closest_hit(h, org, dr)
    assert np.array_equal(a_id, r["id"]) and np.array_equal(a_t.view(np.uint32), r["t"].view(np.uint32))
    p2 = rng.uniform(-1.2, 1.2, (m, 3)).astype(np.float32)
    v_bvh, _, _ = hs.intersect(org, p2, 1)
    v_bf, _, _ = hs.intersect(org, p2, 17)
    assert np.array_equal(v_bvh, v_bf) and np.array_equal(v_bvh, P.visibility(h, org, p2).astype(np.int32))
    if n >= 500:
        assert cnt[1] / m < 0.25 * n                    # the tree actually culls
    P.destroy(h); hs.close()


def test_threaded_flatten_of_large_triangle_scenes_equals_the_general_path(monkeypatch):
    """csrc/b2_host_scene.hpp: triangle-only scenes of >= 2^15 triangles take flatten_triangles_fast (two threaded passes, no
    per-object tables); B2_FLATTEN_GENERAL=1 forces the general path.  Same group boxes, visiting ranks and records <=> the same
    tree, the same hits (ids, t bits, visibility) and the same visit counters; and both equal the CPU intersector."""
    sd = random_soup(40_000, 77, lights=1)
    rng = np.random.default_rng(5)
    m = 4000
    org = rng.uniform(-1.2, 1.2, (m, 3)).astype(np.float32)
    dr = rng.normal(size=(m, 3)).astype(np.float32); dr /= np.linalg.norm(dr, axis=1, keepdims=True)
    p2 = rng.uniform(-1.2, 1.2, (m, 3)).astype(np.float32)
    out = []
    for general in ("0", "1"):
        monkeypatch.setenv("B2_FLATTEN_GENERAL", general)
        for threads in (("1", "3") if general == "0" else ("1",)):
            monkeypatch.setenv("B2_FLATTEN_THREADS", threads)
            hs = HostSim(sd)
            ids, t, cnt = hs.intersect(org, dr, 0)
            vis, _, cnt2 = hs.intersect(org, p2, 1)
            out.append((ids, t.view(np.uint32), vis, tuple(cnt), tuple(cnt2), hs.bvh_info()))
            hs.close()
    for o in out[1:]:
        assert all(np.array_equal(a, b) for a, b in zip(o[:3], out[0][:3])) and o[3:] == out[0][3:]
    P = O.Port(); h = P.scene(desc(sd))
    k = 300
    r = P.closest_hit(h, org[:k], dr[:k])
    assert np.array_equal(out[0][0][:k], r["id"]) and np.array_equal(out[0][1][:k], r["t"].view(np.uint32))
    assert np.array_equal(out[0][2][:k], P.visibility(h, org[:k], p2[:k]).astype(np.int32))
    P.destroy(h)


@pytest.mark.parametrize("name,depth", [("cornell", 6), ("mis", 8), ("mis_variant", 8)])
def test_wavefront_bodies_reproduce_the_oracle_render(name, depth):
    sd = reference_scene(name)
    hs = HostSim(sd)
    P = O.Port(); d = desc(sd); h = P.scene(d)
    opts = P.opts(rng_mode=P.PHILOX, seed=5, max_delta=256)
    pool_o = P.light_pool(h, d.n_lights, depth, 8, opts)
    pool_h = hs.light_pool(8, depth, seed=5)
    assert np.allclose(pool_h, pool_o, rtol=1e-5, atol=1e-7)
    W = H = 24
    fo, st = P.render(h, d.cam, depth, depth, W, H, 3, 8, opts, want_stats=True)
    fh, cnt = hs.render(W, H, 3, 8, depth, depth, seed=5, max_batch_paths=500)    # several batches
    assert cnt["eye_samples"] == st.eye_samples == W * H * 3
    assert cnt["closest_rays"] == st.closest_rays and cnt["shadow_rays"] == st.shadow_rays and cnt["path_samples"] == st.path_samples
    assert cnt["pretest_mismatch"] <= 1e-5 * cnt["pairs"]     # division-free record pre-test == the reference-order pre-test
    assert np.allclose(fh, fo, rtol=1e-4, atol=1e-6)
    # sample sharding: two shards (s = 0,2 and s = 1) summed == the whole film
    a, _ = hs.render(W, H, 3, 8, depth, depth, seed=5, sample_offset=0, sample_stride=2, film_mode=1)
    b, _ = hs.render(W, H, 3, 8, depth, depth, seed=5, sample_offset=1, sample_stride=2, film_mode=1)
    assert np.allclose((a + b) / 3.0, fo, rtol=1e-4, atol=1e-6)
    # the MIS weight with both walks hoisted out of the pair loop (what k_pool_active + k_connect run): same terms, fp32 rounding only
    ff, cf = hs.render(W, H, 3, 8, depth, depth, seed=5, max_batch_paths=500, factored=True)
    assert cf["shadow_rays"] == cnt["shadow_rays"] and cf["path_samples"] == cnt["path_samples"]
    assert np.allclose(ff, fh, rtol=2e-5, atol=1e-7)
    P.destroy(h); hs.close()


def test_factored_mis_weight_reproduces_the_reference_golden_weights():
    """The connection stage evaluates cpu_calculate_mis_weight (cpu_bdpt.cpp:112-167) with both subpath walks hoisted out of the
    pair loop: each walk is linear in the one pdf that depends on the pair (mis_eye_sum / mis_light_sum, csrc/b2_bdpt.cuh).
    On the 3 000 chains the unmodified reference produced (arbitrary pdf_fwd / pdf_rev at every vertex, delta vertices,
    pseudo-source vertices): the sequential form on the host, the factored form, and the golden weights agree - the same
    vertices give a zero weight, and the non-zero ones are within fp32 rounding of each other."""
    from hostsim_py import mis_weight
    g = np.load(os.path.join(GOLD, "mis_tables.npz"))
    args = (g["eye"], g["s_idx"], g["lit"], g["t_idx"], g["d_vec"], g["dist2"], g["cam_pos"])
    seq, fac = mis_weight(*args), mis_weight(*args, factored=True)
    want = g["w"]
    assert ((seq > 0) == (want > 0)).all() and ((fac > 0) == (want > 0)).mean() > 0.999
    live = (want > 0) & (fac > 0)
    assert live.mean() > 0.3
    assert np.allclose(seq[live], want[live], rtol=2e-5, atol=0.0)
    assert np.allclose(fac[live], seq[live], rtol=2e-5, atol=0.0)


def test_traversal_stacks_stay_inside_the_short_shared_memory_stacks(tmp_path):
    """The kernels give every thread bvh_depth - 2 (+ 1 spare) stack entries of shared memory (stack_entries, csrc/b2_cuda_util.hpp):
    a walk stacks at most one group of unvisited siblings per level between the root's children and the current node's parent.
    The host compilation records the deepest stack of every closest-hit and any-hit walk of a BDPT render on a generated
    mesh scene (tree of depth >= 8) and of random rays through a triangle soup."""
    from path_tracing_b200 import gen
    sd = SceneData.from_file(gen.write_obj_scene(str(tmp_path), n_tris=1 << 15, width=48, height=32, tex_size=64), obj_directive=True)
    hs = HostSim(sd)
    depth = hs.bvh_info()["depth"]
    assert depth >= 8
    hs.max_stack_depth(reset=True)
    _, cnt = hs.render(48, 32, 1, 8, 8, 8, seed=2)
    assert cnt["shadow_rays"] > 100_000
    assert 1 <= hs.max_stack_depth() <= depth - 2
    hs.close()
    soup = HostSim(random_soup(20_000, 7))
    d2 = soup.bvh_info()["depth"]
    rng = np.random.default_rng(1)
    org = rng.uniform(-0.2, 1.2, (20_000, 3)).astype(np.float32); dr = rng.normal(size=(20_000, 3)).astype(np.float32)
    soup.max_stack_depth(reset=True)
    soup.intersect(org, dr, 0); soup.intersect(org, org + dr, 1)
    assert 1 <= soup.max_stack_depth() <= d2 - 2
    soup.close()

