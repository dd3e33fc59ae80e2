"""B2_STRATEGIES_* beyond the reference set (csrc/b2_strategies.cuh): the t = 1 light-tracing splats the north_star names are
not part of cpu_bdpt.cpp, so there is no oracle for them; they are validated against an independent unbiased estimator
of the same integral built from the parity-tested pieces -- next-event estimation on the eye subpaths.  Both must
converge to the same image; their 1/2 + 1/2 combination (B2_STRATEGIES_FULL) too."""
import numpy as np
import pytest

from path_tracing_b200 import api
from path_tracing_b200.scene import SceneData

pytestmark = pytest.mark.gpu


def quad(a, b, c, d):
    return [a + b + c, a + c + d]


def box_scene(tmp_path):
    L = ["E 0 0 -2.2", "V 0 0 1  0 1 0", "F 40", "R 48 48"]
    X0, X1, Y0, Y1, Z0, Z1 = -1.0, 1.0, -1.0, 1.0, 0.0, 3.0

    def add(mat, quads):
        L.append("M %g %g %g %g %g 0" % mat)
        for q in quads:
            for t in quad(*q):
                L.append("T " + " ".join("%g" % x for x in t))

    add((0.7, 0.7, 0.7, 1.0, 0.0), [((X0, Y0, Z0), (X1, Y0, Z0), (X1, Y0, Z1), (X0, Y0, Z1)),        # floor
                                    ((X0, Y1, Z0), (X0, Y1, Z1), (X1, Y1, Z1), (X1, Y1, Z0)),        # ceiling
                                    ((X0, Y0, Z1), (X1, Y0, Z1), (X1, Y1, Z1), (X0, Y1, Z1))])       # back
    add((0.7, 0.15, 0.15, 0.8, 0.0), [((X0, Y0, Z0), (X0, Y0, Z1), (X0, Y1, Z1), (X0, Y1, Z0))])      # left, red
    add((0.15, 0.6, 0.2, 0.5, 0.0), [((X1, Y0, Z0), (X1, Y1, Z0), (X1, Y1, Z1), (X1, Y0, Z1))])       # right, green, glossier
    # a glossy half-metal block and a rough block standing on the floor
    for (cx, cz, h, w, mat) in ((-0.4, 1.9, 0.9, 0.3, (0.8, 0.7, 0.4, 0.35, 0.5)), (0.45, 1.2, 0.5, 0.28, (0.5, 0.55, 0.8, 0.9, 0.0))):
        x0, x1, z0, z1, y1 = cx - w, cx + w, cz - w, cz + w, Y0 + h
        add(mat, [((x0, Y0, z0), (x0, y1, z0), (x1, y1, z0), (x1, Y0, z0)), ((x1, Y0, z0), (x1, y1, z0), (x1, y1, z1), (x1, Y0, z1)),
                  ((x1, Y0, z1), (x1, y1, z1), (x0, y1, z1), (x0, Y0, z1)), ((x0, Y0, z1), (x0, y1, z1), (x0, y1, z0), (x0, Y0, z0)),
                  ((x0, y1, z0), (x0, y1, z1), (x1, y1, z1), (x1, y1, z0))])
    L.append("L 0 0.93 1.5  0 -1 0  9 9 9  75 0 0.001")
    L.append("L -0.7 0.6 0.6  0.5 -1 0.6  3 4 6  50 0 0.001")
    p = tmp_path / "box.txt"
    p.write_text("\n".join(L) + "\n")
    return SceneData.from_file(str(p))


def blocks(img, b=6):
    H, W, _ = img.shape
    return img.reshape(H // b, b, W // b, b, 3).mean(axis=(1, 3))


def compare(sc, kw, tol_total, tol_block):
    nee, st_n = sc.render_bdpt(strategies=api.STRATEGIES_NEE, **kw)
    lt, st_l = sc.render_bdpt(strategies=api.STRATEGIES_LIGHT_TRACING, **kw)
    full, st_f = sc.render_bdpt(strategies=api.STRATEGIES_FULL, **kw)
    for img in (nee, lt, full):
        assert np.isfinite(img).all() and (img >= 0).all() and img.mean() > 1e-3
    assert st_l["path_samples"] > 100_000 and st_l["eye_samples"] == kw["width"] * kw["height"] * kw["spp"]   # one light subpath per sample
    # total flux on the film: two estimators of the same number
    tn, tl, tf = nee.sum(axis=(0, 1)), lt.sum(axis=(0, 1)), full.sum(axis=(0, 1))
    assert np.allclose(tl, tn, rtol=tol_total), (tl, tn)
    assert np.allclose(tf, 0.5 * (tn + tl), rtol=0.02), (tf, tn, tl)
    # and the same image: 6x6-pixel block means, wherever there is signal
    bn, bl, bf = blocks(nee), blocks(lt), blocks(full)
    lum = bn.mean(axis=2)
    sel = lum > 0.25 * lum.mean()
    rel = np.abs(bl - bn)[sel].mean(axis=1) / np.maximum(bn[sel].mean(axis=1), 1e-6)
    assert sel.sum() >= 40 and np.median(rel) < tol_block and np.quantile(rel, 0.95) < 4 * tol_block, (np.median(rel), np.quantile(rel, 0.95))
    relf = np.abs(bf - 0.5 * (bn + bl))[sel].mean(axis=1) / np.maximum(bn[sel].mean(axis=1), 1e-6)
    assert np.median(relf) < 0.05
    return bl[sel].mean(axis=0) / bn[sel].mean(axis=0)


def test_light_tracing_splats_agree_with_next_event_estimation(tmp_path):
    sd = box_scene(tmp_path)
    sc = api.Scene(sd)
    # (a) one surface vertex (direct light): no BSDF sampling on either side, only the t = 1 machinery -- camera importance,
    #     geometry term, pixel projection, visibility, normalisation of the W*H*spp subpaths -- against I cos / d^2: strict
    ratio = compare(sc, dict(width=48, height=48, spp=512, spl=1, eye_depth=1, light_depth=2, seed=3), tol_total=0.015, tol_block=0.04)
    assert np.allclose(ratio, 1.0, atol=0.02), ratio
    # (b) up to three surface vertices: BSDF sampling on both sides.  These modes use the normalised GGX D (b2_bsdf.cuh,
    #     mat_mark_fix): with the reference's tr_D (tan^2 squared once too often, geometric.cuh:178-185) bsdf_pdf is not the
    #     density of bsdf_sample and the two estimators differ by ~5 % in total and ~18 % per block, each biased on a different subpath.
    ratio = compare(sc, dict(width=48, height=48, spp=512, spl=1, eye_depth=3, light_depth=4, seed=3), tol_total=0.02, tol_block=0.05)
    assert np.allclose(ratio, 1.0, atol=0.03), ratio
    sc.close()


def test_strategies_shard_like_the_reference_set_and_reject_unknown_values(tmp_path):
    sd = box_scene(tmp_path)
    sc = api.Scene(sd)
    kw = dict(width=48, height=48, spp=8, spl=1, eye_depth=3, light_depth=4, seed=5, strategies=api.STRATEGIES_FULL)
    whole, _ = sc.render_bdpt(film_mode=api.FILM_SUM, **kw)
    parts = [sc.render_bdpt(film_mode=api.FILM_SUM, sample_offset=g, sample_stride=2, **kw)[0] for g in range(2)]
    assert np.allclose(parts[0] + parts[1], whole, rtol=1e-4, atol=1e-5)             # same streams on 1 or 2 GPUs, atomics reorder sums
    with pytest.raises(RuntimeError, match="strategies"):
        sc.render_bdpt(width=8, height=8, spp=1, strategies=7)
    ref, _ = sc.render_bdpt(width=48, height=48, spp=8, spl=4, eye_depth=3, light_depth=4, seed=5)   # parity mode is the default and untouched
    assert np.isfinite(ref).all()
    sc.close()


def test_mis_mode_has_the_expectation_of_next_event_estimation_and_less_noise(tmp_path):
    """B2_STRATEGIES_MIS (csrc/b2_mis.cuh): all s/t strategies with balance-heuristic weights over real area
    densities.  Weights that sum to one over the strategies of every path <=> same expectation as any complete single
    strategy; on a scene without specular surfaces that is next-event estimation."""
    sd = box_scene(tmp_path)
    sc = api.Scene(sd)
    for eye_depth, light_depth in ((1, 2), (3, 4), (5, 3), (4, 9)):
        kw = dict(width=48, height=48, spl=1, eye_depth=eye_depth, light_depth=light_depth)
        nee, _ = sc.render_bdpt(strategies=api.STRATEGIES_NEE, spp=1024, seed=3, **kw)
        mis, st = sc.render_bdpt(strategies=api.STRATEGIES_MIS, spp=512, seed=4, **kw)
        mis2, _ = sc.render_bdpt(strategies=api.STRATEGIES_MIS, spp=512, seed=5, **kw)
        nee2, _ = sc.render_bdpt(strategies=api.STRATEGIES_NEE, spp=512, seed=6, **kw)
        assert np.isfinite(mis).all() and (mis >= 0).all() and st["eye_samples"] == 48 * 48 * 512 and st["path_samples"] > 48 * 48 * 256
        tn, tm = nee.sum(axis=(0, 1)), (0.5 * (mis + mis2)).sum(axis=(0, 1))
        assert np.allclose(tm, tn, rtol=0.006), (eye_depth, light_depth, tm / tn)
        bn, bm = blocks(nee), blocks(0.5 * (mis + mis2))
        rel = np.abs(bm - bn).mean(axis=2) / np.maximum(bn.mean(axis=2), 0.25 * bn.mean())
        assert np.median(rel) < 0.015 and rel.max() < 0.08, (eye_depth, light_depth, np.median(rel), rel.max())
        if light_depth > 2 and eye_depth > 1:                           # the point of the exercise: less variance per sample than one strategy alone
            assert np.mean((mis - mis2) ** 2) / 2.0 < np.mean((nee2 - nee) ** 2) / 1.5, (eye_depth, light_depth)      # var at 512 spp: 2 v_mis vs v_nee (1 + 1/2)
    sc.close()


def test_mis_mode_equals_its_host_compilation_and_shards(tmp_path):
    from hostsim_py import HostSim
    sd = box_scene(tmp_path)
    sc = api.Scene(sd)
    hs = HostSim(sd)
    kw = dict(width=32, height=32, spp=8, eye_depth=4, light_depth=5, seed=9)
    want, cw = hs.render_strategies(4, 32, 32, spp=8, eye_depth=4, light_depth=5, seed=9)
    got, st = sc.render_bdpt(strategies=api.STRATEGIES_MIS, spl=1, **kw)
    assert abs(int(st["closest_rays"]) - cw["closest_rays"]) <= 1e-3 * cw["closest_rays"] and abs(int(st["shadow_rays"]) - cw["shadow_rays"]) <= 1e-3 * cw["shadow_rays"]   # a weight that rounds to 0 skips its ray
    assert (np.abs(got - want).max(axis=2) <= 2e-3 * want.mean()).mean() >= 0.99        # same Philox streams; libm vs device sin/cos, atomics reorder the splats
    whole, _ = sc.render_bdpt(strategies=api.STRATEGIES_MIS, spl=1, film_mode=api.FILM_SUM, **kw)
    parts = [sc.render_bdpt(strategies=api.STRATEGIES_MIS, spl=1, film_mode=api.FILM_SUM, sample_offset=g, sample_stride=2, **kw)[0] for g in range(2)]
    assert np.allclose(parts[0] + parts[1], whole, rtol=1e-4, atol=1e-5)
    with pytest.raises(RuntimeError, match="eye_depth"):                  # a subpath stores at most 24 vertices
        sc.render_bdpt(width=8, height=8, spp=1, eye_depth=25, light_depth=4, strategies=api.STRATEGIES_MIS)
    hs.close(); sc.close()


def test_mis_wavefront_equals_the_one_thread_per_sample_kernel(tmp_path, monkeypatch):
    """csrc/b2_mis_wave.cu (the product path: light / eye bounces, pre-test and connection as separate dense kernels over HBM
    records) against csrc/b2_mis.cuh's kernel (B2_MIS_MEGAKERNEL=1) on the same Philox streams: the same paths, so the ray and
    path counters are EQUAL (at most a weight that rounds differently moves one) and the films differ by the order of the film
    atomics only; small batches force several batches and every queue swap."""
    from path_tracing_b200.scene import reference_scene
    for sd, kw in ((box_scene(tmp_path), dict(width=40, height=30, spp=8, eye_depth=4, light_depth=5, seed=3)),
                   (reference_scene("cornell"), dict(width=64, height=64, spp=4, eye_depth=6, light_depth=7, seed=2, max_batch_paths=5000)),
                   (reference_scene("mis_variant"), dict(width=64, height=64, spp=4, eye_depth=8, light_depth=9, seed=1))):
        sc = api.Scene(sd)
        monkeypatch.setenv("B2_MIS_MEGAKERNEL", "1")
        mega, sm = sc.render_bdpt(strategies=api.STRATEGIES_MIS, spl=1, **kw)
        monkeypatch.delenv("B2_MIS_MEGAKERNEL")
        wave, sw = sc.render_bdpt(strategies=api.STRATEGIES_MIS, spl=1, **kw)
        assert sw["n_launches"] > sm["n_launches"]                       # the wavefront really ran
        for k in ("eye_samples", "closest_rays"):
            assert sw[k] == sm[k], k
        for k in ("shadow_rays", "path_samples"):
            assert abs(int(sw[k]) - int(sm[k])) <= 2 + 1e-6 * int(sm[k]), (k, sw[k], sm[k])
        d = np.abs(wave - mega)
        assert (d <= 1e-4 * np.abs(mega) + 1e-6).mean() >= 0.999 and abs(wave.mean() - mega.mean()) <= 1e-5 * mega.mean()
        sc.close()


def test_pt_cli_strategies_flag_selects_the_mis_mode(tmp_path):
    """`pt_cli --mode bdpt --strategies mis` == `render_bdpt(strategies=STRATEGIES_MIS)` on the same scene text and seed."""
    import subprocess
    import __graft_entry__ as g
    sd = box_scene(tmp_path)
    scene_txt = tmp_path / "box.txt"; out = tmp_path / "o.ppm"
    r = subprocess.run([g.CLI, "--input", str(scene_txt), "--output", str(out), "--mode", "bdpt", "--strategies", "mis", "--spp", "16",
                        "--eye-depth", "3", "--light-depth", "4", "--width", "48", "--height", "48"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    b = open(out, "rb").read()
    hdr = b"P6\n48 48\n255\n"
    assert b.startswith(hdr)
    img = np.frombuffer(b[len(hdr):], np.uint8).reshape(48, 48, 3)
    sc = api.Scene(sd)
    film, _ = sc.render_bdpt(width=48, height=48, spp=16, spl=8, eye_depth=3, light_depth=4, seed=1, strategies=api.STRATEGIES_MIS)
    assert (np.abs(img.astype(int) - api.film_to_rgb8(film).astype(int)) <= 1).mean() > 0.995 and img.mean() > 5
    r = subprocess.run([g.CLI, "--input", str(scene_txt), "--strategies", "veach"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 2 and "unknown --strategies" in r.stderr
    sc.close()
