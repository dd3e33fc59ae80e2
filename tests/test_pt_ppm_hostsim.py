"""The pt and ppm integrators (path_tracing_b200/csrc/b2_pt.cuh) on the CPU-only box: the device
bodies compiled for the host (tests/hostsim) against oracle/bdpt_oracle.c's restatement of src/pt_cu.cu / src/ppm_cu.cu on
the same Philox streams.  The reference has no CPU implementation of these integrators (parity unpinned on the CPU); the
GPU-side comparison with the reference's own kernels is tests/test_gpu_pt_ppm.py."""
import numpy as np
import pytest
from hostsim_py import HostSim
from oracle import oracle_py as O
from path_tracing_b200.scene import reference_scene


def desc(sd):
    f = sd.to_flat()
    return O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])


def close_films(a, b, frac, tol=1e-3):
    """`frac` of the pixels within `tol` of the film mean (paths that graze an edge may diverge between two fp32 evaluation orders)"""
    scale = max(float(a.mean()), 1e-6)
    ok = (np.abs(a - b).max(axis=2) <= tol * scale).mean()
    return ok >= frac, ok


@pytest.mark.parametrize("name", ["cornell", "mis"])
def test_pt_bodies_match_the_restated_kernel(name):
    sd = reference_scene(name)
    d = desc(sd)
    P = O.Port(); h = P.scene(d)
    hs = HostSim(sd)
    W, H, spp, depth = 40, 32, 4, 6
    ref, st = P.render_pt(h, d.cam, depth, W, H, spp, P.opts(rng_mode=1, seed=3, max_delta=256), want_stats=True)
    img, cnt = hs.render_other(1, W, H, spp=spp, eye_depth=depth, seed=3)
    hs.close(); P.destroy(h)
    assert st.eye_samples == W * H * spp == cnt["eye_samples"]
    assert abs(cnt["closest_rays"] - st.closest_rays) <= 2e-3 * st.closest_rays + 2
    assert abs(cnt["shadow_rays"] - st.shadow_rays) <= 2e-3 * st.shadow_rays + 2
    assert abs(cnt["path_samples"] - st.path_samples) <= 2e-3 * st.path_samples + 2
    assert np.allclose(img.mean(axis=(0, 1)), ref.mean(axis=(0, 1)), rtol=2e-3, atol=1e-6)
    ok, frac = close_films(ref, img, 0.99)
    assert ok, frac


def test_pt_light_ball_radiance_and_depth_zero():
    """A pixel that looks straight at a light ball gets illum / (4 pi r^2) (pt_cu.cu:60-96, cone fraction 1 at depth 0);
    max_depth 0 renders black (the loop of :48 does not run)."""
    sd = reference_scene("mis")
    d = desc(sd)
    P = O.Port(); h = P.scene(d)
    W = H = 64
    ref = P.render_pt(h, d.cam, 4, W, H, 2, P.opts(rng_mode=1, seed=1, max_delta=256))
    L = sd.lights
    peak = max(float(l["illum"].max()) / (4 * np.pi * float(l["ball_r"]) ** 2) for l in L)
    assert ref.max() <= min(peak, 15.0) * 1.0001 and ref.max() > 0
    hs = HostSim(sd)
    img, _ = hs.render_other(1, W, H, spp=2, eye_depth=4, seed=1)
    black, cnt = hs.render_other(1, 16, 16, spp=2, eye_depth=0, seed=1)
    hs.close(); P.destroy(h)
    assert np.allclose(img, ref, rtol=1e-4, atol=1e-6)
    assert not black.any() and cnt["closest_rays"] == 0


@pytest.mark.parametrize("name", ["cornell", "mis"])
def test_ppm_bodies_match_the_restated_kernels(name):
    sd = reference_scene(name)
    d = desc(sd)
    P = O.Port(); h = P.scene(d)
    hs = HostSim(sd)
    W, H, spl = 40, 32, 1500
    ref, st = P.render_ppm(h, d.cam, 8, 6, W, H, spl, 0.0, P.opts(rng_mode=1, seed=5, max_delta=256), want_stats=True)
    img, cnt = hs.render_other(2, W, H, spl=spl, eye_depth=8, light_depth=6, seed=5)
    hs.close(); P.destroy(h)
    assert cnt["shadow_rays"] == st.shadow_rays                  # hit points
    assert abs(cnt["closest_rays"] - st.closest_rays) <= 2e-3 * st.closest_rays + 2
    assert abs(cnt["path_samples"] - st.path_samples) <= 5e-3 * st.path_samples + 2      # photon -> hit point gathers
    assert np.allclose(img.mean(axis=(0, 1)), ref.mean(axis=(0, 1)), rtol=1e-2, atol=1e-6)
    ok, frac = close_films(ref, img, 0.97)
    assert ok, frac


def test_ppm_radius_scales_the_gather_and_zero_photons_leave_the_direct_term():
    sd = reference_scene("cornell")
    hs = HostSim(sd)
    W, H = 32, 24
    a, ca = hs.render_other(2, W, H, spl=800, eye_depth=8, light_depth=4, seed=2, ppm_radius=0.05)
    b, cb = hs.render_other(2, W, H, spl=800, eye_depth=8, light_depth=4, seed=2, ppm_radius=0.1)
    z, cz = hs.render_other(2, W, H, spl=0, eye_depth=8, light_depth=4, seed=2)
    e0, _ = hs.render_other(2, W, H, spl=800, eye_depth=0, light_depth=4, seed=2)
    hs.close()
    assert ca["shadow_rays"] == cb["shadow_rays"] == cz["shadow_rays"] > 0           # same eye pass, same hit points
    assert 2.5 * ca["path_samples"] < cb["path_samples"] < 6 * ca["path_samples"]    # ~4x the disc area
    assert cz["path_samples"] == 0 and z.sum() <= a.sum()
    assert not e0.any()                                                              # no eye pass, no hit points (ppm_cu.cu:98)
