"""B2_STRATEGIES_MIS (csrc/b2_mis.cuh: one light subpath per eye sample, every s/t strategy, balance heuristic over real
area densities -- SURVEY 8(f) rank 2) on the CPU: the device body compiled for the host (tests/hostsim, TEST-ONLY).
The reference has no such mode and therefore no oracle; the checks are the ones the theory offers:
  * weights of all strategies of a path sum to one  <=>  the image has the expectation of ANY single complete strategy:
    next-event estimation (B2_STRATEGIES_NEE) on scenes without specular surfaces, light tracing on the pixels that see
    no specular surface when there is a mirror (next-event estimation cannot sample mirror caustics);
  * with light_depth = 1 there is no light subpath and the mode IS next-event estimation, sample for sample."""
import numpy as np
import pytest

from hostsim_py import HostSim
from path_tracing_b200.scene import SceneData
from test_gpu_strategies import box_scene

NEE, LT, MIS = 3, 2, 4


@pytest.fixture(scope="module")
def box(tmp_path_factory):
    tmp = tmp_path_factory.mktemp("mis")
    sd = box_scene(tmp)
    txt = (tmp / "box.txt").read_text()
    assert "M 0.8 0.7 0.4 0.35 0.5 0" in txt
    (tmp / "mirror.txt").write_text(txt.replace("M 0.8 0.7 0.4 0.35 0.5 0", "M 0.9 0.9 0.9 0 1 0"))      # the tall block becomes a perfect mirror
    hs, hm = HostSim(sd), HostSim(SceneData.from_file(str(tmp / "mirror.txt")))
    yield hs, hm
    hs.close(); hm.close()


def test_without_light_subpaths_the_mode_is_next_event_estimation(box):
    hs, _ = box
    a, sa = hs.render_strategies(MIS, 16, 16, spp=8, eye_depth=3, light_depth=1, seed=5)
    b, sb = hs.render_strategies(NEE, 16, 16, spp=8, eye_depth=3, light_depth=1, seed=5)
    assert b.mean() > 0.05 and np.allclose(a, b, rtol=1e-4, atol=1e-6)
    assert sa["shadow_rays"] == sb["shadow_rays"] and sa["path_samples"] == sb["path_samples"]


@pytest.mark.parametrize("eye_depth,light_depth", [(1, 2), (3, 4), (4, 2), (3, 9)])
def test_same_expectation_as_next_event_estimation(box, eye_depth, light_depth):
    hs, _ = box
    W = H = 24
    nee, _ = hs.render_strategies(NEE, W, H, spp=384, eye_depth=eye_depth, light_depth=light_depth, seed=11)
    mis, st = hs.render_strategies(MIS, W, H, spp=192, eye_depth=eye_depth, light_depth=light_depth, seed=12)
    assert np.isfinite(mis).all() and (mis >= 0).all() and st["path_samples"] > W * H * 192 // 2
    tn, tm = nee.sum(axis=(0, 1)), mis.sum(axis=(0, 1))
    assert np.allclose(tm, tn, rtol=0.012), (tm / tn)                  # measured 0.998 - 1.002 at 2048 spp; noise of this size 0.3 %
    bn, bm = (x.reshape(6, 4, 6, 4, 3).mean(axis=(1, 3)) for x in (nee, mis))
    rel = np.abs(bm - bn).mean(axis=2) / np.maximum(bn.mean(axis=2), 0.25 * bn.mean())
    assert np.median(rel) < 0.03 and rel.max() < 0.15, (np.median(rel), rel.max())


def test_mirror_caustics_match_light_tracing_where_no_mirror_is_seen(box):
    _, hm = box
    W = H = 24
    kw = dict(eye_depth=3, light_depth=4)
    seen, _ = hm.render_strategies(LT, W, H, spp=128, seed=1, **kw)
    m = seen.sum(axis=2) > 0                                           # light tracing never splats from a delta surface
    m2 = m.copy()                                                      # drop the pixels next to the mirror: partly covered by it
    m2[1:, :] &= m[:-1, :]; m2[:-1, :] &= m[1:, :]; m2[:, 1:] &= m[:, :-1]; m2[:, :-1] &= m[:, 1:]
    assert 0.7 < m2.mean() < 0.98
    lt, _ = hm.render_strategies(LT, W, H, spp=512, seed=2, **kw)
    mis, _ = hm.render_strategies(MIS, W, H, spp=256, seed=3, **kw)
    nee, _ = hm.render_strategies(NEE, W, H, spp=256, seed=4, **kw)
    assert np.allclose(mis[m2].sum(axis=0), lt[m2].sum(axis=0), rtol=0.015), mis[m2].sum(axis=0) / lt[m2].sum(axis=0)
    assert (nee[m2].sum(axis=0) < 0.9 * lt[m2].sum(axis=0)).all()      # the caustic next-event estimation cannot sample is > 10 % of the light
    # on the mirror itself only eye subpaths work: next-event estimation after the reflection, plus the caustic seen in the mirror
    assert (mis[~m].sum(axis=0) >= 0.98 * nee[~m].sum(axis=0)).all() and nee[~m].sum() > 0
