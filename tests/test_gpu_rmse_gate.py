"""The RMSE gate of SURVEY.md 8(d): the GPU film against INDEPENDENT CPU-oracle films (different random streams, unlike
test_render_matches_oracle_on_identical_streams), with the tolerance calibrated from CPU-vs-CPU reruns.

  A_1..A_4  four CPU films (oracle port, Philox keys 11..14; the port is pinned bit for bit to the compiled reference in
            tests/test_oracle_pin.py), G = GPU film (key 99), all at the same (W, H, spp, spl, depths)
  R_1, R_2  two films of the UNMODIFIED reference itself (oracle/_ref = cpu_bdpt.cpp, all host threads; where it is built) join
            the CPU family: floor and sigma are taken over A_1..A_4, R_1, R_2 and G has to pass against every one of them, the
            reference's own mt19937 films included.  (Measured: R_1 / R_2 tested AGAINST a floor from the four A_i alone fail it
            -- raw relRMSE 3.9 vs 1.2-2.7, a channel mean 3.5 sigma off: at 64 x 64 x 16 spp the statistics are heavy-tailed,
            and the reference draws its one-shot light pool from seed 1337 + (the thread that grabs the chunk), so its runs
            often share a pool the A_i do not have.  A family without the reference would be too narrow, not the GPU wrong.)
  relRMSE_c(X, Y) = sqrt(mean((X_c - Y_c)^2)) / mean((X_c + Y_c) / 2)          per channel c
  gate 1  relRMSE_c(G, A_1) <= 1.25 * max over CPU pairs of relRMSE_c(A_i, A_j)                (raw HDR)
  gate 2  the same after an 8 x 8 box filter, <= max(1.25 * floor, 0.05)
  gate 3  channel means of G within 3 sigma of the CPU run-to-run spread (sigma from the 4 CPU films; the light pool
          is drawn once per render, so this spread does not fall with spp -- SURVEY.md hard part 2)
  gate 4  (converged check) with spl = 256 the pool noise is gone: filtered relRMSE(G, A) <= max(1.25 * floor, 0.05) with a
          floor of 0.06-0.07, and the film means within 2 %"""
import itertools
import os

import numpy as np
import pytest

from oracle import oracle_py as O
from path_tracing_b200 import api
from path_tracing_b200.scene import reference_scene

pytestmark = pytest.mark.gpu


def desc(sd):
    f = sd.to_flat()
    return O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])


def rel_rmse(x, y):
    return np.sqrt(((x - y) ** 2).mean(axis=(0, 1))) / np.maximum((0.5 * (x + y)).mean(axis=(0, 1)), 1e-12)


def box8(x):
    H, W, _ = x.shape
    return x.reshape(H // 8, 8, W // 8, 8, 3).mean(axis=(1, 3))


@pytest.mark.parametrize("name,depth", [("cornell", 4), ("mis_variant", 4)])
def test_rmse_gate_against_independent_cpu_films(name, depth):
    sd = reference_scene(name)
    P = O.Port(); d = desc(sd); h = P.scene(d)
    W = H = 64; spp, spl = 16, 8
    A = [P.render(h, d.cam, depth, depth, W, H, spp, spl, P.opts(rng_mode=P.PHILOX, seed=k, max_delta=256)) for k in (11, 12, 13, 14)]
    sc = api.Scene(sd)
    G, _ = sc.render_bdpt(width=W, height=H, spp=spp, spl=spl, eye_depth=depth, light_depth=depth, seed=99)
    if O.have_ref():
        R = O.Ref(); hr = R.scene(d)
        A += [R.render(hr, d.cam, depth, depth, W, H, spp, spl, nthreads=os.cpu_count() or 1) for _ in range(2)]
        R.destroy(hr)
    n = len(A)
    pairs = list(itertools.combinations(range(n), 2))
    floor_raw = np.max([rel_rmse(A[i], A[j]) for i, j in pairs], axis=0)
    floor_box = np.max([rel_rmse(box8(A[i]), box8(A[j])) for i, j in pairs], axis=0)
    g_raw = np.max([rel_rmse(G, a) for a in A], axis=0)
    g_box = np.max([rel_rmse(box8(G), box8(a)) for a in A], axis=0)
    assert np.all(g_raw <= 1.25 * floor_raw), (g_raw, floor_raw)                              # gate 1
    assert np.all(g_box <= np.maximum(1.25 * floor_box, 0.05)), (g_box, floor_box)            # gate 2
    means = np.array([a.mean(axis=(0, 1)) for a in A])
    mu, sigma = means.mean(axis=0), means.std(axis=0, ddof=1)
    assert np.all(np.abs(G.mean(axis=(0, 1)) - mu) <= 3.0 * sigma * np.sqrt(1.0 + 1.0 / n) + 1e-3 * mu), (G.mean(axis=(0, 1)), mu, sigma)   # gate 3
    P.destroy(h); sc.close()


def test_converged_check_with_a_large_light_pool():
    sd = reference_scene("cornell")
    P = O.Port(); d = desc(sd); h = P.scene(d)
    W = H = 32; spp, spl, depth = 16, 256, 3
    A1, A2 = (P.render(h, d.cam, depth, depth, W, H, spp, spl, P.opts(rng_mode=P.PHILOX, seed=k, max_delta=256)) for k in (21, 22))
    sc = api.Scene(sd)
    G, _ = sc.render_bdpt(width=W, height=H, spp=spp, spl=spl, eye_depth=depth, light_depth=depth, seed=77)
    floor = rel_rmse(box8(A1), box8(A2))                       # 0.06-0.07 at this size: eye-path noise, the pool's is gone
    assert np.all(floor < 0.12), floor
    for a in (A1, A2):
        assert np.all(rel_rmse(box8(G), box8(a)) <= np.maximum(1.25 * floor, 0.05)), (rel_rmse(box8(G), box8(a)), floor)   # gate 4
    assert np.all(np.abs(G.mean(axis=(0, 1)) / (0.5 * (A1 + A2)).mean(axis=(0, 1)) - 1.0) < 0.02)
    P.destroy(h); sc.close()
