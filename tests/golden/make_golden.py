"""Generates the committed fixtures from the reference ITSELF (run in the build container, where /root/reference
exists; the GPU box only reads the outputs):

    path_tracing_b200/data/reference_scenes.npz   the two scene files of the reference, parsed (flat arrays)
    tests/golden/hits_<scene>.npz                 rays -> (obj id, t) of cpu_find_closest_hit, segments -> cpu_check_visibility
    tests/golden/bsdf_tables.npz                  bsdf_evaluate / bsdf_pdf / bsdf_sample on a grid of inputs
    tests/golden/mis_tables.npz                   cpu_calculate_mis_weight on random vertex chains
    tests/golden/films.npz                        single-thread run_cpu_bdpt films (std::mt19937 streams)

All outputs come from oracle/_ref/libref_bdpt.so, i.e. the UNMODIFIED reference sources compiled by oracle/Makefile.
Usage: python tests/golden/make_golden.py
"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle_py as O  # noqa: E402

REF_DIR = "/root/reference"
GOLD = os.path.join(ROOT, "tests", "golden")
DATA = os.path.join(ROOT, "path_tracing_b200", "data")
SCENES = {"cornell": "input.txt", "mis": "mis_test.txt"}


def unit(v):
    return (v / np.linalg.norm(v, axis=-1, keepdims=True)).astype(np.float32)


def scene_rays(d, rng, n):
    """Rays that exercise the scene: camera rays, rays from random interior points, rays aimed at triangle
    vertices / edge midpoints (grazing cases), rays starting on surfaces."""
    tri = d.geom[d.kind == 1].reshape(-1, 3, 3)
    lo, hi = tri.reshape(-1, 3).min(0), tri.reshape(-1, 3).max(0)
    eye = d.cam[0:3]
    org, dr = [], []
    # camera-like rays
    tgt = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    org.append(np.tile(eye, (n, 1))); dr.append(unit(tgt - eye))
    # interior random
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    org.append(o); dr.append(unit(rng.normal(size=(n, 3))))
    # aimed at vertices and edge midpoints (exact float targets -> edge/vertex grazing, tie t between neighbours)
    k = rng.integers(0, len(tri), n)
    a, b = rng.integers(0, 3, n), rng.integers(0, 3, n)
    tg = 0.5 * (tri[k, a] + tri[k, b])
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    org.append(o); dr.append(unit(tg - o))
    # starting exactly on a triangle (self-intersection epsilon handling)
    w = rng.dirichlet([1, 1, 1], n).astype(np.float32)
    o = (tri[k] * w[:, :, None]).sum(1).astype(np.float32)
    org.append(o); dr.append(unit(rng.normal(size=(n, 3))))
    # towards the light balls
    if d.n_lights:
        li = rng.integers(0, d.n_lights, n)
        o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
        tg = d.lights[li, 0:3] + rng.normal(size=(n, 3)).astype(np.float32) * d.lights[li, 11:12]
        org.append(o); dr.append(unit(tg - o))
    return np.concatenate(org).astype(np.float32), np.concatenate(dr).astype(np.float32)


def main():
    O.build()
    P, R = O.Port(), O.Ref()
    os.makedirs(GOLD, exist_ok=True); os.makedirs(DATA, exist_ok=True)
    rng = np.random.default_rng(20261104)
    scenes = {}
    films = {}
    for name, fn in SCENES.items():
        d = P.parse(os.path.join(REF_DIR, fn))
        for k in ("kind", "group", "geom", "mtl", "lights", "cam"):
            scenes[f"{name}_{k}"] = getattr(d, k)
        scenes[f"{name}_res"] = np.array([d.W, d.H], np.int32)
        h = R.scene(d)
        org, dr = scene_rays(d, rng, 4000)
        hit = R.closest_hit(h, org, dr)
        assert hit["bad"] == 0
        p1 = org.copy()
        p2 = (org + dr * rng.uniform(0.05, 4.0, (len(org), 1)).astype(np.float32)).astype(np.float32)
        vis = R.visibility(h, p1, p2)
        np.savez_compressed(os.path.join(GOLD, f"hits_{name}.npz"), org=org, dir=dr, id=hit["id"], t=hit["t"], pos=hit["pos"],
                            normal=hit["normal"], mtl=hit["mtl"], p1=p1, p2=p2, vis=vis)
        print(name, "rays", len(org), "hit fraction", float((hit["id"] != -1).mean()), "visible fraction", float(vis.mean()))
        # single-thread films (OMP_NUM_THREADS=1 semantics): the reference's own mt19937 streams, reproducible
        cfg = dict(W=40, H=40, spp=2, spl=4, eye_depth=4, light_depth=4)
        films[f"{name}_film"] = R.render(h, d.cam, cfg["eye_depth"], cfg["light_depth"], cfg["W"], cfg["H"], cfg["spp"], cfg["spl"], nthreads=1)
        films[f"{name}_cfg"] = np.array([cfg["W"], cfg["H"], cfg["spp"], cfg["spl"], cfg["eye_depth"], cfg["light_depth"]], np.int32)
        R.destroy(h)
    np.savez_compressed(os.path.join(DATA, "reference_scenes.npz"), **scenes)
    np.savez_compressed(os.path.join(GOLD, "films.npz"), **films)

    # ---- BSDF tables ----
    n = 6000
    mats = np.array([[0.2, 0.2, 0.2, 0.3, 0.0, 0.0], [1, 0, 0, 1.0, 0.0, 0.0], [1, 1, 1, 0.0, 1.0, 0.0], [0.9, 0.6, 0.2, 0.25, 1.0, 0.0],
                     [1, 1, 1, 0.0, 0.0, 1.5], [0.8, 0.9, 1.0, 0.0, 0.0, 2.42], [1, 1, 1, 0.0, 0.9, 0.0], [0.5, 0.5, 0.9, 0.6, 0.5, 0.0],
                     [0.7, 0.7, 0.7, 0.2, 0.0, 1.3], [5, 5, 5, 1.0, 0.0, 0.0]], np.float32)
    mtl = mats[rng.integers(0, len(mats), n)]
    N = unit(rng.normal(size=(n, 3)))
    N[:200] = np.array([0, 0, 1], np.float32); N[200:400] = np.array([0, 0, -1], np.float32); N[400:500] = np.array([0, 1, 0], np.float32)
    wo = unit(rng.normal(size=(n, 3))); wi = unit(rng.normal(size=(n, 3)))
    flip = (np.sum(wo * N, 1) < 0) & (rng.uniform(size=n) < 0.7)
    wo[flip] *= -1
    flip = (np.sum(wi * N, 1) < 0) & (rng.uniform(size=n) < 0.6)
    wi[flip] *= -1
    # near-specular pairs (wi ~ reflect(wo)) so the GGX lobe of smooth materials is exercised
    refl = (2 * np.sum(wo * N, 1, keepdims=True) * N - wo).astype(np.float32)
    near = rng.uniform(size=n) < 0.3
    wi[near] = unit(refl[near] + rng.normal(size=(near.sum(), 3)).astype(np.float32) * 1e-3)
    u = rng.uniform(size=(n, 3)).astype(np.float32)
    cur_eta = np.where(rng.uniform(size=n) < 0.7, 1.0, mtl[:, 5]).astype(np.float32)
    cur_eta[cur_eta <= 0] = 1.0
    smp = R.bsdf_sample(mtl, wo, N, u, cur_eta)
    np.savez_compressed(os.path.join(GOLD, "bsdf_tables.npz"), mtl=mtl, wo=wo, wi=wi, N=N, u=u, cur_eta=cur_eta,
                        evaluate=R.bsdf_evaluate(mtl, wo, wi, N), pdf=R.bsdf_pdf(mtl, wo, wi, N),
                        s_wi=smp["wi"], s_f=smp["f"], s_pdf=smp["pdf"], s_delta=smp["is_delta"], s_eta=smp["new_eta"])

    # ---- MIS tables: random vertex chains of up to 6 vertices ----
    m, V = 3000, 6
    eye = np.zeros((m, V, 17), np.float32); lit = np.zeros((m, V, 20), np.float32)
    eye[..., 0:3] = rng.uniform(-1, 1, (m, V, 3)); lit[..., 0:3] = rng.uniform(-1, 1, (m, V, 3))
    eye[..., 3:6] = unit(rng.normal(size=(m, V, 3))); lit[..., 3:6] = unit(rng.normal(size=(m, V, 3)))
    eye[..., 6:9] = rng.uniform(0, 1, (m, V, 3)); lit[..., 6:9] = rng.uniform(0, 1, (m, V, 3))
    eye[..., 9:15] = mats[rng.integers(0, len(mats), (m, V))]; lit[..., 9:15] = mats[rng.integers(0, len(mats), (m, V))]
    eye[..., 15] = rng.uniform(0, 3, (m, V)) * (rng.uniform(size=(m, V)) < 0.8); eye[..., 16] = rng.uniform(0, 3, (m, V))
    lit[..., 15] = (rng.uniform(size=(m, V)) < 0.15); lit[:, 0, 15] = 1.0
    lit[..., 16] = (rng.uniform(size=(m, V)) < 0.1)
    lit[..., 17] = rng.uniform(0, 1, (m, V)); lit[..., 18] = rng.uniform(0, 3, (m, V)); lit[..., 19] = rng.uniform(0, 3, (m, V))
    s_idx = rng.integers(0, V, m).astype(np.int32); t_idx = rng.integers(0, V, m).astype(np.int32)
    cam_pos = rng.uniform(-2, 2, (m, 3)).astype(np.float32)
    d_vec = np.zeros((m, 3), np.float32); dist2 = np.zeros(m, np.float32); w = np.zeros(m, np.float32)
    for i in range(m):
        dv = lit[i, t_idx[i], 0:3] - eye[i, s_idx[i], 0:3]
        if rng.uniform() < 0.6:   # orient the normals so the connection is front-facing on both ends
            if np.dot(eye[i, s_idx[i], 3:6], dv) < 0: eye[i, s_idx[i], 3:6] *= -1
            if np.dot(lit[i, t_idx[i], 3:6], -dv) < 0: lit[i, t_idx[i], 3:6] *= -1
        d_vec[i] = dv; dist2[i] = np.float32(np.dot(dv, dv))
        w[i] = R.mis_weight(eye[i], int(s_idx[i]), lit[i], int(t_idx[i]), d_vec[i], float(dist2[i]), cam_pos[i])
    np.savez_compressed(os.path.join(GOLD, "mis_tables.npz"), eye=eye, lit=lit, s_idx=s_idx, t_idx=t_idx, d_vec=d_vec, dist2=dist2,
                        cam_pos=cam_pos, w=w)
    print("mis weights: nonzero fraction", float((w > 0).mean()))


if __name__ == "__main__":
    main()
