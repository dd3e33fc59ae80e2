"""C3 of BASELINE.json / SURVEY.md 8(d): synthetic OBJ + MTL + PPM textures (~1 M triangles) loaded through the OBJ
adapter, 1920x1080, 256 spp, eye/light depth 16, spl 1 and spl 8.  Prints one JSON object; writes the spl-8 image.
usage: python tools/c3_render.py [--tris N] [--spp S] [--out gpurun_out/c3.png]"""
import argparse, json, sys, tempfile, time
import numpy as np
sys.path.insert(0, ".")
from path_tracing_b200 import api, gen
from path_tracing_b200.scene import SceneData

ap = argparse.ArgumentParser()
ap.add_argument("--tris", type=int, default=1 << 20)
ap.add_argument("--spp", type=int, default=256)
ap.add_argument("--depth", type=int, default=16)
ap.add_argument("--width", type=int, default=1920)
ap.add_argument("--height", type=int, default=1080)
ap.add_argument("--out", default=None)
a = ap.parse_args()
res = {"config": f"C3: generated OBJ scene, {a.width}x{a.height}, {a.spp} spp, depth {a.depth}/{a.depth}"}
with tempfile.TemporaryDirectory() as d:
    t0 = time.perf_counter(); scene_txt = gen.write_obj_scene(d, n_tris=a.tris, width=a.width, height=a.height); t1 = time.perf_counter()
    sd = SceneData.from_file(scene_txt, obj_directive=True); t2 = time.perf_counter()
res.update(generate_s=t1 - t0, parse_s=t2 - t1, n_tris=len(sd.tris), n_materials=len(sd.materials), n_lights=len(sd.lights))
t0 = time.perf_counter(); sc = api.Scene(sd); t1 = time.perf_counter()
res.update(upload_and_build_s=t1 - t0, bvh=sc.bvh_info())
for spl in (1, 8):
    kw = dict(width=a.width, height=a.height, spp=a.spp, spl=spl, eye_depth=a.depth, light_depth=a.depth, seed=1)
    sc.render_bdpt(**dict(kw, spp=1))                                     # warm-up (allocations, first launches)
    t0 = time.perf_counter(); film, st = sc.render_bdpt(**kw); wall = time.perf_counter() - t0
    _, sn = sc.render_bdpt(**dict(kw, spp=max(a.spp // 16, 1)), count_traversal=True)
    rays = st["closest_rays"] + st["shadow_rays"]
    res[f"spl{spl}"] = dict(ms_device=st["ms_total"], wall_s=wall, eye_msamples_per_s=st["eye_samples"] / st["ms_total"] / 1e3,
                            mrays_per_s=rays / st["ms_total"] / 1e3, mpath_samples_per_s=st["path_samples"] / st["ms_total"] / 1e3,
                            closest_rays_per_sample=st["closest_rays"] / st["eye_samples"], shadow_rays_per_sample=st["shadow_rays"] / st["eye_samples"],
                            pairs_per_sample=st["pairs_tested"] / st["eye_samples"], nodes_per_ray=sn["nodes_visited"] / max(sn["closest_rays"] + sn["shadow_rays"], 1),
                            tris_per_ray=sn["tris_tested"] / max(sn["closest_rays"] + sn["shadow_rays"], 1), n_waves=st["n_waves"], n_batches=st["n_batches"],
                            film_mean=[float(x) for x in film.reshape(-1, 3).mean(0)], finite=bool(np.isfinite(film).all()), nonneg=bool((film >= 0).all()))
    if a.out and spl == 8:
        api.write_image(a.out, film)
print(json.dumps(res))
