"""One pt (or ppm) render for profiling / timing: python tools/pt_once.py [scene=cornell|mis|c3] [mode=pt|ppm] [W H spp depth spl]
Prints one JSON line (device time, rays, per-ray traversal counts from a counting pass)."""
import json, sys, tempfile
sys.path.insert(0, ".")
from path_tracing_b200 import api, gen
from path_tracing_b200.scene import SceneData, reference_scene

a = sys.argv[1:]
scene = a[0] if len(a) > 0 else "cornell"
mode = a[1] if len(a) > 1 else "pt"
W, H, spp, depth, spl = [int(x) for x in (a[2:7] + ["512", "512", "64", "8", "1000000"][len(a[2:7]):])]
if scene == "c3":
    with tempfile.TemporaryDirectory() as d:
        sd = SceneData.from_file(gen.write_obj_scene(d, n_tris=1 << 20, width=W, height=H), obj_directive=True)
else:
    sd = reference_scene(scene)
sc = api.Scene(sd)
run = (lambda **kw: sc.render_pt(max_depth=depth, width=W, height=H, spp=spp, seed=1, **kw)) if mode == "pt" else \
      (lambda **kw: sc.render_ppm(spl=spl, eye_depth=depth, light_depth=depth, width=W, height=H, seed=1, **kw))
run()
film, st = run()
_, sn = run(count_traversal=True)
rays = st["closest_rays"] + st["shadow_rays"]
print(json.dumps({"scene": scene, "mode": mode, "W": W, "H": H, "spp": spp, "depth": depth, "spl": spl, "n_tris": len(sd.tris), "ms_device": st["ms_total"],
                  "eye_msamples_per_s": st["eye_samples"] / st["ms_total"] / 1e3, "mrays_per_s": rays / st["ms_total"] / 1e3,
                  "closest_rays": st["closest_rays"], "shadow_rays": st["shadow_rays"], "path_samples": st["path_samples"], "hit_points": st["eye_vertices"],
                  "n_waves": st["n_waves"], "n_launches": st["n_launches"], "nodes_per_ray": sn["nodes_visited"] / max(rays, 1), "tris_per_ray": sn["tris_tested"] / max(rays, 1),
                  "film_mean": [float(x) for x in film.reshape(-1, 3).mean(0)]}))
