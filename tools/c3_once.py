"""One reduced C3 render (generated OBJ scene, 1 M triangles) for profiling: python tools/c3_once.py [W H spp spl depth]"""
import json, sys, tempfile
sys.path.insert(0, ".")
from path_tracing_b200 import api, gen
from path_tracing_b200.scene import SceneData

W, H, spp, spl, depth = [int(x) for x in (sys.argv[1:6] + ["480", "270", "16", "8", "16"][len(sys.argv) - 1:])]
with tempfile.TemporaryDirectory() as d:
    sd = SceneData.from_file(gen.write_obj_scene(d, n_tris=1 << 20, width=W, height=H), obj_directive=True)
sc = api.Scene(sd)
film, st = sc.render_bdpt(width=W, height=H, spp=spp, spl=spl, eye_depth=depth, light_depth=depth, seed=1, profile_stages=True)
print(json.dumps({k: st[k] for k in ("ms_total", "ms_primary", "ms_extend", "ms_connect", "eye_samples", "shadow_rays", "path_samples", "pairs_tested", "eye_vertices", "n_waves", "n_launches")} | {"film_mean": float(film.mean(dtype="float64"))}))
