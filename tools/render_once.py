"""One C2 render (BASELINE.json configs[1]) through the C ABI: the command the ncu captures under profiles/ are taken from."""
import json, sys
sys.path.insert(0, ".")
from path_tracing_b200 import api as b2
from path_tracing_b200.scene import reference_scene

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
scene = sys.argv[2] if len(sys.argv) > 2 else "mis"          # "mis_variant": the documented variant, "cornell": input.txt
s = b2.Scene(reference_scene(scene))
for _ in range(2):
    film, st = s.render_bdpt(width=1024, height=1024, spp=spp, spl=8, eye_depth=8, light_depth=8, seed=1)
out = {k: st[k] for k in ("ms_total", "eye_samples", "path_samples", "closest_rays", "shadow_rays", "pairs_tested", "n_launches", "n_waves", "n_batches")}
out.update(scene=scene, eye_msamples_per_s=st["eye_samples"] / st["ms_total"] / 1e3, mpath_samples_per_s=st["path_samples"] / st["ms_total"] / 1e3,
           mrays_per_s=(st["closest_rays"] + st["shadow_rays"]) / st["ms_total"] / 1e3, film_mean=float(film.mean()))
print(json.dumps(out))
