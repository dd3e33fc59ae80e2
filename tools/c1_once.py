"""One C1 render (BASELINE.json configs[0]: input.txt, 512 x 512, 16 spp, spl 8, depth 8) through the C ABI: python tools/c1_once.py [spp]"""
import json, sys
sys.path.insert(0, ".")
from path_tracing_b200 import api as b2
from path_tracing_b200.scene import reference_scene

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 16
s = b2.Scene(reference_scene("cornell"))
for _ in range(2):
    film, st = s.render_bdpt(width=512, height=512, spp=spp, spl=8, eye_depth=8, light_depth=8, seed=1, profile_stages=True)
out = {k: st[k] for k in ("ms_total", "ms_primary", "ms_extend", "ms_connect", "eye_samples", "path_samples", "closest_rays", "shadow_rays", "pairs_tested", "n_launches", "n_waves")}
out.update(eye_msamples_per_s=st["eye_samples"] / st["ms_total"] / 1e3, mpath_samples_per_s=st["path_samples"] / st["ms_total"] / 1e3,
           mrays_per_s=(st["closest_rays"] + st["shadow_rays"]) / st["ms_total"] / 1e3, film_mean=float(film.mean()))
print(json.dumps(out))
