"""One C2 render (BASELINE.json configs[1]) through the C ABI: the command the ncu captures under profiles/ are taken from."""
import json, sys
sys.path.insert(0, ".")
from path_tracing_b200 import api as b2
from path_tracing_b200.scene import reference_scene

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
s = b2.Scene(reference_scene("mis"))
for _ in range(2):
    film, st = s.render_bdpt(width=1024, height=1024, spp=spp, spl=8, eye_depth=8, light_depth=8, seed=1)
print(json.dumps({k: st[k] for k in ("ms_total", "eye_samples", "n_launches", "n_waves", "n_batches")}), float(film.mean()))
