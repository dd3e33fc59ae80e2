"""One small render for ncu captures: python tools/render_small.py scene size spp [spl depth]"""
import json, sys
sys.path.insert(0, ".")
from path_tracing_b200 import api as b2
from path_tracing_b200.scene import reference_scene
scene, size, spp = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
spl = int(sys.argv[4]) if len(sys.argv) > 4 else 8
depth = int(sys.argv[5]) if len(sys.argv) > 5 else 8
s = b2.Scene(reference_scene(scene))
film, st = s.render_bdpt(width=size, height=size, spp=spp, spl=spl, eye_depth=depth, light_depth=depth, seed=1, profile_stages=True)
print(json.dumps({k: st[k] for k in st if k.startswith("ms_") or k in ("eye_samples", "path_samples", "shadow_rays", "pairs_tested", "n_waves")}))
