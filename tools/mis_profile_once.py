"""One B2_STRATEGIES_MIS render for an ncu capture.  usage: mis_profile_once.py scene W H spp depth"""
import sys
sys.path.insert(0, ".")
from path_tracing_b200 import api as b2
from path_tracing_b200.scene import reference_scene
scene, W, H, spp, depth = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
s = b2.Scene(reference_scene(scene))
_, st = s.render_bdpt(width=W, height=H, spp=spp, spl=1, eye_depth=depth, light_depth=depth + 1, strategies=b2.STRATEGIES_MIS, seed=2)
print(st["ms_total"], st["closest_rays"], st["shadow_rays"])
