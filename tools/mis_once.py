"""One render per strategy set on a reference scene: device time, rays, and the per-pixel noise (RMS difference of two seeds)
-- the numbers behind DESIGN.md section 9's table for B2_STRATEGIES_MIS.  usage: mis_once.py scene W H spp depth"""
import json, sys
import numpy as np
sys.path.insert(0, ".")
from path_tracing_b200 import api as b2
from path_tracing_b200.scene import reference_scene

scene, W, H, spp, depth = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
s = b2.Scene(reference_scene(scene))
for name, strat in (("nee", b2.STRATEGIES_NEE), ("light", b2.STRATEGIES_LIGHT_TRACING), ("full", b2.STRATEGIES_FULL), ("mis", b2.STRATEGIES_MIS)):
    kw = dict(width=W, height=H, spp=spp, spl=1, eye_depth=depth, light_depth=depth + 1, strategies=strat)
    s.render_bdpt(seed=1, **kw)
    a, st = s.render_bdpt(seed=2, **kw)
    b, _ = s.render_bdpt(seed=3, **kw)
    m = 0.5 * (a + b)
    print(json.dumps({"scene": scene, "strategies": name, "W": W, "H": H, "spp": spp, "depth": depth, "ms_device": st["ms_total"],
                      "eye_msamples_per_s": st["eye_samples"] / st["ms_total"] / 1e3, "mrays_per_s": (st["closest_rays"] + st["shadow_rays"]) / st["ms_total"] / 1e3,
                      "closest_rays": st["closest_rays"], "shadow_rays": st["shadow_rays"], "path_samples": st["path_samples"],
                      "film_mean": [float(x) for x in m.mean(axis=(0, 1))], "rms_two_seeds": float(np.sqrt(np.mean((a - b) ** 2) / 2.0)),
                      "rms_x_sqrt_ms": float(np.sqrt(np.mean((a - b) ** 2) / 2.0) * np.sqrt(st["ms_total"]))}), flush=True)
