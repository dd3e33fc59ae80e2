"""B2_STRATEGIES_MIS: the wavefront (b2_mis_wave.cu, default) against the one-thread-per-sample kernel (B2_MIS_MEGAKERNEL=1) on the
same Philox streams: films must agree up to the order of the film atomics, counters exactly.  usage: mis_wave_ab.py [scene W H spp depth]"""
import json, os, sys
import numpy as np
sys.path.insert(0, ".")
from path_tracing_b200 import api as b2
from path_tracing_b200.scene import reference_scene

cases = [tuple(sys.argv[1:6])] if len(sys.argv) > 5 else [("cornell", 256, 256, 16, 8), ("mis_variant", 512, 512, 64, 8), ("cornell", 512, 512, 64, 8)]
for scene, W, H, spp, depth in cases:
    W, H, spp, depth = int(W), int(H), int(spp), int(depth)
    s = b2.Scene(reference_scene(scene))
    kw = dict(width=W, height=H, spp=spp, spl=1, eye_depth=depth, light_depth=depth + 1, strategies=b2.STRATEGIES_MIS, seed=2)
    out = {}
    for name, env in (("wave", "0"), ("mega", "1")):
        os.environ["B2_MIS_MEGAKERNEL"] = env
        s.render_bdpt(**kw)
        film, st = s.render_bdpt(**kw)
        out[name] = (film, st)
    fw, sw = out["wave"]; fm, sm = out["mega"]
    d = np.abs(fw - fm)
    print(json.dumps({"scene": scene, "W": W, "H": H, "spp": spp, "depth": depth, "ms_wave": sw["ms_total"], "ms_mega": sm["ms_total"],
                      "speedup": sm["ms_total"] / sw["ms_total"],
                      "counters_wave": [int(sw[k]) for k in ("closest_rays", "shadow_rays", "path_samples")],
                      "counters_mega": [int(sm[k]) for k in ("closest_rays", "shadow_rays", "path_samples")],
                      "n_launches": int(sw["n_launches"]), "n_waves": int(sw["n_waves"]),
                      "film_mean": [float(fw.mean()), float(fm.mean())], "max_abs_diff": float(d.max()), "max_film": float(fm.max()),
                      "frac_within_1e-4_rel": float((d <= 1e-4 * np.abs(fm) + 1e-6).mean())}), flush=True)
    s.close()
os.environ.pop("B2_MIS_MEGAKERNEL", None)
