"""pt / ppm throughput of this repo next to the reference's own kernels (oracle/_ref/libref_cuda.so) on the same GPU.
    python tools/pt_ppm_probe.py [scene] [size] [spp] [depth] [spl]
One JSON line per (implementation, mode).  Baseline evidence only (the oracle directory is test / baseline infrastructure)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from oracle import oracle_py as O                      # noqa: E402
from path_tracing_b200 import api as b2                # noqa: E402
from path_tracing_b200.scene import reference_scene    # noqa: E402


def main():
    a = sys.argv[1:]
    scene = a[0] if len(a) > 0 else "cornell"
    size = int(a[1]) if len(a) > 1 else 512
    spp = int(a[2]) if len(a) > 2 else 64
    depth = int(a[3]) if len(a) > 3 else 8
    spl = int(a[4]) if len(a) > 4 else 1000000
    sd = reference_scene(scene)
    f = sd.to_flat()
    d = O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])
    n = size * size * spp
    if O.have_ref_cuda():
        rc = O.RefCuda()
        h = rc.scene(d)
        for m, mid, s_spp, s_spl in (("pt", rc.PT, spp, 8), ("ppm", rc.PPM, 1, spl)):
            rc.render(h, mid, d.cam, depth, depth, 64, 64, 1, 8)
            img, ms = rc.render(h, mid, d.cam, depth, depth, size, size, s_spp, s_spl)
            unit = n if m == "pt" else len(sd.lights) * spl
            print(json.dumps({"impl": "reference-cuda", "mode": m, "scene": scene, "size": size, "spp": s_spp, "spl": s_spl, "depth": depth, "ms": ms,
                              "msamples_or_mphotons_per_s": unit / ms / 1e3, "mean_rgb": [float(x) for x in img.mean(axis=(0, 1))]}), flush=True)
        rc.destroy(h)
    s = b2.Scene(sd)
    for m in ("pt", "ppm"):
        run = (lambda: s.render_pt(max_depth=depth, width=size, height=size, spp=spp, seed=1)) if m == "pt" else \
              (lambda: s.render_ppm(spl=spl, eye_depth=depth, light_depth=depth, width=size, height=size, seed=1))
        run()
        t0 = time.perf_counter()
        img, st = run()
        wall = (time.perf_counter() - t0) * 1e3
        unit = n if m == "pt" else len(sd.lights) * spl
        print(json.dumps({"impl": "b2", "mode": m, "scene": scene, "size": size, "spp": spp if m == "pt" else 1, "spl": spl, "depth": depth, "ms": wall,
                          "ms_device": st["ms_total"], "msamples_or_mphotons_per_s": unit / wall / 1e3, "closest_rays": st["closest_rays"],
                          "shadow_rays": st["shadow_rays"], "path_samples": st["path_samples"], "hit_points": st["eye_vertices"],
                          "mean_rgb": [float(x) for x in img.mean(axis=(0, 1))]}), flush=True)
    s.close()


if __name__ == "__main__":
    main()
