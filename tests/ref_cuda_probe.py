"""Run the reference's own CUDA integrators (oracle/_ref/libref_cuda.so: bdpt_cu.cu / pt_cu.cu / ppm_cu.cu compiled
unmodified for sm_100a) on this GPU beside this repo's path and print one JSON line per configuration.

    python tests/ref_cuda_probe.py [scene] [size] [spp] [spl] [depth] [modes] [ref-only]     e.g.  mis 256 4 8 8 bdpt,pt

Baseline evidence only (SURVEY.md 2b "(ii) honesty number"): the CUDA BDPT is not the parity oracle.
Lives under tests/ because only tests/, smoke() and bench.py's baseline legs may execute oracle/."""
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
    scene = a[0] if len(a) > 0 else "mis"
    size = int(a[1]) if len(a) > 1 else 256
    spp = int(a[2]) if len(a) > 2 else 4
    spl = int(a[3]) if len(a) > 3 else 8
    depth = int(a[4]) if len(a) > 4 else 8
    modes = (a[5] if len(a) > 5 else "bdpt").split(",")
    sd = reference_scene(scene)
    f = sd.to_flat()
    d = O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])
    rc = O.RefCuda()
    h = rc.scene(d)
    n = size * size * spp
    mode_id = {"bdpt": rc.BDPT, "pt": rc.PT, "ppm": rc.PPM}
    for m in modes:
        rc.render(h, mode_id[m], d.cam, depth, depth, min(size, 64), min(size, 64), 1, spl)        # context + module load
        img, ms = rc.render(h, mode_id[m], d.cam, depth, depth, size, size, spp, spl)
        print(json.dumps({"impl": "reference-cuda", "mode": m, "scene": scene, "size": size, "spp": spp, "spl": spl, "depth": depth,
                          "ms": ms, "eye_msamples_per_s": n / ms / 1e3, "mean_rgb": [float(x) for x in img.mean(axis=(0, 1))],
                          "finite": bool(np.isfinite(img).all())}), flush=True)
    if len(a) > 6 and a[6] == "ref-only":
        rc.destroy(h)
        return
    s = b2.Scene(sd)
    kw = dict(width=size, height=size, spp=spp, spl=spl, eye_depth=depth, light_depth=depth, seed=1)
    s.render_bdpt(**kw)
    t0 = time.perf_counter()
    img, st = s.render_bdpt(**kw)
    wall = (time.perf_counter() - t0) * 1e3
    print(json.dumps({"impl": "b2", "mode": "bdpt", "scene": scene, "size": size, "spp": spp, "spl": spl, "depth": depth,
                      "ms": wall, "ms_device": st["ms_total"], "eye_msamples_per_s": n / wall / 1e3,
                      "mean_rgb": [float(x) for x in img.mean(axis=(0, 1))]}), flush=True)
    s.close()
    rc.destroy(h)


if __name__ == "__main__":
    main()
