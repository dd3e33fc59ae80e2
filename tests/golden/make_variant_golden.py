"""Golden film of the documented mis_test.txt variant (path_tracing_b200.scene.reference_scene("mis_variant"), SURVEY.md 8d
C2): the UNMODIFIED reference (oracle/_ref/libref_bdpt.so = cpu_bdpt.cpp) rendered with one thread, so its std::mt19937
streams are reproducible.  Run here, where /root/reference exists:  python tests/golden/make_variant_golden.py
-> tests/golden/film_mis_variant.npz"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import oracle_py as O                      # noqa: E402
from path_tracing_b200.scene import reference_scene    # noqa: E402

O.build()
R = O.Ref()
f = reference_scene("mis_variant").to_flat()
d = O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])
h = R.scene(d)
cfg = dict(W=40, H=40, spp=2, spl=4, eye_depth=4, light_depth=4)
film = R.render(h, d.cam, cfg["eye_depth"], cfg["light_depth"], cfg["W"], cfg["H"], cfg["spp"], cfg["spl"], nthreads=1)
R.destroy(h)
np.savez_compressed(os.path.join(HERE, "film_mis_variant.npz"), film=film,
                    cfg=np.array([cfg["W"], cfg["H"], cfg["spp"], cfg["spl"], cfg["eye_depth"], cfg["light_depth"]], np.int32))
print("film mean", film.mean(axis=(0, 1)), "lit pixels", float((film.sum(axis=2) > 0).mean()))
