"""Where the end-to-end step (host scene arrays -> host film) spends its wall time.  Run on a GPU box."""
import json, sys, time
import numpy as np
sys.path.insert(0, ".")
from path_tracing_b200 import api as b2
from path_tracing_b200.scene import reference_scene

data = reference_scene("mis")
import torch
host = torch.empty((1024, 1024, 3), dtype=torch.float32).pin_memory().numpy()      # the caller's film, pinned as in bench.py
out = {}
for it in range(5):
    t0 = time.perf_counter(); s = b2.Scene(data, build=False); t1 = time.perf_counter()
    s.build_bvh(); t2 = time.perf_counter()
    film, st = s.render_bdpt(width=1024, height=1024, spp=64, spl=8, eye_depth=8, light_depth=8, seed=1, out=host); t3 = time.perf_counter()
    film, st2 = s.render_bdpt(width=1024, height=1024, spp=64, spl=8, eye_depth=8, light_depth=8, seed=1, out=host); t4 = time.perf_counter()
    s.close(); t5 = time.perf_counter()
    out[f"iter{it}"] = dict(create_ms=(t1 - t0) * 1e3, build_ms=(t2 - t1) * 1e3, render1_ms=(t3 - t2) * 1e3, render1_dev_ms=st["ms_total"],
                            render2_ms=(t4 - t3) * 1e3, render2_dev_ms=st2["ms_total"], destroy_ms=(t5 - t4) * 1e3)
print(json.dumps(out, indent=1))
