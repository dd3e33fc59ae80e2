"""Where the end-to-end step of bench.py goes beyond the device render: wall times of b2_scene_create, b2_scene_build_bvh, the first
render of a fresh scene (workspace allocation inside), a second render, film read-back and b2_scene_destroy on the C3 scene.
usage: python tools/e2e_breakdown.py [spp=4]"""
import json, sys, time
sys.path.insert(0, ".")
import numpy as np
import torch
import bench
from path_tracing_b200 import api

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 4
sd = bench.c3_scene()
film = torch.empty((1080, 1920, 3), dtype=torch.float32).pin_memory().numpy()
kw = dict(width=1920, height=1080, spp=spp, spl=8, eye_depth=16, light_depth=16, seed=1)
for rep in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    sc = api.Scene(sd); torch.cuda.synchronize(); t1 = time.perf_counter()
    _, st1 = sc.render_bdpt(**kw); torch.cuda.synchronize(); t2 = time.perf_counter()
    _, st2 = sc.render_bdpt(**kw); torch.cuda.synchronize(); t3 = time.perf_counter()
    sc.close(); torch.cuda.synchronize(); t4 = time.perf_counter()
    print(json.dumps({"rep": rep, "create_plus_build_ms": 1e3 * (t1 - t0), "bvh_build_device_ms": sc_info if (sc_info := None) else None,
                      "first_render_wall_ms": 1e3 * (t2 - t1), "first_render_device_ms": st1["ms_total"],
                      "second_render_wall_ms": 1e3 * (t3 - t2), "second_render_device_ms": st2["ms_total"], "destroy_ms": 1e3 * (t4 - t3)}), flush=True)
