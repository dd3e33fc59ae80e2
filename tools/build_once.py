"""BVH build only (for profiling): python tools/build_once.py [n_tris] [soup|terrain]"""
import json, sys, time
sys.path.insert(0, ".")
from path_tracing_b200 import api, gen
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
kind = sys.argv[2] if len(sys.argv) > 2 else "soup"
sd = gen.soup_scene(n) if kind == "soup" else gen.terrain_scene(n)
t0 = time.perf_counter(); sc = api.Scene(sd, build=False); t1 = time.perf_counter()
sc.build_bvh(); t2 = time.perf_counter()
sc.build_bvh(); t3 = time.perf_counter()
print(json.dumps(dict(n=len(sd.tris), create_s=t1 - t0, build1_s=t2 - t1, build2_s=t3 - t2, info=sc.bvh_info())))
