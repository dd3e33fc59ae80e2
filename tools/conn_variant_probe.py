"""BDPT render of a generated OBJ scene of ~n triangles, counters against the host compilation of the same stage bodies:
python tools/conn_variant_probe.py n_tris W H spp [hostsim]   (the non-pooled connection kernel serves scenes below 2048 triangles)"""
import json, sys, tempfile
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from path_tracing_b200 import api, gen
from path_tracing_b200.scene import SceneData

n, W, H, spp = [int(x) for x in sys.argv[1:5]]
with tempfile.TemporaryDirectory() as d:
    sd = SceneData.from_file(gen.write_obj_scene(d, n_tris=n, width=W, height=H, tex_size=64), obj_directive=True)
sc = api.Scene(sd)
film, st = sc.render_bdpt(width=W, height=H, spp=spp, spl=8, eye_depth=8, light_depth=8, seed=1)
out = {"n_tris": int(sc.bvh_info()["n_tris"]), "bvh_depth": int(sc.bvh_info()["depth"]), "film_mean": float(film.mean(dtype="float64"))}
out.update({k: int(st[k]) for k in ("eye_samples", "closest_rays", "shadow_rays", "path_samples", "pairs_tested")})
if len(sys.argv) > 5:
    from hostsim_py import HostSim
    hs = HostSim(sd)
    fh, cnt = hs.render(W, H, spp, 8, 8, 8, seed=1)
    out.update({"hs_" + k: int(cnt[k]) for k in ("closest_rays", "shadow_rays", "path_samples")}, hs_film_mean=float(fh.mean(dtype="float64")))
print(json.dumps(out))
