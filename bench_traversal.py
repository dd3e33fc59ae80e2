#!/usr/bin/env python
"""C5 of BASELINE.json: traversal microbench.  N incoherent random rays against an M-triangle BVH (uniform triangle soup,
SURVEY.md 8d), closest-hit and any-hit, rays resident in HBM (torch tensors), timed with CUDA events on the launching
stream.  Prints one JSON line: Mrays/s, measured nodes / triangles per ray, the roofline legs (algorithmic bytes
rays x (24 B ray + 8 B result + nodes x 80 B + tris x 48 B) against the measured HBM / L2 bandwidth, 160 / 51 flop per node /
triangle against the in-run FP32 peak; bound = the slower applicable leg), and the bit-exactness check of hit ids against
the exhaustive kernel and the CPU intersector on a subset.  bench.py calls run() for its `extra.C5` sub-record."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402


def run(tris=10_000_000, rays=100_000_000, steps=3, check=2000, cpu_check=50, pk=None, device=0):
    import torch
    from path_tracing_b200 import api, gen
    import bench
    dev = torch.device("cuda", device)
    if pk is None:
        pk = bench.Peaks(device)
    t0 = time.perf_counter()
    sd = gen.soup_scene(tris)
    t_gen = time.perf_counter() - t0
    t0 = time.perf_counter()
    sc = api.Scene(sd, device=device)
    t_build_wall = time.perf_counter() - t0
    info = sc.bvh_info()
    g = torch.Generator(device=dev); g.manual_seed(7)
    n = rays
    org = torch.rand((n, 3), device=dev, generator=g)
    d = torch.randn((n, 3), device=dev, generator=g); d = d / d.norm(dim=1, keepdim=True)
    p2 = torch.rand((n, 3), device=dev, generator=g)
    ids = torch.empty(n, dtype=torch.int32, device=dev); t = torch.empty(n, dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    out = {"metric": "traversal_mrays_per_s", "unit": "Mrays/s",
           "workload": f"C5: {n} incoherent random rays (origins uniform in the box, directions uniform on the sphere / segment ends uniform in the box) vs "
                       f"{info['n_tris']}-triangle soup BVH",
           "bvh": dict({k: info[k] for k in ("n_tris", "n_nodes", "node_bytes", "tri_bytes", "depth", "build_ms", "build_passes", "sah_cost")},
                       builder="lbvh" if os.environ.get("B2_BVH_BUILDER", "0") == "1" else "ploc"),
           "scene_gen_s": t_gen, "scene_create_plus_build_wall_s": t_build_wall}
    # "closest" / "any": the library's default = rays walked in (origin cell, direction octant) order, results in the caller's
    # order, keys + sort inside the timed call; "*_caller_order": B2_KEEP_RAY_ORDER, the incoherent stream walked as it comes
    for name, mode, second in (("closest", api.CLOSEST, d), ("any", api.ANY, p2),
                               ("closest_caller_order", api.CLOSEST | api.KEEP_RAY_ORDER, d), ("any_caller_order", api.ANY | api.KEEP_RAY_ORDER, p2)):
        sc.intersect_device(org.data_ptr(), second.data_ptr(), n, mode | api.COUNT_VISITS, ids.data_ptr(), t.data_ptr(), stream)
        torch.cuda.synchronize()
        cnt = sc.intersect_counters()
        sc.intersect_device(org.data_ptr(), second.data_ptr(), n, mode, ids.data_ptr(), t.data_ptr(), stream)   # warm-up
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            sc.intersect_device(org.data_ptr(), second.data_ptr(), n, mode, ids.data_ptr(), t.data_ptr(), stream)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        nodes, tris_r = cnt["nodes"] / cnt["rays"], cnt["tris"] / cnt["rays"]
        alg = n * (24 + 8 + nodes * bench.NODE_B + tris_r * bench.TRI_B)
        flops = n * (nodes * bench.NODE_FLOP + tris_r * bench.TRI_FLOP)
        out[name] = {"mrays_per_s": n / ms / 1e3, "ms": ms, "nodes_per_ray": nodes, "tris_per_ray": tris_r,
                     "hit_fraction": float((ids != -1).float().mean()) if not (mode & api.ANY) else float((ids == 0).float().mean()),
                     "roofline": pk.roofline("k_closest_pooled<0>" if not (mode & api.ANY) else "k_any_pooled<0, 0>", alg, flops, ms,
                                             info["node_bytes"] + info["tri_bytes"])}
    # parity on a subset: BVH == exhaustive kernel == CPU intersector (bit-exact ids and t)
    k = check
    if k > 0:
        o_h, d_h, p_h = org[:k].cpu().numpy(), d[:k].cpu().numpy(), p2[:k].cpu().numpy()
        a_id, a_t = sc.closest_hit(o_h, d_h); b_id, b_t = sc.closest_hit(o_h, d_h, brute_force=True)
        v_a, v_b = sc.visibility(o_h, p_h), sc.visibility(o_h, p_h, brute_force=True)
        out["parity"] = {"rays": k, "closest_id_mismatch_vs_exhaustive": int((a_id != b_id).sum()), "closest_t_bit_mismatch": int((a_t.view(np.uint32) != b_t.view(np.uint32)).sum()),
                         "any_mismatch_vs_exhaustive": int((v_a != v_b).sum())}
        if cpu_check > 0:
            from oracle import oracle_py as O
            f = sd.to_flat()
            P = O.Port(); h = P.scene(O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"]))
            c = min(cpu_check, k)
            r = P.closest_hit(h, o_h[:c], d_h[:c])
            out["parity"].update({"cpu_rays": c, "closest_id_mismatch_vs_cpu": int((a_id[:c] != r["id"]).sum()),
                                  "closest_t_bit_mismatch_vs_cpu": int((a_t[:c].view(np.uint32) != r["t"].view(np.uint32)).sum())})
    out["value"] = out["closest"]["mrays_per_s"]
    sc.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=10_000_000)
    ap.add_argument("--rays", type=int, default=100_000_000)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--check", type=int, default=2000, help="rays cross-checked against the exhaustive kernel")
    ap.add_argument("--cpu-check", type=int, default=50, help="rays cross-checked against the CPU intersector")
    args = ap.parse_args()
    import torch
    torch.cuda.set_stream(torch.cuda.Stream(device=torch.device("cuda", 0)))
    print(json.dumps(run(args.tris, args.rays, args.steps, args.check, args.cpu_check)))


if __name__ == "__main__":
    main()
