"""C4 of BASELINE.json / SURVEY.md 8(d): procedural 10 M-triangle scene, 3840x2160, spp sharded s = g (mod G) over the G GPUs
of one node (one process per GPU, torchrun), NCCL reduce of the raw-sum films, resolve on rank 0.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port P tools/c4_multi.py --spp 64
The configured job is 1024 spp; --spp reduces it (rates are per sample, the pool is per render).  With --check the job
is also rendered at a reduced size on rank 0 alone and compared (film(G) == film(1) up to fp32 summation order)."""
import argparse, json, os, sys, time
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, ".")
from path_tracing_b200 import api, gen, multi

ap = argparse.ArgumentParser()
ap.add_argument("--tris", type=int, default=10_000_000)
ap.add_argument("--spp", type=int, default=64)
ap.add_argument("--spl", type=int, default=8)
ap.add_argument("--depth", type=int, default=8)
ap.add_argument("--width", type=int, default=3840)
ap.add_argument("--height", type=int, default=2160)
ap.add_argument("--check", action="store_true")
a = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
os.environ.setdefault("NCCL_DEBUG", "WARN")
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
t0 = time.perf_counter(); sd = gen.terrain_scene(a.tris, width=a.width, height=a.height); t1 = time.perf_counter()
sc = api.Scene(sd, device=local); t2 = time.perf_counter()
film = torch.zeros((a.height, a.width, 3), dtype=torch.float32, device=f"cuda:{local}")
stream = torch.cuda.Stream()
off, stride, n_local = multi.shard(rank, world, a.spp)
kw = dict(width=a.width, height=a.height, spp=a.spp, spl=a.spl, eye_depth=a.depth, light_depth=a.depth, seed=1)


def step():
    with torch.cuda.stream(stream):
        _, st = sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream.cuda_stream, sample_offset=off, sample_stride=stride, film_mode=api.FILM_SUM, **kw)
        multi.reduce_film(film, a.spp)
    return st


sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream.cuda_stream, **dict(kw, spp=1))     # warm-up: allocations, first launches
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(stream); st = step(); e1.record(stream)
torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1)], device=f"cuda:{local}")
cnt = torch.tensor([st["eye_samples"], st["closest_rays"], st["shadow_rays"], st["path_samples"]], dtype=torch.float64, device=f"cuda:{local}")
if world > 1:
    dist.all_reduce(ms, op=dist.ReduceOp.MAX); dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
res = None
if rank == 0:
    s = ms.item() * 1e-3
    f = film.cpu().numpy()
    res = dict(config=f"C4: terrain_scene {len(sd.tris)} triangles, {a.width}x{a.height}, {a.spp} spp job (configured: 1024), spl {a.spl}, depth {a.depth}/{a.depth}",
               n_gpus=world, generate_s=t1 - t0, upload_build_s=t2 - t1, bvh=sc.bvh_info(), ms=ms.item(), eye_msamples_per_s=cnt[0].item() / s / 1e6,
               mrays_per_s=(cnt[1].item() + cnt[2].item()) / s / 1e6, mpath_samples_per_s=cnt[3].item() / s / 1e6,
               shadow_rays_per_sample=cnt[2].item() / cnt[0].item(), film_mean=[float(x) for x in f.reshape(-1, 3).mean(0)], finite=bool(np.isfinite(f).all()))
    if a.check:                                      # same job at 480x270 on this GPU alone vs sharded
        kc = dict(kw, width=480, height=270)
        one, _ = sc.render_bdpt(**kc)
        parts = [sc.render_bdpt(sample_offset=g, sample_stride=world, film_mode=api.FILM_SUM, **kc)[0] for g in range(world)]
        shard_film = np.sum(parts, axis=0) / a.spp
        res["check_480x270_max_rel_diff"] = float(np.max(np.abs(shard_film - one) / np.maximum(np.abs(one), 1e-3)))
    print(json.dumps(res))
sc.close()
if world > 1:
    dist.barrier(); dist.destroy_process_group()
