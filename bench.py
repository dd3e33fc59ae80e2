#!/usr/bin/env python
"""bench.py -- BDPT throughput of the B200 path (and, with --impl reference, of the reference's own CPU code).

Headline workload (BASELINE.json configs[2], "C3", the configuration the north_star target is stated on): a generated
OBJ + MTL + PPM-texture scene of ~1 M triangles (path_tracing_b200.gen.write_obj_scene, fixed seed) loaded through the
reference's vendored tinyobjloader, 1920 x 1080, spl 8 (the reference's --spl default), eye / light depth 16.  A "step"
is a 32-spp slice of the configured 256-spp job (one complete BDPT render: light pool + all eye waves + film resolve).
With N GPUs (torchrun, one rank per GPU) the SAME 32-spp job is sharded over the ranks (sample index s = rank (mod N),
identical Philox streams as the 1-GPU run) and the raw-sum films are reduced to rank 0 with one NCCL reduce: "strong"
scaling of a fixed job.

  value       M eye samples / s of the whole job, scene + BVH resident in HBM, film left on the device (CUDA events)
  e2e         the same job through the public call with HOST buffers on N GPUs: per-rank scene upload + BVH build + shard
              render + NCCL reduce + film read-back to pinned host memory on rank 0, max over ranks
  roofline    dominant kernel of the step: algorithmic bytes and flops (SURVEY.md 8d constants x measured visit counts)
              against the memory leg that applies (L2 when nodes + triangles fit in it, else HBM; both measured) and the
              FP32 leg (in-run FMA microbenchmark); `bound` is whichever leg is slower
  cpu_baseline  oracle/_ref (the reference's cpu_bdpt.cpp, unmodified) on all host cores, same scene / spl / depths on
              the reduced film named in config.cpu_sample
  extra       C1, C2, the documented C2 variant and C5 as sub-records with their own roofline (N = 1 only)
--impl reference: the reference's CPU code alone, same config dict, same steps / warm-up.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

C3 = dict(tris=1 << 20, width=1920, height=1080, spp=32, job_spp=256, spl=8, eye_depth=16, light_depth=16, seed=1)
C3_CPU = dict(width=64, height=36, spp=1)          # bounded sample of the same workload for the CPU arm (seconds per step)
NODE_B, TRI_B, NODE_FLOP, TRI_FLOP = 80, 48, 160, 51


def make_config(n_gpus):
    """The SAME dict in both arms (the driver compares them)."""
    return {
        "workload": f"C3: generated ~1 M-triangle OBJ scene (gen.write_obj_scene, seed 1, {C3['tris']} requested) via tinyobjloader, "
                    f"{C3['width']}x{C3['height']}, {C3['spp']} spp per step (a slice of the {C3['job_spp']}-spp job), spl {C3['spl']}, "
                    f"eye/light depth {C3['eye_depth']}/{C3['light_depth']}, BDPT reference strategy set, Philox seed {C3['seed']}",
        "cpu_sample": f"reference arm / cpu_baseline: same scene, spl and depths on a {C3_CPU['width']}x{C3_CPU['height']} film, {C3_CPU['spp']} spp per step",
        "units": "value = M eye samples/s (one (pixel, s) iteration of cpu_bdpt.cpp:349); mpath_samples_per_s = connections that reach the MIS "
                 "evaluation (cpu_bdpt.cpp:431), the unit the north_star's 2e9 target is read against",
        "l2": "scene tables (nodes + triangles ~ 63 MB) are L2-resident by design; every step streams > 1 GB of path state and eye-vertex records "
              "through HBM between reuses, so no explicit flush is needed",
        "parallelism": f"{n_gpus} GPU(s): scene + BVH replicated, the step's {C3['spp']} spp sharded s = rank (mod N), one NCCL reduce of the fp32 film",
    }


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu):
        self.gpu, self.rows, self.p = gpu, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except OSError:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.p.terminate()
        try:
            self.p.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.p.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def scene_bytes(sd):
    return int(sd.tris.nbytes + sd.spheres.nbytes + sd.materials.nbytes + sd.lights.nbytes + sd.camera.nbytes)


def host_cores():
    # every host core this process may run on, set explicitly: torchrun exports OMP_NUM_THREADS=1 to its children, which
    # would otherwise time the reference on ONE thread
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def c3_scene():
    """The generated C3 scene, parsed through b2_parse_scene_file_ex (`O mesh.obj` -> tinyobjloader).  The files are
    written once per box into a temp directory keyed by the generator's arguments and reused by later processes."""
    import tempfile
    from path_tracing_b200 import gen
    from path_tracing_b200.scene import SceneData
    key = hashlib.sha1(repr((C3["tris"], C3["width"], C3["height"], "v2")).encode()).hexdigest()[:12]
    d = os.path.join(tempfile.gettempdir(), f"b2_c3_scene_{key}")
    scene_txt = os.path.join(d, "scene.txt")
    if not os.path.exists(scene_txt):
        tmp = tempfile.mkdtemp(prefix="b2_c3_gen_")
        gen.write_obj_scene(tmp, n_tris=C3["tris"], width=C3["width"], height=C3["height"])
        try:
            os.rename(tmp, d)                       # atomic: another rank may have won the race
        except OSError:
            import shutil
            shutil.rmtree(tmp, ignore_errors=True)
    return SceneData.from_file(scene_txt, obj_directive=True)


class CpuArm:
    """The reference's own CPU implementation (oracle/_ref = unmodified cpu_bdpt.cpp; else the C port) on all host cores."""

    def __init__(self, sd):
        from oracle import oracle_py as O
        f = sd.to_flat()
        self.d = O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])
        self.cores = host_cores()
        self.kind = "reference" if O.have_ref() else "port"
        self.lib = O.Ref() if self.kind == "reference" else O.Port()
        self.h = self.lib.scene(self.d)

    def render(self, width, height, spp, spl, eye_depth, light_depth):
        """seconds of one run_cpu_bdpt call"""
        if self.kind == "reference":
            devnull = os.open(os.devnull, os.O_WRONLY); saved = os.dup(1)
            sys.stdout.flush(); os.dup2(devnull, 1)       # the reference prints per-thread timings to stdout
            try:
                t0 = time.perf_counter()
                self.lib.render(self.h, self.d.cam, eye_depth, light_depth, width, height, spp, spl, nthreads=self.cores)
                return time.perf_counter() - t0
            finally:
                os.dup2(saved, 1); os.close(devnull); os.close(saved)
        t0 = time.perf_counter()
        self.lib.render(self.h, self.d.cam, eye_depth, light_depth, width, height, spp, spl, self.lib.opts(rng_mode=self.lib.MT19937, nthreads=self.cores))
        return time.perf_counter() - t0

    def timed(self, steps, warmup, width, height, spp, spl, eye_depth, light_depth, label):
        for _ in range(warmup):
            self.render(width, height, spp, spl, eye_depth, light_depth)
        ts = [self.render(width, height, spp, spl, eye_depth, light_depth) for _ in range(max(steps, 1))]
        sec = float(np.mean(ts)); n = width * height * spp
        return {"value": n / sec / 1e6, "unit": "Msamples/s", "cores": int(self.cores), "kind": self.kind, "ms_per_step": sec * 1e3,
                "steps": max(steps, 1), "warmup": warmup,
                "sample": f"{label} {width}x{height}, {spp} spp, spl {spl}, depth {eye_depth}/{light_depth} ({n} eye samples per step)"}


def ref_cuda_arm(scene, width, spp, spl, depth, timeout=180):
    """The reference's EXISTING CUDA BDPT (src/bdpt_cu.cu, compiled unmodified for sm_100a into oracle/_ref/libref_cuda.so)
    on this GPU: the same-hardware kernel of SURVEY.md 2b (brute force: small scenes only).  Runs in a child process
    (tests/ref_cuda_probe.py) so that a failure of that code cannot take the bench line with it."""
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref_cuda.so")):
        return None
    cmd = [sys.executable, os.path.join(ROOT, "tests", "ref_cuda_probe.py"), scene, str(width), str(spp), str(spl), str(depth), "bdpt", "ref-only"]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
        d = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
        return {"value": d["eye_msamples_per_s"], "unit": "Msamples/s", "kind": "reference CUDA kernels (bdpt_cu.cu, unmodified, sm_100a) on this GPU",
                "ms_per_step": d["ms"], "sample": f"{scene} scene {width}x{width}, {spp} spp, spl {spl}, depth {depth}/{depth}"}
    except Exception as e:          # noqa: BLE001 -- baseline evidence only
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}


def measured_traffic(kernel, version, workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of all launches of `kernel` in one step of `workload`, from the committed
    ncu summary profiles/*_traffic.json whose `version` is this library's b2_version() (tools/ncu_traffic.py stamps it);
    None when no summary of this very build is committed."""
    import glob
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")), reverse=True):
        try:
            d = json.load(open(f))
            if d.get("version") != version or d.get("workload") != workload:
                continue
            k = d["kernels"].get(kernel)
            if k:
                scale = float(d.get("per_step_scale", 1.0))
                return (k["dram_read_bytes"] + k["dram_write_bytes"]) * scale, "profiles/" + os.path.basename(f)
        except Exception:      # noqa: BLE001
            continue
    return None, None


class Peaks:
    """Roofline denominators: HBM from MEASURED_PEAKS.json (driver-written), L2 and FP32 measured in this run."""

    def __init__(self, device):
        import torch
        from path_tracing_b200 import api
        self.hbm, self.hbm_src = peaks()
        self.l2 = api.measure_l2_bandwidth(device)
        self.fp32 = api.measure_fp32_peak(device)
        self.l2_bytes = int(torch.cuda.get_device_properties(device).L2_cache_size)

    def roofline(self, kernel, alg_bytes, flops, kernel_ms, resident_bytes, traffic=None, traffic_src=None):
        """T_bound = max(bytes / BW of the memory level the working set lives in, flops / FP32 peak); frac = T_bound / T."""
        t = kernel_ms * 1e-3
        in_l2 = resident_bytes <= 0.6 * self.l2_bytes          # the 126 MB L2 is two 63 MB halves; stay well inside
        legs = {"hbm": alg_bytes / (self.hbm * 1e9), "l2": alg_bytes / (self.l2 * 1e9), "fp32": flops / (self.fp32 * 1e12)}
        mem = "l2" if in_l2 else "hbm"
        bound = mem if legs[mem] >= legs["fp32"] else "fp32"
        if bound == "fp32":
            achieved, peak, unit = flops / t / 1e12, self.fp32, "TFLOP/s"
        else:
            achieved, peak, unit = alg_bytes / t / 1e9, (self.l2 if bound == "l2" else self.hbm), "GB/s"
        return {"bound": bound, "kernel": kernel, "achieved": achieved, "peak": peak, "unit": unit, "frac": achieved / peak if peak else None,
                "traffic": traffic, "traffic_source": traffic_src,
                "legs_ms": {k: v * 1e3 for k, v in legs.items()}, "memory_leg": mem, "working_set_bytes": int(resident_bytes), "l2_cache_bytes": self.l2_bytes,
                "peaks": {"hbm_gbs": self.hbm, "hbm_source": self.hbm_src, "l2_gbs": self.l2, "l2_source": "in-run read of a 32 MiB L2-resident buffer (b2_measure_l2_bandwidth)",
                          "fp32_tflops": self.fp32, "fp32_source": "in-run FMA microbenchmark (b2_measure_fp32_peak)"},
                "algorithmic_bytes_per_step": alg_bytes, "flops_per_step": flops, "kernel_ms_per_step": kernel_ms,
                "frac_of_hbm": alg_bytes / t / 1e9 / self.hbm}


def stage_roofline(pk, sc, sp, sn, version, workload):
    """Roofline of the stage with the largest CUDA-event time.  sp: stats of a profile_stages render, sn: of a count_traversal
    render of the same job.  Node / triangle visits are the counting variant's; shadow-ray visits belong to k_connect,
    closest-hit visits to k_extend."""
    stages = {k: sp[k] for k in ("ms_light_pool", "ms_primary", "ms_extend", "ms_connect")}
    rays = sp["closest_rays"] + sp["shadow_rays"]
    n_primary = sp["eye_samples"]; n_later = max(sp["closest_rays"] - n_primary, 0)
    dom = max(("ms_primary", "ms_extend", "ms_connect"), key=lambda k: stages[k])
    closest_share = sp["closest_rays"] / max(rays, 1)
    trav_bytes = sn["nodes_visited"] * NODE_B + sn["tris_tested"] * TRI_B
    trav_flops = sn["nodes_visited"] * NODE_FLOP + sn["tris_tested"] * TRI_FLOP
    n_ev = sp.get("eye_vertices", 0)
    share = {"ms_primary": closest_share * n_primary / max(sp["closest_rays"], 1), "ms_extend": closest_share * n_later / max(sp["closest_rays"], 1),
             "ms_connect": 1.0 - closest_share}[dom]
    alg = {"ms_primary": n_ev * (80 + 84), "ms_extend": n_later * (4 + 80) + n_ev * (80 + 84),
           "ms_connect": sp["pairs_tested"] * 48 / 32 + sp["shadow_rays"] * (80 + 80)}[dom] + trav_bytes * share
    info = sc.bvh_info()
    kern = {"ms_primary": "k_extend<1, 0>", "ms_extend": "k_extend<0, 0>", "ms_connect": "k_connect<0, 1>" if info["n_tris"] >= 2048 else "k_connect<0, 0>"}[dom]
    traffic, src = measured_traffic(kern, version, workload)
    r = pk.roofline(kern, alg, trav_flops * share, stages[dom], info["node_bytes"] + info["tri_bytes"], traffic, src)
    return r, stages


def render_record(pk, sd, name, workload, kw, version, steps=3, cpu=None, device=0):
    """One sub-record of `extra`: a small BDPT configuration rendered `steps` times (after 2 warm-ups) with its own roofline."""
    import torch
    from path_tracing_b200 import api
    sc = api.Scene(sd, device=device)
    film = torch.zeros((kw["height"], kw["width"], 3), dtype=torch.float32, device=f"cuda:{device}")
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(2):
        sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream, **kw)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(steps):
        _, st = sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream, **kw)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    _, sp = sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream, profile_stages=True, **kw)
    _, sn = sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream, count_traversal=True, **kw)
    roof, stages = stage_roofline(pk, sc, sp, sn, version, name)
    rays = st["closest_rays"] + st["shadow_rays"]
    rec = {"workload": workload, "value": st["eye_samples"] / ms / 1e3, "unit": "Msamples/s", "ms_per_step": ms, "steps": steps,
           "mrays_per_s": rays / ms / 1e3, "mpath_samples_per_s": st["path_samples"] / ms / 1e3,
           "per_sample": {"closest_rays": st["closest_rays"] / st["eye_samples"], "shadow_rays": st["shadow_rays"] / st["eye_samples"],
                          "nodes_per_ray": sn["nodes_visited"] / max(rays, 1), "tris_per_ray": sn["tris_tested"] / max(rays, 1)},
           "stages_ms": stages, "roofline": roof, "film_mean": float(film.mean().item())}
    if cpu is not None:
        rec["cpu_baseline"] = cpu
    sc.close()
    return rec


def extras(pk, version, device, with_cpu):
    """C1, C2, the documented C2 variant, C5 (BASELINE.json configs[0], [1], [4]) as sub-records of the bench line."""
    from path_tracing_b200.scene import reference_scene
    out = {}
    c1 = dict(width=512, height=512, spp=16, spl=8, eye_depth=8, light_depth=8, seed=1)
    c2 = dict(width=1024, height=1024, spp=64, spl=8, eye_depth=8, light_depth=8, seed=1)
    try:
        sd = reference_scene("cornell")
        cpu = None
        if with_cpu:        # the reference needs minutes for the full C1 film: same scene / spl / depth at 128 x 128, 4 spp
            cpu = CpuArm(sd).timed(2, 1, 128, 128, 4, 8, 8, 8, "input.txt scene")
        out["C1"] = render_record(pk, sd, "C1", "C1: reference input.txt scene, 512x512, 16 spp, spl 8, depth 8/8 (the loop-C workload: ~450 shadow rays per eye sample)",
                                  c1, version, steps=3, cpu=cpu, device=device)
    except Exception as e:      # noqa: BLE001 -- a failing sub-record must not take the headline with it
        out["C1"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    for key, scene, label in (("C2", "mis", "reference mis_test.txt scene as parsed"), ("C2_variant", "mis_variant", "documented variant of mis_test.txt (6-number M lines, cutoff 180)")):
        try:
            sd = reference_scene(scene)
            cpu = CpuArm(sd).timed(2, 1, 1024, 1024, 64 if scene == "mis" else 4, 8, 8, 8, label) if with_cpu else None    # C2 full size on both arms
            out[key] = render_record(pk, sd, key, f"{key}: {label}, 1024x1024, 64 spp, spl 8, depth 8/8", c2, version, steps=5, cpu=cpu, device=device)
            if key == "C2" and with_cpu:
                rc = ref_cuda_arm("mis", 512, 16, 8, 8)
                if rc is not None:
                    out[key]["reference_cuda_baseline"] = rc
        except Exception as e:      # noqa: BLE001
            out[key] = {"error": f"{type(e).__name__}: {e}"[:300]}
    try:
        import bench_traversal
        out["C5"] = bench_traversal.run(10_000_000, 100_000_000, steps=2, check=2000, cpu_check=0, pk=pk, device=device)
    except Exception as e:      # noqa: BLE001
        out["C5"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    return out


def gpu_arm(args):
    import torch
    import torch.distributed as dist
    from path_tracing_b200 import api, multi, _abi
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)          # NCCL_DEBUG is left as the caller set it
    torch.cuda.set_stream(torch.cuda.Stream(device=dev))      # a real (non-default) stream: the library launches on it, the events time it
    version = _abi.lib().b2_version().decode()
    w = dict(C3)
    for k in ("width", "height", "spp", "spl"):
        if getattr(args, k, None):
            w[k] = getattr(args, k)
    W, H = w["width"], w["height"]
    sd = c3_scene()
    sc = api.Scene(sd, device=local)
    film = torch.zeros((H, W, 3), dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    off, stride, n_local = multi.shard(rank, world, w["spp"])
    kw = dict(width=W, height=H, spp=w["spp"], spl=w["spl"], eye_depth=w["eye_depth"], light_depth=w["light_depth"], seed=w["seed"],
              sample_offset=off, sample_stride=stride, film_mode=api.FILM_SUM if world > 1 else api.FILM_MEAN)

    def step():
        _, st = sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream, **kw)
        if world > 1:
            multi.reduce_film(film, w["spp"], dst=0)                # one NCCL reduce of the raw-sum films over NVLink, mean on rank 0
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    warm = max(args.warmup, 3)
    for _ in range(warm):
        st = step()
    sampler = ClockSampler(local); sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    launches = 0
    for _ in range(args.steps):
        st = step()
        launches += st["n_launches"] + (1 if world > 1 else 0)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / args.steps
    n_timed = len(sampler.rows)
    clocks = sampler.stop()
    clocks["samples_in_timed_region"] = n_timed
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    cnt = torch.tensor([st["eye_samples"], st["closest_rays"], st["shadow_rays"], st["path_samples"], launches], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    ms = float(t.item())
    eye_samples, closest, shadow, paths, launches = [float(x) for x in cnt.tolist()]
    value = eye_samples / (ms * 1e-3) / 1e6

    # ---- e2e on N GPUs: what a caller pays from HOST buffers, every rank: scene upload + BVH build + shard render + NCCL
    #      reduce, film read-back to pinned host memory on rank 0; max over ranks (device-synchronised wall clock) ----
    host_film = torch.zeros((H, W, 3), dtype=torch.float32, pin_memory=True)
    e2e_ms, e2e_parts = [], []
    for i in range(3 if args.steps >= 3 else 2):
        barrier()
        t0 = time.perf_counter()
        s2 = api.Scene(sd, device=local)                      # b2_scene_create (flatten + upload) + b2_scene_build_bvh; synchronous
        t1 = time.perf_counter()
        s2.render_bdpt(device_ptr=film.data_ptr(), stream=stream, **kw)   # returns when the stream has drained the render
        t2 = time.perf_counter()
        if world > 1:
            multi.reduce_film(film, w["spp"], dst=0)
        if rank == 0:
            host_film.copy_(film, non_blocking=True)
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        s2.close()
        t4 = time.perf_counter()
        dt = torch.tensor([(t4 - t0) * 1e3, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        if i > 0:
            e2e_ms.append(float(dt[0].item())); e2e_parts.append([float(x) for x in dt[1:].tolist()])
    e2e_iters = [[round(x, 3) for x in it] for it in e2e_parts]
    e2e_parts = dict(zip(("scene_create_plus_bvh_build", "render_wall", "reduce_plus_film_readback", "scene_destroy"), [float(x) for x in np.mean(np.array(e2e_parts), axis=0)]))
    e2e_parts["per_iteration"] = e2e_iters
    e2e = float(np.mean(e2e_ms))

    out = None
    if rank == 0:
        # untimed diagnostics on rank 0: stage shares and traversal visit counts of this rank's shard
        _, sp = sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream, profile_stages=True, **kw)
        _, sn = sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream, count_traversal=True, **kw)
        pk = Peaks(local)
        roof, stages = stage_roofline(pk, sc, sp, sn, version, "C3")
        rays = closest + shadow
        info = sc.bvh_info()
        out = {
            "metric": "bdpt_eye_msamples_per_s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": make_config(world),
            "mrays_per_s": rays / (ms * 1e-3) / 1e6, "mpath_samples_per_s": paths / (ms * 1e-3) / 1e6,
            "target": {"north_star": "2e9 path samples/s on 8 GPUs at C3", "unit": "mpath_samples_per_s", "this_run_fraction_of_target": paths / (ms * 1e-3) / 2e9},
            "per_sample": {"closest_rays": closest / eye_samples, "shadow_rays": shadow / eye_samples, "path_samples": paths / eye_samples,
                           "pairs_tested": sp["pairs_tested"] / sp["eye_samples"],
                           "nodes_per_ray": sn["nodes_visited"] / max(sn["closest_rays"] + sn["shadow_rays"], 1),
                           "tris_per_ray": sn["tris_tested"] / max(sn["closest_rays"] + sn["shadow_rays"], 1)},
            "scene": {"n_tris": int(info["n_tris"]), "n_nodes": int(info["n_nodes"]), "node_bytes": int(info["node_bytes"]), "tri_bytes": int(info["tri_bytes"]),
                      "bvh_depth": int(info["depth"]), "bvh_build_ms": float(info["build_ms"]), "sah_cost": float(info["sah_cost"]),
                      "obj_reader": "tinyobjloader" if _abi.lib().b2_obj_has_tinyobj() else "own reader (library built without tiny_obj_loader.h)"},
            "stages_ms": stages, "stage_total_ms": sp["ms_total"], "stages_note": "rank 0's shard, one untimed profiled render",
            "roofline": roof,
            "e2e": {"value": eye_samples / (e2e * 1e-3) / 1e6, "unit": "Msamples/s", "ms_per_step": e2e, "parts_ms_max_over_ranks": e2e_parts, "n_gpus": world,
                    "h2d_bytes_per_step": scene_bytes(sd) * world, "d2h_bytes_per_step": int(host_film.numel() * 4),
                    "note": f"per rank: b2_scene_create + b2_scene_build_bvh + shard render; NCCL reduce; film to pinned host memory on rank 0; max over {world} rank(s)"},
            "gpu_launches": int(launches), "clocks": clocks, "version": version,
        }
        if world == 1 and not args.no_extra:
            if not args.no_cpu:
                out["cpu_baseline"] = CpuArm(sd).timed(3, 1, C3_CPU["width"], C3_CPU["height"], C3_CPU["spp"], w["spl"], w["eye_depth"], w["light_depth"], "C3 scene")
            out["extra"] = extras(pk, version, local, with_cpu=not args.no_cpu)
    sc.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        sys.stdout.flush()
        print(json.dumps(out), flush=True)          # the LAST line of stdout (NCCL's own banner, if any, comes earlier)


def reference_arm(args):
    """The reference's unmodified cpu_bdpt.cpp on all host cores: same config dict, same steps / warm-up, each step the
    bounded sample config.cpu_sample names."""
    sd = c3_scene()
    arm = CpuArm(sd)
    c = arm.timed(args.steps, args.warmup, C3_CPU["width"], C3_CPU["height"], C3_CPU["spp"], C3["spl"], C3["eye_depth"], C3["light_depth"], "C3 scene")
    line = {"impl": "reference", "metric": "bdpt_eye_msamples_per_s", "value": c["value"], "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": c["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": make_config(args.gpus),
            "cpu_baseline": c, "e2e": {"value": c["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baselines")
    ap.add_argument("--no-extra", action="store_true", help="skip cpu_baseline and the C1 / C2 / C5 sub-records")
    for k in ("width", "height", "spp", "spl"):
        ap.add_argument("--" + k, type=int, default=None, dest=k, help="development override of the C3 workload")
    args = ap.parse_args()
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return
        reference_arm(args)
        return
    gpu_arm(args)


if __name__ == "__main__":
    main()
