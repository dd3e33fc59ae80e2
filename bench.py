#!/usr/bin/env python
"""bench.py -- BDPT throughput of the B200 path (and, with --impl reference, of the reference's own CPU code).

Workload (BASELINE.json configs[1], "C2"): the reference's mis_test.txt scene, 1024 x 1024, 64 spp per GPU, spl 8,
eye / light depth 8.  A "step" is one complete render of that configuration.  With N GPUs every rank renders its
own 64 samples per pixel of a 64*N-spp job (sample indices s = rank (mod N), identical Philox streams as a 1-GPU
run of the same job) and the raw-sum films are reduced to rank 0 with one NCCL reduce ("weak" scaling).

  value      M eye samples / s, scene + BVH resident in HBM, film left on the device (CUDA events)
  e2e        the same through the public call with HOST buffers: scene upload + BVH build + render + film read-back
  roofline   dominant kernel of the step, algorithmic bytes / CUDA-event time vs the measured HBM copy bandwidth
  cpu_baseline  oracle/_ref (the reference's cpu_bdpt.cpp, unmodified) on the host cores, bounded sample
  reference_cuda_baseline  the reference's existing CUDA BDPT (bdpt_cu.cu, unmodified, built for sm_100a) on this GPU, same sample
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

WORKLOAD = dict(scene="mis", width=1024, height=1024, spp=64, spl=8, eye_depth=8, light_depth=8, seed=1)
CPU_SAMPLE = dict(width=512, height=512, spp=16)      # bounded sample of the same workload for the CPU arm


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
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
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
        time.sleep(0.15)
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


def cpu_arm(steps, warmup, sample=CPU_SAMPLE):
    """The reference's own CPU implementation (oracle/_ref = unmodified cpu_bdpt.cpp; else the C port) on all host
    cores, on a bounded sample of the workload (same scene, spl, depths; smaller film).  Returns the JSON fields."""
    from oracle import oracle_py as O
    from path_tracing_b200.scene import reference_scene
    w = dict(WORKLOAD); w.update(sample)
    sd = reference_scene(w["scene"]); f = sd.to_flat()
    d = O.SceneDesc(f["kind"], f["group"], f["geom"], f["mtl"], f["lights"], f["cam"], f["W"], f["H"])
    # every host core this process may run on, set explicitly: torchrun exports OMP_NUM_THREADS=1 to its children, which
    # would otherwise time the reference on ONE thread (measured: 2.6 instead of 37 Msamples/s)
    try:
        cores = len(os.sched_getaffinity(0))
    except AttributeError:
        cores = os.cpu_count() or 1
    n_samples = w["width"] * w["height"] * w["spp"]
    if O.have_ref():
        kind, lib = "reference", O.Ref()
        h = lib.scene(d)
        devnull = os.open(os.devnull, os.O_WRONLY); saved = os.dup(1)

        def run():
            sys.stdout.flush(); os.dup2(devnull, 1)       # the reference prints per-thread timings to stdout
            try:
                t0 = time.perf_counter()
                lib.render(h, d.cam, w["eye_depth"], w["light_depth"], w["width"], w["height"], w["spp"], w["spl"], nthreads=cores)
                return time.perf_counter() - t0
            finally:
                os.dup2(saved, 1)
    else:
        kind, lib = "port", O.Port()
        h = lib.scene(d)

        def run():
            t0 = time.perf_counter()
            lib.render(h, d.cam, w["eye_depth"], w["light_depth"], w["width"], w["height"], w["spp"], w["spl"], lib.opts(rng_mode=lib.MT19937, nthreads=cores))
            return time.perf_counter() - t0
    for _ in range(warmup):
        run()
    ts = [run() for _ in range(max(steps, 1))]
    sec = float(np.mean(ts))
    val = n_samples / sec / 1e6
    return {"value": val, "unit": "Msamples/s", "cores": int(cores), "kind": kind, "ms_per_step": sec * 1e3,
            "sample": f"{w['scene']} scene {w['width']}x{w['height']}, {w['spp']} spp, spl {w['spl']}, depth {w['eye_depth']}/{w['light_depth']} "
                      f"({n_samples / 1e6:.2f} M eye samples per step; full workload is {WORKLOAD['width']}x{WORKLOAD['height']} x {WORKLOAD['spp']} spp)"}


def ref_cuda_arm(sample=CPU_SAMPLE, timeout=180):
    """The reference's EXISTING CUDA BDPT (src/bdpt_cu.cu, compiled unmodified for sm_100a into oracle/_ref/libref_cuda.so)
    on this GPU, on the same bounded sample as the CPU arm: the same-hardware "kernel to beat" of SURVEY.md 2b.  Runs in
    a child process (tests/ref_cuda_probe.py) so that a failure of that code cannot take the bench line with it.
    Wall time of run_cuda_bdpt as main_cli.cpp:207-219 reports it.  None when the library is not there."""
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libref_cuda.so")):
        return None
    w = dict(WORKLOAD); w.update(sample)
    cmd = [sys.executable, os.path.join(ROOT, "tests", "ref_cuda_probe.py"), w["scene"], str(w["width"]), str(w["spp"]), str(w["spl"]),
           str(w["eye_depth"]), "bdpt", "ref-only"]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
        d = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
        return {"value": d["eye_msamples_per_s"], "unit": "Msamples/s", "kind": "reference CUDA kernels (bdpt_cu.cu, unmodified, sm_100a) on this GPU",
                "ms_per_step": d["ms"], "sample": f"{w['scene']} scene {w['width']}x{w['height']}, {w['spp']} spp, spl {w['spl']}, depth {w['eye_depth']}/{w['light_depth']}"}
    except Exception as e:          # noqa: BLE001 -- baseline evidence only
        return {"unavailable": f"{type(e).__name__}: {e}"[:200]}


def measured_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of all launches of `kernel` in ONE step of this workload, from the newest
    committed ncu summary (profiles/*_traffic.json, written by tools/ncu_traffic.py from an ncu launch list of
    tools/render_once.py 64 = the same render as a bench step).  None when no summary is committed."""
    import glob
    import re
    files = sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "*_traffic.json")),
                   key=lambda f: [int(x) for x in re.findall(r"\d+", os.path.basename(f))])      # r01_v7 < r01_v11
    if not files:
        return None, None
    try:
        k = json.load(open(files[-1]))["kernels"].get(kernel)
        return (k["dram_read_bytes"] + k["dram_write_bytes"], "profiles/" + os.path.basename(files[-1])) if k else (None, None)
    except Exception:
        return None, None


def gpu_arm(args):
    import torch
    import torch.distributed as dist
    from path_tracing_b200 import api, multi
    from path_tracing_b200.scene import reference_scene
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        if not os.environ.get("B2_KEEP_NCCL_DEBUG"):
            os.environ["NCCL_DEBUG"] = "WARN"        # NCCL's "NCCL version ..." banner goes to stdout and would precede the JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    torch.cuda.set_stream(torch.cuda.Stream(device=dev))      # a real (non-default) stream: the library launches on it, the events time it
    w = dict(WORKLOAD)
    for k in ("width", "height", "spp", "spl", "eye_depth", "light_depth"):
        if getattr(args, k, None):
            w[k] = getattr(args, k)
    if args.scene:
        w["scene"] = args.scene
    W, H = w["width"], w["height"]
    spp_total = w["spp"] * world                       # weak scaling: every GPU renders w["spp"] samples per pixel
    sd = reference_scene(w["scene"])
    sc = api.Scene(sd, device=local)
    film = torch.zeros((H, W, 3), dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    kw = dict(width=W, height=H, spp=spp_total, spl=w["spl"], eye_depth=w["eye_depth"], light_depth=w["light_depth"], seed=w["seed"],
              sample_offset=rank, sample_stride=world, film_mode=api.FILM_SUM if world > 1 else api.FILM_MEAN)

    def step(profile=False, count=False):
        _, st = sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream, profile_stages=profile, count_traversal=count, **kw)
        if world > 1:
            multi.reduce_film(film, spp_total, dst=0)                # one NCCL reduce of the raw-sum films over NVLink, mean on rank 0
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        st = step()
    # L2: every step writes > 1 GB of path state (max_batch_paths x ~400 B) between reuses of any line: no explicit flush needed
    sampler = ClockSampler(local); sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    launches = 0
    for _ in range(args.steps):
        st = step()
        launches += st["n_launches"]
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / args.steps
    # K steps take ~70 ms and nvidia-smi needs longer than that to print its first row (several hundred ms with 8 GPUs on the
    # box): when the timed region ended before two rows arrived, the same step keeps the GPU under the same load, untimed,
    # until they have (at most 4 s)
    n_timed, t_wait = len(sampler.rows), time.time()
    while sampler.p and len(sampler.rows) < 2 and time.time() - t_wait < 4.0:
        sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream, **kw)      # the render alone: ranks leave this loop at different times, no collective here
        torch.cuda.synchronize()
    clocks = sampler.stop()
    clocks["samples_in_timed_region"] = n_timed
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    n_samples = W * H * w["spp"] * world
    value = n_samples / (ms * 1e-3) / 1e6

    # ---- untimed diagnostics on rank 0: stage shares, traversal visit counts, e2e through host buffers ----
    out = None
    if rank == 0:
        kw1 = dict(kw); kw1.update(spp=w["spp"], sample_offset=0, sample_stride=1, film_mode=api.FILM_MEAN)
        _, sp = sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream, profile_stages=True, **kw1)
        _, sn = sc.render_bdpt(device_ptr=film.data_ptr(), stream=stream, count_traversal=True, **kw1)
        stages = {k: sp[k] for k in ("ms_light_pool", "ms_primary", "ms_extend", "ms_connect")}
        rays = sp["closest_rays"] + sp["shadow_rays"]
        n_primary = sp["eye_samples"]; n_later = sp["closest_rays"] - n_primary
        n_cv = sp["pairs_tested"] / max(sn["n_pool_active"], 1) if "n_pool_active" in sn else None
        # dominant kernel and its algorithmic bytes per step: node / triangle fetches use the measured
        # visit counts of the counting variant (80 B per wide node, 48 B per triangle); path state and eye-vertex records
        # are counted once per use; shadow-ray visits are attributed to k_connect, closest-hit visits to k_extend.
        dom = max(("ms_primary", "ms_extend", "ms_connect"), key=lambda k: stages[k])
        per_node, per_tri = 80, 48
        trav_bytes = sn["nodes_visited"] * per_node + sn["tris_tested"] * per_tri
        closest_share = sp["closest_rays"] / max(rays, 1)
        n_ev = sp["eye_vertices"] if "eye_vertices" in sp else 0
        alg = {
            "ms_primary": trav_bytes * closest_share * (n_primary / max(sp["closest_rays"], 1)) + n_ev * (80 + 84),
            "ms_extend": n_later * (4 + 80) + trav_bytes * closest_share * (n_later / max(sp["closest_rays"], 1)) + n_ev * (80 + 84),
            "ms_connect": sp["pairs_tested"] * 48 / 32 + sp["shadow_rays"] * (80 + 80) + trav_bytes * (1 - closest_share),
        }[dom]
        hbm, hbm_src = peaks()
        traffic, traffic_src = measured_traffic({"ms_primary": "k_extend<1, 0>", "ms_extend": "k_extend<0, 0>", "ms_connect": "k_connect<0, 0>"}[dom])
        # e2e: public host-buffer call; scene upload + BVH build + render + film read-back inside the timed region
        # the caller's film buffer is page-locked host memory (the contract's "pinned host memory"): b2_render_bdpt then reads
        # the film straight into it; a pageable numpy array works too and goes through the library's staging buffer
        host_film = torch.zeros((H, W, 3), dtype=torch.float32, pin_memory=True).numpy()
        e2e_ms = []
        for i in range(max(2, min(args.steps, 5)) + 1):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            s2 = api.Scene(sd, device=local)
            s2.render_bdpt(out=host_film, **kw1)
            s2.close()
            dtm = (time.perf_counter() - t0) * 1e3
            if i > 0:
                e2e_ms.append(dtm)
        e2e = float(np.mean(e2e_ms))
        dom_ms = stages[dom]
        out = {
            "metric": "bdpt_eye_msamples_per_s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"C2: reference mis_test.txt scene, {W}x{H}, {w['spp']} spp per GPU ({spp_total} spp job), spl {w['spl']}, "
                                   f"eye/light depth {w['eye_depth']}/{w['light_depth']}, BDPT, Philox seed {w['seed']}",
                       "l2": "inputs larger than L2: each step streams > 1 GB of path state", "parallelism": f"spp-sharded x{world}, NCCL film reduce"},
            "mrays_per_s": rays * world / (ms * 1e-3) / 1e6, "mpath_samples_per_s": sp["path_samples"] * world / (ms * 1e-3) / 1e6,
            "per_sample": {"closest_rays": sp["closest_rays"] / sp["eye_samples"], "shadow_rays": sp["shadow_rays"] / sp["eye_samples"],
                           "pairs_tested": sp["pairs_tested"] / sp["eye_samples"], "nodes_per_ray": sn["nodes_visited"] / max(rays, 1),
                           "tris_per_ray": sn["tris_tested"] / max(rays, 1)},
            "stages_ms": stages, "stage_total_ms": sp["ms_total"],
            "roofline": {"bound": "hbm", "kernel": {"ms_primary": "k_extend<PRIMARY>", "ms_extend": "k_extend", "ms_connect": "k_connect"}[dom],
                         "achieved": alg / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else None, "peak": hbm, "unit": "GB/s",
                         "frac": (alg / (dom_ms * 1e-3) / 1e9) / hbm if dom_ms > 0 else None, "traffic": traffic, "traffic_source": traffic_src, "peak_source": hbm_src,
                         "algorithmic_bytes_per_step": alg, "kernel_ms_per_step": dom_ms},
            "e2e": {"value": W * H * w["spp"] / (e2e * 1e-3) / 1e6, "unit": "Msamples/s", "ms_per_step": e2e,
                    "h2d_bytes_per_step": scene_bytes(sd), "d2h_bytes_per_step": int(host_film.nbytes),
                    "note": "b2_scene_create + b2_scene_build_bvh + b2_render_bdpt(host film) per step, 1 GPU"},
            "gpu_launches": int(launches), "clocks": clocks,
        }
        if world == 1 and not args.no_cpu:
            out["cpu_baseline"] = cpu_arm(1, 0)
            rc = ref_cuda_arm()
            if rc is not None:
                out["reference_cuda_baseline"] = rc
    sc.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if out is not None:
        print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--scene", default=None)
    ap.add_argument("--no-cpu", action="store_true")
    for k in ("width", "height", "spp", "spl", "eye_depth", "light_depth"):
        ap.add_argument("--" + k.replace("_", "-"), type=int, default=None, dest=k)
    args = ap.parse_args()
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return
        c = cpu_arm(min(args.steps, 5), min(args.warmup, 1))
        W, H = WORKLOAD["width"], WORKLOAD["height"]
        line = {"impl": "reference", "metric": "bdpt_eye_msamples_per_s", "value": c["value"], "unit": "Msamples/s", "n_gpus": args.gpus,
                "steps": min(args.steps, 5), "warmup": min(args.warmup, 1), "ms_per_step": c["ms_per_step"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"C2: reference mis_test.txt scene, {W}x{H}, {WORKLOAD['spp']} spp, spl {WORKLOAD['spl']}, eye/light depth "
                                       f"{WORKLOAD['eye_depth']}/{WORKLOAD['light_depth']}, BDPT (timed on the bounded sample in cpu_baseline.sample)"},
                "cpu_baseline": c, "e2e": {"value": c["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return
    gpu_arm(args)


if __name__ == "__main__":
    main()
