"""ctypes loader of tests/hostsim/libhostsim.so: a TEST-ONLY host compilation of the device headers (b2_*.cuh), used to
check the BVH build / traversal / wavefront stage logic against the oracle on machines without a GPU.  Not product code."""
import ctypes as C
import os
import subprocess
import numpy as np
from path_tracing_b200 import _abi as A

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hostsim")
SO = os.path.join(HERE, "libhostsim.so")
ROOT = os.path.dirname(os.path.dirname(HERE))


def build():
    srcs = [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith((".cpp", ".hpp"))]
    csrc = os.path.join(ROOT, "path_tracing_b200", "csrc")
    srcs += [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith((".cuh", ".hpp"))]
    if os.path.exists(SO) and all(os.path.getmtime(s) <= os.path.getmtime(SO) for s in srcs):
        return
    subprocess.run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-Wno-unknown-pragmas",
                    "-I/usr/local/cuda/include", "-I" + os.path.join(ROOT, "include"),
                    os.path.join(HERE, "hostsim.cpp"), "-o", SO], check=True)


def mis_weight(eye, s_idx, lit, t_idx, d_vec, dist2, cam_pos, factored=False):
    """mis_weight_core of csrc/b2_bdpt.cuh compiled for the host, over flat chains (eye [n, V, 17], lit [n, V, 20]);
    factored: the two walks through mis_eye_sum / mis_light_sum (what k_pool_active + k_connect evaluate)."""
    build()
    L = C.CDLL(SO)
    eye = np.ascontiguousarray(eye, np.float32); lit = np.ascontiguousarray(lit, np.float32)
    n, V = eye.shape[0], eye.shape[1]
    s_idx = np.ascontiguousarray(s_idx, np.int32); t_idx = np.ascontiguousarray(t_idx, np.int32)
    d_vec = np.ascontiguousarray(d_vec, np.float32); dist2 = np.ascontiguousarray(dist2, np.float32); cam_pos = np.ascontiguousarray(cam_pos, np.float32)
    out = np.zeros(n, np.float32)
    vp = C.c_void_p
    L.hs_mis_weight.argtypes = [C.c_size_t, C.c_int, vp, vp, vp, vp, vp, vp, vp, C.c_int, vp]
    L.hs_mis_weight(n, V, A.ptr(eye), A.ptr(s_idx), A.ptr(lit), A.ptr(t_idx), A.ptr(d_vec), A.ptr(dist2), A.ptr(cam_pos), int(factored), A.ptr(out))
    return out


class HostSim:
    def __init__(self, data, max_leaf=0, pad_rel=0.0):
        build()
        L = self.L = C.CDLL(SO)
        vp, sz = C.c_void_p, C.c_size_t
        L.hs_scene_create.restype = vp
        L.hs_scene_create.argtypes = [vp, sz, vp, sz, vp, sz, vp, sz, C.c_int, C.c_float]
        L.hs_scene_destroy.argtypes = [vp]
        L.hs_bvh_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.hs_intersect.argtypes = [vp, vp, vp, sz, C.c_int, vp, vp, vp]
        L.hs_render_bdpt.argtypes = [vp, vp, C.POINTER(A.RenderParams), vp, vp]
        L.hs_render_bdpt_factored.argtypes = [vp, vp, C.POINTER(A.RenderParams), vp, vp]
        L.hs_light_pool.argtypes = [vp, C.POINTER(A.RenderParams), vp]
        L.hs_render_other.argtypes = [vp, vp, C.POINTER(A.RenderParams), vp, vp]
        L.hs_render_strategies_c.argtypes = [vp, vp, C.POINTER(A.RenderParams), vp, vp]
        self.data = data
        self.h = L.hs_scene_create(A.ptr(data.tris), len(data.tris), A.ptr(data.spheres), len(data.spheres), A.ptr(data.materials),
                                   len(data.materials), A.ptr(data.lights), len(data.lights), max_leaf, pad_rel)
        assert self.h

    def max_stack_depth(self, reset=False):
        """Deepest traversal stack (entries) any walk of the host library reached since the last reset (process-wide)."""
        self.L.hs_max_stack_depth.argtypes = [C.c_int]; self.L.hs_max_stack_depth.restype = C.c_int
        return int(self.L.hs_max_stack_depth(1 if reset else 0))

    def close(self):
        if self.h:
            self.L.hs_scene_destroy(self.h); self.h = None

    def bvh_info(self):
        a, b = C.c_int(), C.c_int()
        self.L.hs_bvh_info(self.h, C.byref(a), C.byref(b))
        return dict(n_nodes=a.value, depth=b.value)

    def intersect(self, org, dirs, mode):
        org = np.ascontiguousarray(org, np.float32).reshape(-1, 3); dirs = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        n = len(org)
        ids = np.zeros(n, np.int32); t = np.zeros(n, np.float32); cnt = np.zeros(2, np.uint64)
        self.L.hs_intersect(self.h, A.ptr(org), A.ptr(dirs), n, mode, A.ptr(ids), A.ptr(t), A.ptr(cnt))
        return ids, t, cnt

    def render(self, W, H, spp, spl, eye_depth, light_depth, seed=1, sample_offset=0, sample_stride=1, film_mode=0, max_batch_paths=0, camera=None, factored=False):
        """factored: the MIS weight as k_pool_active / k_connect compute it (walks hoisted out of the pair loop)."""
        cam = np.ascontiguousarray(self.data.camera if camera is None else camera, dtype=A.Camera).reshape(())
        p = A.RenderParams(W, H, spp, spl, eye_depth, light_depth, seed, sample_offset, sample_stride, film_mode, 0, 0, max_batch_paths, 1, 0)
        film = np.zeros((H, W, 3), np.float32); cnt = np.zeros(16, np.uint64)
        (self.L.hs_render_bdpt_factored if factored else self.L.hs_render_bdpt)(self.h, cam.ctypes.data_as(C.c_void_p), C.byref(p), A.ptr(film), A.ptr(cnt))
        names = ["eye_samples", "closest_rays", "shadow_rays", "path_samples", "delta_cap_hits", "nodes", "tris", "pairs", "pretest_mismatch"]
        return film, {k: int(cnt[i]) for i, k in enumerate(names)}

    def render_other(self, integrator, W, H, spp=1, spl=8, eye_depth=8, light_depth=8, seed=1, ppm_radius=0.0, camera=None,
                     sample_offset=0, sample_stride=1, film_mode=0):
        """integrator 1 = pt (eye_depth = max_depth), 2 = ppm."""
        cam = np.ascontiguousarray(self.data.camera if camera is None else camera, dtype=A.Camera).reshape(())
        p = A.RenderParams(W, H, spp, spl, eye_depth, light_depth, seed, sample_offset, sample_stride, film_mode, 0, 0, 0, 1, 0, integrator, ppm_radius)
        film = np.zeros((H, W, 3), np.float32); cnt = np.zeros(16, np.uint64)
        self.L.hs_render_other(self.h, cam.ctypes.data_as(C.c_void_p), C.byref(p), A.ptr(film), A.ptr(cnt))
        names = ["eye_samples", "closest_rays", "shadow_rays", "path_samples", "delta_cap_hits", "nodes", "tris"]
        return film, {k: int(cnt[i]) for i, k in enumerate(names)}

    def render_strategies(self, strategies, W, H, spp=1, eye_depth=8, light_depth=8, seed=1, camera=None, sample_offset=0, sample_stride=1, film_mode=0):
        """strategies 1 = 1/2 nee + 1/2 light tracing, 2 = light tracing, 3 = nee, 4 = every s/t strategy with the balance heuristic."""
        cam = np.ascontiguousarray(self.data.camera if camera is None else camera, dtype=A.Camera).reshape(())
        p = A.RenderParams(W, H, spp, 1, eye_depth, light_depth, seed, sample_offset, sample_stride, film_mode, strategies, 0, 0, 1, 0, 0, 0.0)
        film = np.zeros((H, W, 3), np.float32); cnt = np.zeros(16, np.uint64)
        self.L.hs_render_strategies_c(self.h, cam.ctypes.data_as(C.c_void_p), C.byref(p), A.ptr(film), A.ptr(cnt))
        names = ["eye_samples", "closest_rays", "shadow_rays", "path_samples", "delta_cap_hits", "nodes", "tris"]
        return film, {k: int(cnt[i]) for i, k in enumerate(names)}

    def light_pool(self, spl, light_depth, seed=1):
        p = A.RenderParams(1, 1, 1, spl, 1, light_depth, seed, 0, 1, 0, 0, 0, 0, 0, 0)
        out = np.zeros((len(self.data.lights) * spl * light_depth, 20), np.float32)
        self.L.hs_light_pool(self.h, C.byref(p), A.ptr(out))
        return out
