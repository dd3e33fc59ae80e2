"""Python host side of the C ABI (include/b2pt.h).  Mirrors the reference's call surface for the BDPT path:

    move_data_to_cuda_bdpt(groups, lights, spl)   -> Scene(scene_data)             (include/bdpt_cu_helper.h:5)
    run_cuda_bdpt / run_cpu_bdpt(camera, ..., image, eye_depth, light_depth, W, H, spp, spl)
                                                  -> Scene.render_bdpt(...)        (include/cpu_bdpt.h:9-17)
    cpu_find_closest_hit / cpu_check_visibility   -> Scene.closest_hit / Scene.visibility   (src/cpu_bdpt.cpp:30,82)

Everything computes in libb2pt.so (hand-written CUDA for sm_100a); there is no CPU fallback here.
"""
import ctypes as C
import os
import numpy as np
from . import _abi as A
from .scene import SceneData

CLOSEST, ANY, BRUTE_FORCE, COUNT_VISITS, FAST_SHADOW, KEEP_RAY_ORDER = 0, 1, 16, 32, 64, 128
FILM_MEAN, FILM_SUM = 0, 1
STRATEGIES_REFERENCE, STRATEGIES_FULL, STRATEGIES_LIGHT_TRACING, STRATEGIES_NEE, STRATEGIES_MIS = 0, 1, 2, 3, 4
NOALLOC_MAX_POOL_VERTICES = 32768                            # B2_NOALLOC_MAX_POOL_VERTICES
INTEGRATOR_BDPT, INTEGRATOR_PT, INTEGRATOR_PPM = 0, 1, 2      # main_cli's --mode (src/main_cli.cpp:201-213)


def device_count():
    n = C.c_int(0)
    A.check(A.lib().b2_device_count(C.byref(n)))
    return n.value


def measure_fp32_peak(device=0):
    """Non-tensor FP32 FMA peak of `device` in TFLOP/s (in-run microbenchmark)."""
    v = C.c_float(0)
    A.check(A.lib().b2_measure_fp32_peak(int(device), C.byref(v)))
    return v.value


def measure_l2_bandwidth(device=0, nbytes=0):
    """Read bandwidth of an L2-resident buffer in GB/s (the roof of traversals whose nodes + triangles fit in L2)."""
    v = C.c_float(0)
    A.check(A.lib().b2_measure_l2_bandwidth(int(device), int(nbytes), C.byref(v)))
    return v.value


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a if shape is None else a.reshape(shape)


class Scene:
    """Device-resident scene + wide BVH (opaque B2Scene handle)."""

    def __init__(self, data: SceneData, device=0, max_leaf_tris=0, pad_rel=0.0, build=True, builder=None, ploc_radius=0):
        if builder is None:                                    # A/B runs of the tools: B2_BVH_BUILDER=1 selects the radix-tree builder
            builder = int(os.environ.get("B2_BVH_BUILDER", "0"))
        if not ploc_radius:
            ploc_radius = int(os.environ.get("B2_PLOC_RADIUS", "0"))
        self.data = data
        self._h = C.c_void_p()
        L = A.lib()
        A.check(L.b2_scene_create(A.ptr(data.tris), len(data.tris), A.ptr(data.spheres), len(data.spheres),
                                  A.ptr(data.materials), len(data.materials), A.ptr(data.lights), len(data.lights),
                                  int(device), C.byref(self._h)))
        if build:
            self.build_bvh(max_leaf_tris, pad_rel, builder, ploc_radius)

    def build_bvh(self, max_leaf_tris=0, pad_rel=0.0, builder=0, ploc_radius=0):
        """builder 0 = PLOC (default), 1 = LBVH (B2_BUILDER_*)."""
        p = A.BvhParams(int(max_leaf_tris), float(pad_rel), int(builder), int(ploc_radius))
        A.check(A.lib().b2_scene_build_bvh(self._h, C.byref(p)))

    def bvh_info(self):
        info = A.BvhInfo()
        A.check(A.lib().b2_scene_bvh_info(self._h, C.byref(info)))
        return {n: getattr(info, n) for n, _ in info._fields_}

    def close(self):
        if self._h:
            A.lib().b2_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- ray queries ----
    def closest_hit(self, org, dirs, brute_force=False, count=False, keep_order=False):
        org, dirs = _f32(org, (-1, 3)), _f32(dirs, (-1, 3))
        n = len(org)
        ids = np.full(n, -1, np.int32)
        t = np.full(n, 1e20, np.float32)
        mode = CLOSEST | (BRUTE_FORCE if brute_force else 0) | (COUNT_VISITS if count else 0) | (KEEP_RAY_ORDER if keep_order else 0)
        A.check(A.lib().b2_intersect(self._h, A.ptr(org), A.ptr(dirs), n, mode, A.ptr(ids), A.ptr(t)))
        return ids, t

    def visibility(self, p1, p2, brute_force=False, count=False, fast_shadow=False, keep_order=False):
        p1, p2 = _f32(p1, (-1, 3)), _f32(p2, (-1, 3))
        n = len(p1)
        vis = np.zeros(n, np.int32)
        mode = ANY | (BRUTE_FORCE if brute_force else 0) | (COUNT_VISITS if count else 0) | (FAST_SHADOW if fast_shadow else 0) | (KEEP_RAY_ORDER if keep_order else 0)
        A.check(A.lib().b2_intersect(self._h, A.ptr(p1), A.ptr(p2), n, mode, A.ptr(vis), None))
        return vis

    def intersect_device(self, d_org, d_dir, n, mode, d_prim_id, d_t=0, stream=0):
        """Raw device pointers (ints), asynchronous on `stream`."""
        A.check(A.lib().b2_intersect_device(self._h, C.c_void_p(d_org), C.c_void_p(d_dir), n, mode, C.c_void_p(d_prim_id),
                                            C.c_void_p(d_t) if d_t else None, C.c_void_p(stream) if stream else None))

    def intersect_counters(self, reset=True):
        out = (C.c_uint64 * 3)()
        A.check(A.lib().b2_intersect_counters(self._h, C.byref(out), 1 if reset else 0))
        return dict(rays=out[0], nodes=out[1], tris=out[2])

    # ---- render ----
    def _params(self, width, height, spp, spl, eye_depth, light_depth, seed, sample_offset, sample_stride, film_mode,
                max_delta, max_batch_paths, count_traversal, profile_stages=False, strategies=0, integrator=0, ppm_radius=0.0):
        return A.RenderParams(int(width), int(height), int(spp), int(spl), int(eye_depth), int(light_depth), int(seed),
                              int(sample_offset), int(sample_stride), int(film_mode), int(strategies), int(max_delta), int(max_batch_paths),
                              1 if count_traversal else 0, 1 if profile_stages else 0, int(integrator), float(ppm_radius))

    def render_bdpt(self, camera=None, width=None, height=None, spp=8, spl=8, eye_depth=4, light_depth=4, seed=1,
                    sample_offset=0, sample_stride=1, film_mode=FILM_MEAN, max_delta=0, max_batch_paths=0,
                    count_traversal=False, profile_stages=False, out=None, device_ptr=None, stream=0, strategies=STRATEGIES_REFERENCE,
                    integrator=INTEGRATOR_BDPT, ppm_radius=0.0):
        """Returns (film[H,W,3] float32, stats dict).  With device_ptr (int) the film stays on the device and None is returned for it."""
        cam = np.ascontiguousarray(self.data.camera if camera is None else camera, dtype=A.Camera).reshape(())
        W = int(width if width else self.data.width)
        H = int(height if height else self.data.height)
        p = self._params(W, H, spp, spl, eye_depth, light_depth, seed, sample_offset, sample_stride, film_mode, max_delta,
                         max_batch_paths, count_traversal, profile_stages, strategies, integrator, ppm_radius)
        st = A.Stats()
        if device_ptr is not None:
            A.check(A.lib().b2_render_bdpt_device(self._h, cam.ctypes.data_as(C.c_void_p), C.byref(p), C.c_void_p(int(device_ptr)),
                                                  C.c_void_p(stream) if stream else None, C.byref(st)))
            return None, st.as_dict()
        film = out if out is not None else np.zeros((max(H, 0), max(W, 0), 3), np.float32)
        assert film.dtype == np.float32 and film.size == 3 * max(W, 0) * max(H, 0) and film.flags["C_CONTIGUOUS"]
        A.check(A.lib().b2_render_bdpt(self._h, cam.ctypes.data_as(C.c_void_p), C.byref(p), A.ptr(film), C.byref(st)))
        return film, st.as_dict()

    def render_pt(self, max_depth=8, **kw):
        """run_cuda_pt (include/pt_cu_helper.h:6): path tracing with next-event estimation, max_depth non-delta bounces."""
        kw.pop("integrator", None)
        return self.render_bdpt(eye_depth=max_depth, integrator=INTEGRATOR_PT, **kw)

    def render_ppm(self, spl=8, eye_depth=8, light_depth=8, ppm_radius=0.0, **kw):
        """run_cuda_ppm (include/ppm_cu_helper.h): one photon-mapping pass of n_lights * spl photons; spp is ignored."""
        kw.pop("integrator", None)
        return self.render_bdpt(spl=spl, eye_depth=eye_depth, light_depth=light_depth, integrator=INTEGRATOR_PPM, ppm_radius=ppm_radius, **kw)

    def render_progressive(self, passes, camera=None, width=None, height=None, spp=8, spl=8, eye_depth=4, light_depth=4, seed=1,
                           on_pass=None, want_rms=True, integrator=INTEGRATOR_BDPT):
        """The viewer's progressive accumulation (src/main.cpp:399-531): `passes` renders of `spp` samples (seed + pass)
        averaged.  Returns (film, rms[passes] or None, stats).  on_pass(pass, film_avg[H,W,3]) is called after every pass."""
        cam = np.ascontiguousarray(self.data.camera if camera is None else camera, dtype=A.Camera).reshape(())
        W = int(width if width else self.data.width); H = int(height if height else self.data.height)
        p = self._params(W, H, spp, spl, eye_depth, light_depth, seed, 0, 1, FILM_MEAN, 0, 0, False, integrator=integrator)
        st = A.Stats()
        film = np.zeros((max(H, 0), max(W, 0), 3), np.float32)
        rms = np.zeros(max(int(passes), 0), np.float32) if want_rms else None
        CB = C.CFUNCTYPE(None, C.c_int32, C.POINTER(C.c_float), C.c_int32, C.c_int32, C.c_void_p)
        cb = CB(lambda k, ptr, w, h, user: on_pass(int(k), np.ctypeslib.as_array(ptr, shape=(h, w, 3)).copy())) if on_pass else None
        A.check(A.lib().b2_render_bdpt_progressive(self._h, cam.ctypes.data_as(C.c_void_p), C.byref(p), int(passes), A.ptr(film),
                                                   A.ptr(rms) if rms is not None else None, C.cast(cb, C.c_void_p) if cb else None, None, C.byref(st)))
        return film, rms, st.as_dict()

    def light_pool_device_noalloc(self, device_ptr, spl=8, light_depth=4, seed=1, max_delta=0, stream=0):
        """b2_light_pool_device_noalloc: the pool into a caller-owned DEVICE buffer (int pointer, n_lights*spl*light_depth x 20
        floats); no allocation inside the library (static device scratch, at most NOALLOC_MAX_POOL_VERTICES vertices)."""
        p = self._params(1, 1, 1, spl, 1, light_depth, seed, 0, 1, 0, max_delta, 0, False)
        A.check(A.lib().b2_light_pool_device_noalloc(self._h, C.byref(p), C.c_void_p(int(device_ptr)), C.c_void_p(stream) if stream else None))

    def light_pool(self, spl=8, light_depth=4, seed=1, max_delta=0):
        n = len(self.data.lights) * spl * light_depth
        out = np.zeros((n, 20), np.float32)
        p = self._params(1, 1, 1, spl, 1, light_depth, seed, 0, 1, 0, max_delta, 0, False)
        A.check(A.lib().b2_light_pool(self._h, C.byref(p), A.ptr(out)))
        return out


def multi_init(devices):
    """NCCL communicators for `devices` ahead of the first render_bdpt_multi (b2_multi_init): seconds on an 8-GPU node."""
    arr = (C.c_int32 * len(devices))(*[int(d) for d in devices])
    A.check(A.lib().b2_multi_init(arr, len(devices)))


def render_bdpt_multi(scenes, camera=None, width=None, height=None, spp=8, spl=8, eye_depth=4, light_depth=4, seed=1,
                      film_mode=FILM_MEAN, max_batch_paths=0, integrator=INTEGRATOR_BDPT):
    """One process, len(scenes) GPUs (b2_render_bdpt_multi): scenes[g] is the same SceneData created on device g."""
    s0 = scenes[0]
    cam = np.ascontiguousarray(s0.data.camera if camera is None else camera, dtype=A.Camera).reshape(())
    W = int(width if width else s0.data.width); H = int(height if height else s0.data.height)
    p = s0._params(W, H, spp, spl, eye_depth, light_depth, seed, 0, 1, film_mode, 0, max_batch_paths, False, integrator=integrator)
    st = A.Stats()
    film = np.zeros((max(H, 0), max(W, 0), 3), np.float32)
    handles = (C.c_void_p * len(scenes))(*[s._h for s in scenes])
    A.check(A.lib().b2_render_bdpt_multi(handles, len(scenes), cam.ctypes.data_as(C.c_void_p), C.byref(p), A.ptr(film), C.byref(st)))
    return film, st.as_dict()


# ---- BSDF / MIS / RNG tables (device kernels over host arrays) ----
def bsdf_evaluate(mtl, wo, wi, N):
    mtl, wo, wi, N = _f32(mtl, (-1, 6)), _f32(wo, (-1, 3)), _f32(wi, (-1, 3)), _f32(N, (-1, 3))
    out = np.zeros((len(mtl), 3), np.float32)
    A.check(A.lib().b2_bsdf_evaluate(len(mtl), A.ptr(mtl), A.ptr(wo), A.ptr(wi), A.ptr(N), A.ptr(out)))
    return out


def bsdf_pdf(mtl, wo, wi, N):
    mtl, wo, wi, N = _f32(mtl, (-1, 6)), _f32(wo, (-1, 3)), _f32(wi, (-1, 3)), _f32(N, (-1, 3))
    out = np.zeros(len(mtl), np.float32)
    A.check(A.lib().b2_bsdf_pdf(len(mtl), A.ptr(mtl), A.ptr(wo), A.ptr(wi), A.ptr(N), A.ptr(out)))
    return out


def bsdf_sample(mtl, wo, N, u, cur_eta):
    mtl, wo, N, u, cur_eta = _f32(mtl, (-1, 6)), _f32(wo, (-1, 3)), _f32(N, (-1, 3)), _f32(u, (-1, 3)), _f32(cur_eta)
    n = len(mtl)
    wi = np.zeros((n, 3), np.float32); f = np.zeros((n, 3), np.float32)
    pdf = np.zeros(n, np.float32); dl = np.zeros(n, np.int32); ne = np.zeros(n, np.float32)
    A.check(A.lib().b2_bsdf_sample(n, A.ptr(mtl), A.ptr(wo), A.ptr(N), A.ptr(u), A.ptr(cur_eta), A.ptr(wi), A.ptr(f), A.ptr(pdf), A.ptr(dl), A.ptr(ne)))
    return dict(wi=wi, f=f, pdf=pdf, is_delta=dl, new_eta=ne)


def mis_weight(eye17, n_eye, s_idx, lit20, n_lit, t_idx, d_vec, dist2, cam_pos, max_verts):
    """eye17: [n, max_verts, 17], lit20: [n, max_verts, 20]."""
    eye17, lit20 = _f32(eye17, (-1, max_verts, 17)), _f32(lit20, (-1, max_verts, 20))
    n = len(eye17)
    i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)
    n_eye, s_idx, n_lit, t_idx = i32(n_eye), i32(s_idx), i32(n_lit), i32(t_idx)
    d_vec, dist2, cam_pos = _f32(d_vec, (-1, 3)), _f32(dist2), _f32(cam_pos, (-1, 3))
    out = np.zeros(n, np.float32)
    A.check(A.lib().b2_mis_weight(n, A.ptr(eye17), A.ptr(n_eye), A.ptr(s_idx), A.ptr(lit20), A.ptr(n_lit), A.ptr(t_idx),
                                  A.ptr(d_vec), A.ptr(dist2), A.ptr(cam_pos), int(max_verts), A.ptr(out)))
    return out


def philox(ctr4, key2):
    ctr4 = np.ascontiguousarray(ctr4, dtype=np.uint32).reshape(-1, 4)
    key2 = np.ascontiguousarray(key2, dtype=np.uint32).reshape(-1, 2)
    out = np.zeros_like(ctr4); unit = np.zeros(ctr4.shape, np.float32)
    A.check(A.lib().b2_philox(len(ctr4), A.ptr(ctr4), A.ptr(key2), A.ptr(out), A.ptr(unit)))
    return out, unit


# ---- film output (main_cli.cpp:221-254) ----
def film_to_rgb8(film):
    film = _f32(film)
    H, W = film.shape[0], film.shape[1]
    out = np.zeros((H, W, 3), np.uint8)
    A.check(A.lib().b2_film_to_rgb8(A.ptr(film), W, H, A.ptr(out)))
    return out


def frame_rms(cur_rgb8, last_rgb8):
    """sqrt(sum (cur - last)^2) / 255 over all bytes: the viewer's convergence curve (main.cpp:502-531)."""
    a = np.ascontiguousarray(cur_rgb8, np.uint8); b = np.ascontiguousarray(last_rgb8, np.uint8)
    assert a.size == b.size
    out = C.c_float(0)
    A.check(A.lib().b2_frame_rms(A.ptr(a), A.ptr(b), a.size, C.byref(out)))
    return out.value


def write_image(path, film):
    film = _f32(film)
    A.check(A.lib().b2_write_image(str(path).encode(), A.ptr(film), film.shape[1], film.shape[0]))
