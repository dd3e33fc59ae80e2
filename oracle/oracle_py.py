"""ctypes bindings for the CHECKERS (test infrastructure only).

Two libraries with the same flat-array interface:
  * ``Port``  -- oracle/liboracle_bdpt.so, the plain-C restatement (oracle/bdpt_oracle.c)
  * ``Ref``   -- oracle/_ref/libref_bdpt.so, the UNMODIFIED reference sources behind oracle/ref_shim.cpp
plus ``RefCuda`` -- oracle/_ref/libref_cuda.so, the reference's existing CUDA integrators (bdpt_cu / pt_cu / ppm_cu)
compiled unmodified for sm_100a behind oracle/ref_cuda_shim.cpp: a same-GPU baseline, not a parity oracle.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "liboracle_bdpt.so")
REF_SO = os.path.join(HERE, "_ref", "libref_bdpt.so")
REF_CUDA_SO = os.path.join(HERE, "_ref", "libref_cuda.so")

_f = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def build(quiet=True):
    """Compile the checkers (the C port always; _ref only where /root/reference exists)."""
    subprocess.run(["make", "-C", HERE, "all"], check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


def _fa(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a if shape is None else a.reshape(shape)


class OrRenderOpts(C.Structure):
    _fields_ = [("rng_mode", C.c_int), ("seed", C.c_uint64), ("nthreads", C.c_int),
                ("max_delta", C.c_int), ("sample_begin", C.c_int)]


class OrStats(C.Structure):
    _fields_ = [("eye_samples", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("path_samples", C.c_uint64), ("delta_cap_hits", C.c_uint64), ("seconds", C.c_double)]


class OrParsed(C.Structure):
    _fields_ = [("n_obj", C.c_int), ("n_lights", C.c_int), ("kind", C.POINTER(C.c_int)), ("group", C.POINTER(C.c_int)),
                ("geom", C.POINTER(C.c_float)), ("mtl", C.POINTER(C.c_float)), ("lights", C.POINTER(C.c_float)),
                ("cam", C.c_float * 10), ("W", C.c_int), ("H", C.c_int)]


class SceneDesc:
    """Flat scene description (what main_cli.cpp:99-141 produces)."""

    def __init__(self, kind, group, geom, mtl, lights, cam=None, W=0, H=0):
        self.kind = np.ascontiguousarray(kind, np.int32)
        self.group = np.ascontiguousarray(group, np.int32)
        self.geom = _fa(geom, (-1, 9))
        self.mtl = _fa(mtl, (-1, 6))
        self.lights = _fa(lights, (-1, 12))
        self.cam = _fa(cam if cam is not None else np.zeros(10))
        self.W, self.H = int(W), int(H)

    @property
    def n_obj(self):
        return len(self.kind)

    @property
    def n_lights(self):
        return len(self.lights)


class _Lib:
    prefix = ""

    def __init__(self, path):
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        self.lib = C.CDLL(path)
        p = self.prefix
        L = self.lib
        g = lambda n: getattr(L, p + n)
        g("scene_create").restype = C.c_void_p
        g("scene_create").argtypes = [C.c_int, _i, _i, _f, _f, C.c_int, _f]
        g("scene_destroy").argtypes = [C.c_void_p]
        g("group_bounds").argtypes = [C.c_void_p, _f, C.c_int]
        g("group_bounds").restype = C.c_int
        g("closest_hit").argtypes = [C.c_void_p, C.c_int, _f, _f, _i, _f, _f, _f, _f]
        g("visibility").argtypes = [C.c_void_p, C.c_int, _f, _f, _f]
        g("bsdf_evaluate").argtypes = [C.c_int, _f, _f, _f, _f, _f]
        g("bsdf_pdf").argtypes = [C.c_int, _f, _f, _f, _f, _f]
        g("bsdf_sample").argtypes = [C.c_int, _f, _f, _f, _f, _f, _f, _f, _f, _i, _f]
        g("mis_weight").argtypes = [_f, C.c_int, C.c_int, _f, C.c_int, C.c_int, _f, C.c_float, _f]
        g("mis_weight").restype = C.c_float
        self._g = g

    # ---- scene ----
    def scene(self, d: SceneDesc):
        return self._g("scene_create")(d.n_obj, d.kind, d.group, d.geom.reshape(-1), d.mtl.reshape(-1),
                                       d.n_lights, d.lights.reshape(-1) if d.n_lights else np.zeros(1, np.float32))

    def destroy(self, h):
        self._g("scene_destroy")(h)

    def group_bounds(self, h):
        out = np.zeros((65536, 7), np.float32)
        n = min(self._g("group_bounds")(h, out.reshape(-1), 65536), 65536)
        return out[:n]

    def closest_hit(self, h, org, dirs):
        org, dirs = _fa(org, (-1, 3)), _fa(dirs, (-1, 3))
        n = len(org)
        ids = np.zeros(n, np.int32); t = np.zeros(n, np.float32)
        pos = np.zeros((n, 3), np.float32); nrm = np.zeros((n, 3), np.float32); mtl = np.zeros((n, 6), np.float32)
        r = self._g("closest_hit")(h, n, org.reshape(-1), dirs.reshape(-1), ids, t, pos.reshape(-1), nrm.reshape(-1), mtl.reshape(-1))
        return dict(id=ids, t=t, pos=pos, normal=nrm, mtl=mtl, bad=r)

    def visibility(self, h, p1, p2):
        p1, p2 = _fa(p1, (-1, 3)), _fa(p2, (-1, 3))
        out = np.zeros(len(p1), np.float32)
        self._g("visibility")(h, len(p1), p1.reshape(-1), p2.reshape(-1), out)
        return out

    # ---- BSDF / MIS ----
    def bsdf_evaluate(self, mtl, wo, wi, N):
        mtl, wo, wi, N = _fa(mtl, (-1, 6)), _fa(wo, (-1, 3)), _fa(wi, (-1, 3)), _fa(N, (-1, 3))
        out = np.zeros((len(mtl), 3), np.float32)
        self._g("bsdf_evaluate")(len(mtl), mtl.reshape(-1), wo.reshape(-1), wi.reshape(-1), N.reshape(-1), out.reshape(-1))
        return out

    def bsdf_pdf(self, mtl, wo, wi, N):
        mtl, wo, wi, N = _fa(mtl, (-1, 6)), _fa(wo, (-1, 3)), _fa(wi, (-1, 3)), _fa(N, (-1, 3))
        out = np.zeros(len(mtl), np.float32)
        self._g("bsdf_pdf")(len(mtl), mtl.reshape(-1), wo.reshape(-1), wi.reshape(-1), N.reshape(-1), out)
        return out

    def bsdf_sample(self, mtl, wo, N, u, cur_eta):
        mtl, wo, N, u = _fa(mtl, (-1, 6)), _fa(wo, (-1, 3)), _fa(N, (-1, 3)), _fa(u, (-1, 3))
        cur_eta = _fa(cur_eta)
        n = len(mtl)
        wi = np.zeros((n, 3), np.float32); f = np.zeros((n, 3), np.float32)
        pdf = np.zeros(n, np.float32); dl = np.zeros(n, np.int32); ne = np.zeros(n, np.float32)
        self._g("bsdf_sample")(n, mtl.reshape(-1), wo.reshape(-1), N.reshape(-1), u.reshape(-1), cur_eta,
                               wi.reshape(-1), f.reshape(-1), pdf, dl, ne)
        return dict(wi=wi, f=f, pdf=pdf, is_delta=dl, new_eta=ne)

    def mis_weight(self, eye17, s_idx, lit20, t_idx, d_vec, dist2, cam_pos):
        eye17, lit20 = _fa(eye17, (-1, 17)), _fa(lit20, (-1, 20))
        return float(self._g("mis_weight")(eye17.reshape(-1), len(eye17), s_idx, lit20.reshape(-1), len(lit20), t_idx,
                                          _fa(d_vec), C.c_float(dist2), _fa(cam_pos)))


class Ref(_Lib):
    """The unmodified reference (oracle/_ref)."""
    prefix = "ref_"

    def __init__(self):
        super().__init__(REF_SO)
        self.lib.ref_render.argtypes = [C.c_void_p, _f, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f]
        self.lib.ref_set_threads.argtypes = [C.c_int]
        self.lib.ref_max_threads.restype = C.c_int

    def render(self, h, cam, eye_depth, light_depth, W, H, spp, spl, nthreads=0):
        if nthreads > 0:
            self.lib.ref_set_threads(nthreads)
        img = np.zeros((H, W, 3), np.float32)
        self.lib.ref_render(h, _fa(cam), eye_depth, light_depth, W, H, spp, spl, img.reshape(-1))
        return img

    def max_threads(self):
        return int(self.lib.ref_max_threads())


class Port(_Lib):
    """The plain-C restatement (oracle/bdpt_oracle.c)."""
    prefix = "or_"
    MT19937, PHILOX = 0, 1

    def __init__(self):
        super().__init__(PORT_SO)
        L = self.lib
        L.or_render.argtypes = [C.c_void_p, _f, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                C.POINTER(OrRenderOpts), _f, C.POINTER(OrStats)]
        L.or_light_pool.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(OrRenderOpts), _f]
        L.or_render_pt.argtypes = [C.c_void_p, _f, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(OrRenderOpts), _f, C.POINTER(OrStats)]
        L.or_render_ppm.argtypes = [C.c_void_p, _f, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                    C.POINTER(OrRenderOpts), _f, C.POINTER(OrStats)]
        L.or_parse_scene_file.restype = C.POINTER(OrParsed)
        L.or_parse_scene_file.argtypes = [C.c_char_p]
        L.or_parsed_free.argtypes = [C.POINTER(OrParsed)]
        L.or_philox4x32_10.argtypes = [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.or_u32_to_unit_float.restype = C.c_float
        L.or_u32_to_unit_float.argtypes = [C.c_uint32]
        L.or_mt19937_floats.argtypes = [C.c_uint32, C.c_int, _f]

    def parse(self, path) -> SceneDesc:
        p = self.lib.or_parse_scene_file(path.encode())
        if not p:
            raise FileNotFoundError(path)
        c = p.contents
        n, nl = c.n_obj, c.n_lights
        d = SceneDesc(np.ctypeslib.as_array(c.kind, (n,)).copy() if n else np.zeros(0, np.int32),
                      np.ctypeslib.as_array(c.group, (n,)).copy() if n else np.zeros(0, np.int32),
                      np.ctypeslib.as_array(c.geom, (n * 9,)).copy() if n else np.zeros(0, np.float32),
                      np.ctypeslib.as_array(c.mtl, (n * 6,)).copy() if n else np.zeros(0, np.float32),
                      np.ctypeslib.as_array(c.lights, (nl * 12,)).copy() if nl else np.zeros(0, np.float32),
                      np.array(list(c.cam), np.float32), c.W, c.H)
        self.lib.or_parsed_free(p)
        return d

    @staticmethod
    def opts(rng_mode=0, seed=1, nthreads=0, max_delta=0, sample_begin=0):
        return OrRenderOpts(rng_mode, seed, nthreads, max_delta, sample_begin)

    def render(self, h, cam, eye_depth, light_depth, W, H, spp, spl, opts=None, want_stats=False):
        img = np.zeros((H, W, 3), np.float32)
        st = OrStats()
        o = opts if opts is not None else self.opts()
        self.lib.or_render(h, _fa(cam), eye_depth, light_depth, W, H, spp, spl, C.byref(o), img.reshape(-1), C.byref(st))
        return (img, st) if want_stats else img

    def render_pt(self, h, cam, max_depth, W, H, spp, opts=None, want_stats=False):
        """cuda_path_trace_kernel (src/pt_cu.cu:20-252) restated on Philox streams (opts.seed / sample_begin / max_delta)."""
        img = np.zeros((H, W, 3), np.float32)
        st = OrStats()
        o = opts if opts is not None else self.opts(rng_mode=1)
        self.lib.or_render_pt(h, _fa(cam), max_depth, W, H, spp, C.byref(o), img.reshape(-1), C.byref(st))
        return (img, st) if want_stats else img

    def render_ppm(self, h, cam, eye_depth, light_depth, W, H, spl, radius=0.0, opts=None, want_stats=False):
        """ppm_render_wrapper (src/ppm_cu.cu:325-398) restated on Philox streams; one pass, spp ignored as in the reference."""
        img = np.zeros((H, W, 3), np.float32)
        st = OrStats()
        o = opts if opts is not None else self.opts(rng_mode=1)
        self.lib.or_render_ppm(h, _fa(cam), eye_depth, light_depth, W, H, spl, float(radius), C.byref(o), img.reshape(-1), C.byref(st))
        return (img, st) if want_stats else img

    def light_pool(self, h, n_lights, light_depth, spl, opts=None):
        out = np.zeros((n_lights * spl * light_depth, 20), np.float32)
        o = opts if opts is not None else self.opts()
        self.lib.or_light_pool(h, light_depth, spl, C.byref(o), out.reshape(-1))
        return out

    def philox(self, ctr, key):
        c = (C.c_uint32 * 4)(*ctr); k = (C.c_uint32 * 2)(*key); o = (C.c_uint32 * 4)()
        self.lib.or_philox4x32_10(c, k, o)
        return list(o)

    def unit_float(self, x):
        return float(self.lib.or_u32_to_unit_float(x))

    def mt19937_floats(self, seed, n):
        out = np.zeros(n, np.float32)
        self.lib.or_mt19937_floats(seed, n, out)
        return out


class RefCuda:
    """The reference's own CUDA integrators on this GPU (needs a device; loads libcudart)."""
    BDPT, PT, PPM = 0, 1, 2

    def __init__(self):
        if not os.path.exists(REF_CUDA_SO):
            raise FileNotFoundError(REF_CUDA_SO)
        self.lib = C.CDLL(REF_CUDA_SO)
        self.lib.refcuda_scene_create.restype = C.c_void_p
        self.lib.refcuda_scene_create.argtypes = [C.c_int, _i, _i, _f, _f, C.c_int, _f]
        self.lib.refcuda_scene_destroy.argtypes = [C.c_void_p]
        self.lib.refcuda_render.restype = C.c_double
        self.lib.refcuda_render.argtypes = [C.c_void_p, C.c_int, _f, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f, C.c_int]

    def scene(self, d: SceneDesc):
        return self.lib.refcuda_scene_create(d.n_obj, d.kind, d.group, d.geom.reshape(-1), d.mtl.reshape(-1),
                                             d.n_lights, d.lights.reshape(-1) if d.n_lights else np.zeros(1, np.float32))

    def destroy(self, h):
        self.lib.refcuda_scene_destroy(h)

    @staticmethod
    def camera12(cam10, W, H):
        """eye, UL, dx, dy the way main_cli.cpp:25-40 (init_camera) derives them, in float32."""
        f = np.float32
        c = np.asarray(cam10, np.float32)
        eye, look, up, fov = c[0:3], c[3:6], c[6:9], c[9]
        half_h = f(np.tan(f(f(fov * f(np.pi)) / f(180.0)) / f(2.0)))
        half_w = f(f(W) / f(H)) * half_h
        nrm = lambda v: (v / np.sqrt(np.dot(v, v), dtype=np.float32)).astype(np.float32)
        w = nrm(eye - look)
        u = nrm(np.cross(up, w).astype(np.float32))
        v = np.cross(w, u).astype(np.float32)
        UL = eye - half_w * u + half_h * v - w
        dx = (f(2) * half_w * u) / f(W)
        dy = (f(-2) * half_h * v) / f(H)
        return np.concatenate([eye, UL, dx, dy]).astype(np.float32)

    def render(self, h, mode, cam10, eye_depth, light_depth, W, H, spp, spl, quiet=True):
        """-> (film[H,W,3] mean radiance, wall milliseconds of run_cuda_* as main_cli.cpp:207-219 times it)"""
        img = np.zeros((H, W, 3), np.float32)
        ms = self.lib.refcuda_render(h, mode, self.camera12(cam10, W, H), W, H, spp, spl, eye_depth, light_depth, img.reshape(-1), 1 if quiet else 0)
        if ms < 0:
            raise ValueError("bad mode")
        return img, float(ms)


def have_ref_cuda():
    return os.path.exists(REF_CUDA_SO)


def have_ref():
    return os.path.exists(REF_SO)
