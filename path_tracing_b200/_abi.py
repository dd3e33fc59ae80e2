"""ctypes view of include/b2pt.h: POD layouts (numpy structured dtypes) and the library loader.

The product is libb2pt.so (hand-written CUDA for sm_100a behind a C ABI).  There is no CPU fallback:
if the library is missing, loading raises; if no CUDA device is present, compute calls return
B2_ERR_NO_DEVICE and the wrappers raise.
"""
import ctypes as C
import os
import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "libb2pt.so")

Material = np.dtype([("base_color", np.float32, 3), ("roughness", np.float32), ("metallic", np.float32), ("eta", np.float32)])
Tri = np.dtype([("v0", np.float32, 3), ("v1", np.float32, 3), ("v2", np.float32, 3),
                ("obj_id", np.int32), ("group", np.int32), ("material", np.int32)])
Sphere = np.dtype([("center", np.float32, 3), ("r", np.float32), ("obj_id", np.int32), ("group", np.int32), ("material", np.int32)])
Light = np.dtype([("pos", np.float32, 3), ("dir", np.float32, 3), ("illum", np.float32, 3),
                  ("cutoff", np.float32), ("is_parallel", np.int32), ("ball_r", np.float32)])
Camera = np.dtype([("eye", np.float32, 3), ("look_at", np.float32, 3), ("view_up", np.float32, 3), ("fov", np.float32)])
assert Material.itemsize == 24 and Tri.itemsize == 48 and Sphere.itemsize == 28 and Light.itemsize == 48 and Camera.itemsize == 40


class BvhParams(C.Structure):
    _fields_ = [("max_leaf_tris", C.c_int32), ("pad_rel", C.c_float), ("builder", C.c_int32), ("ploc_radius", C.c_int32)]


class BvhInfo(C.Structure):
    _fields_ = [("n_tris", C.c_uint64), ("n_nodes", C.c_uint64), ("node_bytes", C.c_uint64), ("tri_bytes", C.c_uint64),
                ("depth", C.c_int32), ("build_ms", C.c_float), ("sah_cost", C.c_float), ("build_passes", C.c_int32), ("reserved", C.c_int32)]


class RenderParams(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp", C.c_int32), ("spl", C.c_int32),
                ("eye_depth", C.c_int32), ("light_depth", C.c_int32), ("seed", C.c_uint64),
                ("sample_offset", C.c_int32), ("sample_stride", C.c_int32), ("film_mode", C.c_int32),
                ("strategies", C.c_int32), ("max_delta", C.c_int32), ("max_batch_paths", C.c_int32),
                ("count_traversal", C.c_int32), ("profile_stages", C.c_int32),
                ("integrator", C.c_int32), ("ppm_radius", C.c_float), ("reserved", C.c_int32 * 2)]


class Stats(C.Structure):
    _fields_ = [("eye_samples", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("path_samples", C.c_uint64), ("delta_cap_hits", C.c_uint64), ("nodes_visited", C.c_uint64),
                ("tris_tested", C.c_uint64), ("n_batches", C.c_uint64), ("n_waves", C.c_uint64), ("n_launches", C.c_uint64), ("pairs_tested", C.c_uint64), ("eye_vertices", C.c_uint64),
                ("ms_total", C.c_float), ("ms_light_pool", C.c_float), ("ms_primary", C.c_float), ("ms_extend", C.c_float),
                ("ms_connect", C.c_float), ("ms_reserved", C.c_float * 2)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_ if n != "ms_reserved"}


class ParsedScene(C.Structure):
    _fields_ = [("n_tris", C.c_size_t), ("n_spheres", C.c_size_t), ("n_materials", C.c_size_t), ("n_lights", C.c_size_t),
                ("tris", C.c_void_p), ("spheres", C.c_void_p), ("materials", C.c_void_p), ("lights", C.c_void_p),
                ("camera", C.c_float * 10), ("width", C.c_int32), ("height", C.c_int32)]


# every symbol include/b2pt.h declares (tests check that the built library exports all of them)
class ObjOptions(C.Structure):
    _fields_ = [("first_obj_id", C.c_int32), ("first_group", C.c_int32), ("group_grid", C.c_int32),
                ("default_color", C.c_float * 3), ("default_roughness", C.c_float), ("reader", C.c_int32), ("reserved", C.c_int32 * 3)]


PARSE_OBJ_DIRECTIVE = 1

SYMBOLS = [
    "b2_last_error", "b2_device_count", "b2_version", "b2_scene_create", "b2_scene_build_bvh", "b2_scene_bvh_info",
    "b2_scene_destroy", "b2_intersect", "b2_intersect_device", "b2_intersect_counters", "b2_render_bdpt",
    "b2_render_bdpt_device", "b2_render_bdpt_multi", "b2_multi_init", "b2_light_pool", "b2_light_pool_device_noalloc", "b2_bsdf_evaluate", "b2_bsdf_pdf", "b2_bsdf_sample", "b2_mis_weight",
    "b2_philox", "b2_parse_scene_file", "b2_parsed_scene_free", "b2_film_to_rgb8", "b2_write_image",
    "b2_obj_default_options", "b2_obj_has_tinyobj", "b2_load_obj", "b2_parse_scene_file_ex", "b2_render_bdpt_progressive", "b2_frame_rms",
    "b2_measure_fp32_peak", "b2_measure_l2_bandwidth",
]

_lib = None


class B2Error(RuntimeError):
    pass


def lib():
    """Load libb2pt.so (built in-tree by __graft_entry__.build()).  Fails loudly if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B2Error(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(the CUDA extension is required; there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.b2_last_error.restype = C.c_char_p
        L.b2_version.restype = C.c_char_p
        vp, sz, i32 = C.c_void_p, C.c_size_t, C.c_int
        L.b2_device_count.argtypes = [C.POINTER(C.c_int)]
        L.b2_scene_create.argtypes = [vp, sz, vp, sz, vp, sz, vp, sz, i32, C.POINTER(vp)]
        L.b2_scene_build_bvh.argtypes = [vp, C.POINTER(BvhParams)]
        L.b2_scene_bvh_info.argtypes = [vp, C.POINTER(BvhInfo)]
        L.b2_scene_destroy.argtypes = [vp]
        L.b2_intersect.argtypes = [vp, vp, vp, sz, i32, vp, vp]
        L.b2_intersect_device.argtypes = [vp, vp, vp, sz, i32, vp, vp, vp]
        L.b2_intersect_counters.argtypes = [vp, C.POINTER(C.c_uint64 * 3), i32]
        L.b2_render_bdpt.argtypes = [vp, vp, C.POINTER(RenderParams), vp, C.POINTER(Stats)]
        L.b2_render_bdpt_device.argtypes = [vp, vp, C.POINTER(RenderParams), vp, vp, C.POINTER(Stats)]
        L.b2_render_bdpt_multi.argtypes = [C.POINTER(vp), i32, vp, C.POINTER(RenderParams), vp, C.POINTER(Stats)]
        L.b2_multi_init.argtypes = [C.POINTER(i32), i32]
        L.b2_light_pool.argtypes = [vp, C.POINTER(RenderParams), vp]
        L.b2_light_pool_device_noalloc.argtypes = [vp, C.POINTER(RenderParams), vp, vp]
        L.b2_bsdf_evaluate.argtypes = [sz, vp, vp, vp, vp, vp]
        L.b2_bsdf_pdf.argtypes = [sz, vp, vp, vp, vp, vp]
        L.b2_bsdf_sample.argtypes = [sz, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
        L.b2_mis_weight.argtypes = [sz, vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_int32, vp]
        L.b2_philox.argtypes = [sz, vp, vp, vp, vp]
        L.b2_parse_scene_file.argtypes = [C.c_char_p, C.POINTER(C.POINTER(ParsedScene))]
        L.b2_parsed_scene_free.argtypes = [C.POINTER(ParsedScene)]
        L.b2_parsed_scene_free.restype = None
        L.b2_obj_default_options.argtypes = [C.POINTER(ObjOptions)]
        L.b2_obj_default_options.restype = None
        L.b2_load_obj.argtypes = [C.c_char_p, C.POINTER(ObjOptions), C.POINTER(C.POINTER(ParsedScene))]
        L.b2_render_bdpt_progressive.argtypes = [vp, vp, vp, C.c_int32, vp, vp, vp, vp, vp]
        L.b2_frame_rms.argtypes = [vp, vp, C.c_size_t, C.POINTER(C.c_float)]
        L.b2_parse_scene_file_ex.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.POINTER(ParsedScene))]
        L.b2_film_to_rgb8.argtypes = [vp, C.c_int32, C.c_int32, vp]
        L.b2_write_image.argtypes = [C.c_char_p, vp, C.c_int32, C.c_int32]
        L.b2_measure_fp32_peak.argtypes = [i32, C.POINTER(C.c_float)]
        L.b2_measure_l2_bandwidth.argtypes = [i32, sz, C.POINTER(C.c_float)]
        L.b2_obj_has_tinyobj.restype = C.c_int
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise B2Error(f"b2pt error {rc}: {lib().b2_last_error().decode(errors='replace')}")


def ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None
