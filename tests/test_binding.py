"""The reference-side binding of INTEGRATION.md section 1 as COMPILED code: examples/bdpt_b200_helper.cpp (move_data_to_cuda_bdpt /
run_cuda_bdpt of include/bdpt_cu_helper.h:5-6 over libb2pt.so) built against the reference's own headers and object.cpp into
oracle/_ref/libbinding_b200.so, driven the way main_cli.cpp:99-141,200-213 drives it (oracle/binding_shim.cpp builds the
std::map<int,AABB> groups with the reference's own add_obj)."""
import ctypes as C
import os

import numpy as np
import pytest

from path_tracing_b200 import _abi as A, api
from path_tracing_b200.scene import reference_scene

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "..", "oracle", "_ref", "libbinding_b200.so")
needs_binding = pytest.mark.skipif(not os.path.exists(SO), reason="oracle/_ref/libbinding_b200.so not built (needs /root/reference at build time)")
_f = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def load():
    A.lib()                                    # libb2pt.so first: the binding links to it
    L = C.CDLL(SO)
    L.binding_render.argtypes = [C.c_int, _i, _i, _f, _f, C.c_int, _f, _f, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f]
    L.binding_last_error.restype = C.c_char_p
    return L


def render_through_binding(L, sd, W, H, spp, spl, eye_depth, light_depth):
    f = sd.to_flat()
    img = np.zeros((H, W, 3), np.float32)
    rc = L.binding_render(len(f["kind"]), f["kind"], f["group"], np.ascontiguousarray(f["geom"].reshape(-1)), np.ascontiguousarray(f["mtl"].reshape(-1)),
                          len(f["lights"]), np.ascontiguousarray(f["lights"].reshape(-1)), f["cam"], eye_depth, light_depth, W, H, spp, spl, img.reshape(-1))
    return rc, img


@needs_binding
def test_binding_exports_the_reference_entry_points_and_has_no_cpu_fallback():
    L = load()
    for sym in ("_Z22move_data_to_cuda_bdptSt3mapIi4AABBSt4lessIiESaISt4pairIKiS0_EEERSt6vectorI9CudaLightSaIS9_EEi", "binding_render"):
        assert hasattr(L, sym), sym                        # the mangled name main_cli.cpp links against (include/bdpt_cu_helper.h:5)
    if api.device_count() == 0:
        rc, _ = render_through_binding(L, reference_scene("cornell"), 8, 8, 1, 1, 2, 2)
        assert rc != 0 and b"device" in L.binding_last_error().lower()


@pytest.mark.gpu
@needs_binding
@pytest.mark.parametrize("name", ["cornell", "mis_variant"])
def test_binding_film_equals_the_python_surface(name):
    """Same scene, same parameters, same seed (the binding fixes seed 1): run_cuda_bdpt through the compiled binding returns the film
    Scene.render_bdpt returns -- counters are stream-exact, the film up to the order of the float atomics."""
    L = load()
    sd = reference_scene(name)
    W, H, spp, spl, de, dl = 96, 64, 8, 8, 6, 5
    rc, img = render_through_binding(L, sd, W, H, spp, spl, de, dl)
    assert rc == 0, L.binding_last_error()
    sc = api.Scene(sd)
    film, _ = sc.render_bdpt(width=W, height=H, spp=spp, spl=spl, eye_depth=de, light_depth=dl, seed=1)
    sc.close()
    assert img.mean() > 1e-4 and np.isfinite(img).all()
    assert np.allclose(img, film, rtol=1e-4, atol=1e-6), float(np.abs(img - film).max())
