"""b2_light_pool_device_noalloc (include/b2pt.h): the light-vertex pool through the entry point that allocates nothing
(static device scratch), against the existing b2_light_pool, through the C ABI.  Collected last on purpose.
Passes alone in a fresh process and after the other tests: it owns its scene handle, and the static
array is cleared by every call."""
import numpy as np
import pytest

from path_tracing_b200 import api
from path_tracing_b200.scene import reference_scene

pytestmark = pytest.mark.gpu


def _pool_via_noalloc(sc, n, out_t, **kw):
    out_t.fill_(float("nan"))                                   # whatever was there must be overwritten
    sc.light_pool_device_noalloc(out_t.data_ptr(), **kw)        # NULL stream: the scene's own stream, blocking
    return out_t.cpu().numpy().reshape(n, 20)


def test_noalloc_light_pool_equals_light_pool_on_two_calls_in_a_row():
    import torch
    sd = reference_scene("cornell")
    sc = api.Scene(sd)
    try:
        n_l = len(sd.lights)
        # the caller owns inputs and outputs: one device buffer large enough for both calls, made BEFORE the calls
        out_t = torch.empty(n_l * 64 * 8 * 20, dtype=torch.float32, device="cuda")
        # call 1: a large pool; call 2, straight after, a SMALLER one with another seed -- it finds call 1's records in the
        # shared static array and must not show any of them (zero vertices of terminated subpaths stay zero)
        for call, kw in enumerate((dict(spl=64, light_depth=8, seed=11), dict(spl=8, light_depth=5, seed=3))):
            n = n_l * kw["spl"] * kw["light_depth"]
            ref = sc.light_pool(**kw)                            # existing entry point (workspace scratch, host output)
            free0 = torch.cuda.mem_get_info()[0]
            got = _pool_via_noalloc(sc, n, out_t[: n * 20], **kw)
            if call > 0:      # (the first call of a process may load the module lazily, which maps the static array itself)
                assert torch.cuda.mem_get_info()[0] == free0     # nothing was allocated on the device by the call
            assert np.isfinite(got).all()
            # same kernel body, same Philox streams: bit-identical (the tolerance of the pool tests is 2e-4; equality is stricter)
            assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), float(np.abs(got - ref).max())
            assert (np.abs(ref[:, 6:9]).sum(1) > 0).sum() > n // 4      # the comparison is not between two empty pools
    finally:
        sc.close()


def test_noalloc_light_pool_on_a_caller_stream_and_limits():
    import torch
    sd = reference_scene("mis")
    sc = api.Scene(sd)
    try:
        kw = dict(spl=16, light_depth=6, seed=5)
        n = len(sd.lights) * kw["spl"] * kw["light_depth"]
        out_t = torch.full((n * 20,), float("nan"), dtype=torch.float32, device="cuda")
        st = torch.cuda.Stream()
        sc.light_pool_device_noalloc(out_t.data_ptr(), stream=st.cuda_stream, **kw)     # asynchronous on the caller's stream
        st.synchronize()
        ref = sc.light_pool(**kw)
        assert np.array_equal(out_t.cpu().numpy().reshape(n, 20).view(np.uint32), ref.view(np.uint32))
        # beyond the static scratch: a clear error, nothing written
        too_many = api.NOALLOC_MAX_POOL_VERTICES // (len(sd.lights) * 8) + 1
        with pytest.raises(RuntimeError, match="exceeds the static scratch"):
            sc.light_pool_device_noalloc(out_t.data_ptr(), spl=too_many, light_depth=8)
        with pytest.raises(RuntimeError, match="NULL"):
            sc.light_pool_device_noalloc(0, **kw)
    finally:
        sc.close()
