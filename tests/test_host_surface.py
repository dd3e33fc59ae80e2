"""Host-side surface kept from the reference (no GPU needed): scene grammar parser (main_cli.cpp:99-141), film
conversion and writers (main_cli.cpp:221-254), the C-ABI symbol table, and loud failure without a device."""
import ctypes as C
import os
import re
import zlib
import numpy as np
import pytest
from oracle import oracle_py as O
from path_tracing_b200 import _abi as A
from path_tracing_b200 import api
from path_tracing_b200.scene import SceneData, reference_scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "b2pt.h")).read()
    declared = set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(A.SYMBOLS), declared ^ set(A.SYMBOLS)
    L = A.lib()
    for s in declared:
        assert hasattr(L, s), s
    assert b"sm_100a" in L.b2_version()


def test_pod_layouts_match_header():
    assert C.sizeof(A.RenderParams) == 80 and C.sizeof(A.BvhParams) == 16 and C.sizeof(A.BvhInfo) == 56
    assert C.sizeof(A.Stats) == 12 * 8 + 7 * 4 + 4    # 7 floats, padded to 8-byte alignment
    assert C.sizeof(A.ParsedScene) == 4 * 8 + 4 * 8 + 40 + 8


GRAMMAR = """// comment line with letters E V F that must be ignored
E 0 1 -10
V 0 -1 0   0 1 0
F 50
R 320 200
M 1 1 1 0.0 0.9 0.0 0.1 -1 64
T -3 1.787 1.578  -3 1.927 1.721  3 1.927 1.721
G 2
M 0.8 0.9 1 0 0 1.5
S 0.25 -0.3 0.4 0.15
G 0
T 0 0 0 1 0 0 0 1 0   // trailing comment
L -2 3 2   0 -1 0   1 1 1   360 0 0.5
L 2 10 2   0 -1 0   5 5 5   30 1 5
xyz stray characters are skipped one at a time
"""


def test_parser_follows_the_reference_grammar(tmp_path):
    p = tmp_path / "scene.txt"
    p.write_text(GRAMMAR)
    s = SceneData.from_file(p)
    assert (s.width, s.height) == (320, 200)
    assert np.allclose(s.camera["eye"], [0, 1, -10]) and float(s.camera["fov"]) == 50.0
    assert len(s.tris) == 2 and len(s.spheres) == 1 and len(s.lights) == 2
    # obj_id is one counter shared by spheres and triangles in file order (main_cli.cpp:109,118)
    assert list(s.tris["obj_id"]) == [0, 2] and list(s.spheres["obj_id"]) == [1]
    assert list(s.tris["group"]) == [0, 0] and list(s.spheres["group"]) == [2]
    # surplus numbers on an M line are dropped character by character: roughness 0, metallic 0.9, eta 0
    m0 = s.materials[s.tris["material"][0]]
    assert (float(m0["roughness"]), float(m0["metallic"]), float(m0["eta"])) == (0.0, np.float32(0.9), 0.0)
    # the second triangle uses the material current at that point (the glass one)
    assert float(s.materials[s.tris["material"][1]]["eta"]) == 1.5
    assert s.lights["cutoff"][0] == np.float32(360.0) * np.float32(0.01745329251994329576923690768489)
    assert list(s.lights["is_parallel"]) == [0, 1] and float(s.lights["ball_r"][1]) == 5.0
    # same file through the oracle's parser (restatement of the same loop)
    d = O.Port().parse(str(p))
    f = s.to_flat()
    for k in ("kind", "group", "geom", "mtl", "lights", "cam"):
        assert np.array_equal(np.asarray(f[k]), np.asarray(getattr(d, k))), k


def test_parser_errors_and_empty_inputs(tmp_path):
    with pytest.raises(A.B2Error):
        SceneData.from_file(tmp_path / "missing.txt")
    (tmp_path / "empty.txt").write_text("")
    s = SceneData.from_file(tmp_path / "empty.txt")
    assert len(s.tris) == 0 and len(s.lights) == 0 and s.width == 0
    # a failed numeric extraction ends the loop (stream fail state), as `while(input >> t)` does
    (tmp_path / "bad.txt").write_text("T 0 0 0 1 0 0 0 1 0\nE 1 2 oops\nT 0 0 0 1 0 0 0 1 0\n")
    s = SceneData.from_file(tmp_path / "bad.txt")
    assert len(s.tris) == 1


@pytest.mark.parametrize("name", ["cornell", "mis"])
def test_reference_scenes_roundtrip_through_text(tmp_path, name):
    s = reference_scene(name)
    s.write_text(tmp_path / "s.txt")
    s2 = SceneData.from_file(tmp_path / "s.txt")
    a, b = s.to_flat(), s2.to_flat()
    for k in a:
        assert np.array_equal(np.asarray(a[k]), np.asarray(b[k])), k
    if name == "cornell":
        assert (len(s.tris), len(s.spheres), len(s.lights)) == (36, 5, 4)
    else:
        assert (len(s.tris), len(s.spheres), len(s.lights)) == (48, 0, 5)


def test_film_to_rgb8_matches_reference_tonemap():
    x = np.array([[-1.0, 0.0, 1e-9], [0.18, 0.5, 1.0], [7.5, np.nan, np.inf], [0.999999, 0.0031, 0.25]], np.float32).reshape(2, 2, 3)
    got = api.film_to_rgb8(x)
    c = np.where(np.isnan(x), 0.0, np.clip(np.nan_to_num(x, nan=0.0, posinf=np.inf), 0.0, 1.0))
    want = (np.power(c.astype(np.float32), np.float32(1.0 / 2.2)) * np.float32(255.0)).astype(np.uint8)
    assert np.array_equal(got, want)
    assert got[0, 0, 0] == 0 and got[0, 1, 2] == 255 and got[1, 0, 1] == 0 and got[1, 0, 2] == 255


def _decode_png(path):
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, W, H = 8, b"", 0, 0
    while pos < len(b):
        n = int.from_bytes(b[pos:pos + 4], "big"); typ = b[pos + 4:pos + 8]; data = b[pos + 8:pos + 8 + n]
        assert zlib.crc32(typ + data) == int.from_bytes(b[pos + 8 + n:pos + 12 + n], "big")
        if typ == b"IHDR":
            W, H = int.from_bytes(data[0:4], "big"), int.from_bytes(data[4:8], "big")
            assert data[8] == 8 and data[9] == 2
        elif typ == b"IDAT":
            idat += data
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(H, 1 + 3 * W)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(H, W, 3)


def test_image_writers(tmp_path):
    rng = np.random.default_rng(0)
    film = rng.uniform(0, 1.2, (13, 17, 3)).astype(np.float32)
    rgb = api.film_to_rgb8(film)
    api.write_image(tmp_path / "a.png", film)
    assert np.array_equal(_decode_png(tmp_path / "a.png"), rgb)
    try:
        from PIL import Image
        assert np.array_equal(np.asarray(Image.open(tmp_path / "a.png")), rgb)
    except ImportError:
        pass
    api.write_image(tmp_path / "a.ppm", film)
    b = open(tmp_path / "a.ppm", "rb").read()
    assert b.startswith(b"P6\n17 13\n255\n") and np.array_equal(np.frombuffer(b[len(b"P6\n17 13\n255\n"):], np.uint8).reshape(13, 17, 3), rgb)
    api.write_image(tmp_path / "a.pfm", film)
    b = open(tmp_path / "a.pfm", "rb").read()
    hdr = b"PF\n17 13\n-1.0\n"
    assert b.startswith(hdr) and np.array_equal(np.frombuffer(b[len(hdr):], np.float32).reshape(13, 17, 3)[::-1], film)
    with pytest.raises(A.B2Error):
        api.write_image(tmp_path / "a.bmp", film)


def test_no_cpu_fallback_without_a_device():
    """On a box without a GPU every compute entry point must fail loudly (B2_ERR_NO_DEVICE), never compute on the CPU."""
    if api.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(A.B2Error, match="no CUDA device"):
        api.Scene(reference_scene("cornell"))
    with pytest.raises(A.B2Error, match="no CUDA device"):
        api.bsdf_pdf(np.zeros((1, 6)), np.zeros((1, 3)), np.zeros((1, 3)), np.zeros((1, 3)))
    with pytest.raises(A.B2Error, match="no CUDA device"):
        api.philox(np.zeros((1, 4)), np.zeros((1, 2)))


def test_frame_rms_is_the_viewers_formula():
    """main.cpp:502-531: rms = sqrt(sum over pixels and channels of (cur - last)^2) / 255 on the tone-mapped 8-bit frames."""
    rng = np.random.default_rng(5)
    a = rng.integers(0, 256, (37, 41, 3)).astype(np.uint8); b = rng.integers(0, 256, (37, 41, 3)).astype(np.uint8)
    want = np.sqrt(((a.astype(np.float64) - b.astype(np.float64)) ** 2).sum()) / 255.0
    assert abs(api.frame_rms(a, b) - want) < 1e-4 * want
    assert api.frame_rms(a, a) == 0.0
