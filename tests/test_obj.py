"""OBJ / MTL surface (b2_load_obj, `O` directive).  The library reads OBJ files through the reference's vendored
tinyobjloader v2.0.0 (csrc/b2_tinyobj.cpp calls tinyobj::ObjReader::ParseFromFile) whenever it was built with that header,
and keeps a small own reader for builds without it.  Both readers and the adapter above them are pinned here: against the
committed dump tests/golden/obj/small_tinyobj.txt (made by oracle/obj_ref.cpp, a separate program around the same loader,
see tests/golden/make_obj_golden.sh), live against oracle/_ref/obj_ref on a generated 20 k-triangle mesh, and against each
other on the 1 M-triangle C3 mesh itself (faces, materials, groups bit-identical).  Host-only: no GPU needed."""
import os
import subprocess
import numpy as np
import pytest

from path_tracing_b200 import gen, _abi as A
from path_tracing_b200.scene import SceneData

OWN, TINYOBJ = 1, 2
HAVE_TINYOBJ = bool(A.lib().b2_obj_has_tinyobj())
READERS = [OWN] + ([TINYOBJ] if HAVE_TINYOBJ else [])

HERE = os.path.dirname(os.path.abspath(__file__))
FIX = os.path.join(HERE, "golden", "obj")
OBJ_REF = os.path.join(HERE, "..", "oracle", "_ref", "obj_ref")


def parse_dump(text):
    """-> (pos[n,3,3] f32, uv[n,3,2] f32, has_uv[n], mat_id[n], materials list of dicts)"""
    lines = text.strip().split("\n")
    n = int(lines[0].split()[1])
    tri = np.zeros((n, 15), np.uint32); has_uv = np.zeros(n, np.int32); mid = np.zeros(n, np.int32)
    for i in range(n):
        t = lines[1 + i].split()
        tri[i] = [int(x, 16) for x in t[:15]]
        has_uv[i], mid[i] = int(t[15]), int(t[16])
    f = tri.view(np.float32)
    m0 = 1 + n
    nm = int(lines[m0].split()[1])
    mats = []
    for i in range(nm):
        name, vals, tex = [x.strip() for x in lines[m0 + 1 + i].split("|")]
        v = vals.split()
        fl = np.array([int(x, 16) for x in v[:8]], np.uint32).view(np.float32)
        mats.append(dict(name=name, kd=fl[0:3], pr=fl[3], pm=fl[4], ni=fl[5], ns=fl[6], d=fl[7], illum=int(v[8]), tex=tex))
    return f[:, :9].reshape(n, 3, 3), f[:, 9:15].reshape(n, 3, 2), has_uv, mid, mats


def read_ppm(path):
    b = open(path, "rb").read()
    toks, p = [], 0
    while len(toks) < 4:
        while b[p:p + 1].isspace():
            p += 1
        if b[p:p + 1] == b"#":
            p = b.index(b"\n", p)
            continue
        q = p
        while not b[q:q + 1].isspace():
            q += 1
        toks.append(b[p:q]); p = q
    w, h, mx = int(toks[1]), int(toks[2]), int(toks[3])
    return np.frombuffer(b[p + 1:p + 1 + 3 * w * h], np.uint8).reshape(h, w, 3).astype(np.float32) / np.float32(mx)


def expected_materials(uv, has_uv, mid, mats, rough_of, texdir, default=(0.8, 0.8, 0.8, 1.0)):
    """The adapter rules of csrc/b2_obj.cpp restated in numpy (float32 throughout)."""
    n = len(mid)
    out = np.zeros((n, 6), np.float32)
    tex_cache = {}
    for i in range(n):
        if mid[i] < 0:
            out[i] = [default[0], default[1], default[2], default[3], 0.0, 0.0]
            continue
        m = mats[mid[i]]
        tx = np.ones(3, np.float32)
        if m["tex"] and has_uv[i]:
            img = tex_cache.setdefault(m["tex"], read_ppm(os.path.join(texdir, m["tex"])))
            u = (uv[i, 0, 0] + uv[i, 1, 0] + uv[i, 2, 0]) / np.float32(3.0)
            v = (uv[i, 0, 1] + uv[i, 1, 1] + uv[i, 2, 1]) / np.float32(3.0)
            u = np.float32(u - np.floor(u)); v = np.float32(v - np.floor(v))
            x = min(max(int(np.float32(u * np.float32(img.shape[1]))), 0), img.shape[1] - 1)
            y = min(max(int(np.float32((np.float32(1.0) - v) * np.float32(img.shape[0]))), 0), img.shape[0] - 1)
            tx = img[y, x]
        transparent = m["illum"] in (4, 6, 7) or m["d"] < 1.0
        out[i, 0:3] = m["kd"] * tx
        out[i, 3] = rough_of(m)
        out[i, 4] = m["pm"]
        out[i, 5] = m["ni"] if transparent else 0.0
    return out


def test_default_reader_is_tinyobj_when_the_reference_header_was_there_at_build_time():
    if os.path.exists("/root/reference/include/tiny_obj_loader.h"):
        assert HAVE_TINYOBJ, "libb2pt.so was built without the vendored loader although its header is present"
    if not HAVE_TINYOBJ:
        with pytest.raises(RuntimeError, match="not compiled"):
            SceneData.from_obj(os.path.join(FIX, "small.obj"), reader=TINYOBJ)


@pytest.mark.parametrize("reader", READERS)
def test_obj_reader_matches_tinyobj_golden(reader):
    pos, uv, has_uv, mid, mats = parse_dump(open(os.path.join(FIX, "small_tinyobj.txt")).read())
    sd = SceneData.from_obj(os.path.join(FIX, "small.obj"), group_grid=1, first_obj_id=5, first_group=7, reader=reader)
    assert len(sd.tris) == len(pos) == 11
    got = np.stack([sd.tris["v0"], sd.tris["v1"], sd.tris["v2"]], 1)
    assert np.array_equal(got.view(np.uint32), pos.view(np.uint32))          # same faces, same order, same float bits
    assert np.array_equal(sd.tris["obj_id"], np.arange(5, 5 + 11))
    assert np.all(sd.tris["group"] == 7)
    rough = {"red_rough": 0.7, "textured": 0.35, "shiny_ns": np.sqrt(np.float32(2.0) / np.float32(100.0)), "glass": 0.0, "faded": 0.5}
    exp = expected_materials(uv, has_uv, mid, mats, lambda m: np.float32(rough[m["name"]]), FIX)
    m = sd.materials[sd.tris["material"]]
    gotm = np.concatenate([m["base_color"], m["roughness"][:, None], m["metallic"][:, None], m["eta"][:, None]], 1)
    assert np.allclose(gotm, exp, rtol=0, atol=1e-7), np.abs(gotm - exp).max()
    # spot checks that do not go through the restatement: glass is transparent with eta = Ni, `faded` (d < 1) too,
    # the face with an unknown material falls back to the default colour
    names = [mats[i]["name"] if i >= 0 else None for i in mid]
    assert gotm[names.index("glass"), 5] == np.float32(1.5) and gotm[names.index("faded"), 5] == np.float32(1.33)
    assert np.allclose(gotm[names.index(None)], [0.8, 0.8, 0.8, 1.0, 0.0, 0.0])
    assert len(np.unique(sd.tris["material"])) == len(np.unique(gotm, axis=0))  # identical materials are shared


@pytest.mark.skipif(not os.path.exists(OBJ_REF), reason="oracle/_ref/obj_ref not built (needs /root/reference)")
def test_generated_mesh_against_live_tinyobj(tmp_path):
    scene = gen.write_obj_scene(str(tmp_path), n_tris=20000, tex_size=64)
    dump = subprocess.run([OBJ_REF, str(tmp_path / "mesh.obj")], capture_output=True, text=True, check=True).stdout
    pos, uv, has_uv, mid, mats = parse_dump(dump)
    sd = SceneData.from_file(scene, obj_directive=True)
    assert len(sd.tris) == len(pos) > 19000
    got = np.stack([sd.tris["v0"], sd.tris["v1"], sd.tris["v2"]], 1)
    assert np.array_equal(got.view(np.uint32), pos.view(np.uint32))
    exp = expected_materials(uv, has_uv, mid, mats, lambda m: m["pr"], str(tmp_path))
    m = sd.materials[sd.tris["material"]]
    gotm = np.concatenate([m["base_color"], m["roughness"][:, None], m["metallic"][:, None], m["eta"][:, None]], 1)
    assert np.allclose(gotm, exp, rtol=0, atol=1e-7)
    assert len(sd.lights) == 3 and (sd.width, sd.height) == (1920, 1080)
    g = np.unique(sd.tris["group"])
    assert g.min() >= 0 and len(g) > 4                                            # spatial G groups for the CPU oracle's cull


@pytest.mark.skipif(not HAVE_TINYOBJ, reason="libb2pt.so built without tiny_obj_loader.h")
def test_own_reader_equals_tinyobj_on_the_c3_mesh(tmp_path):
    """The 1 M-triangle OBJ of BASELINE config 3 itself, through both readers: same faces in the same order (float bits),
    same materials, same groups."""
    gen.write_obj_scene(str(tmp_path), n_tris=1 << 20)
    a = SceneData.from_obj(str(tmp_path / "mesh.obj"), reader=TINYOBJ)
    b = SceneData.from_obj(str(tmp_path / "mesh.obj"), reader=OWN)
    assert len(a.tris) == len(b.tris) > 1_000_000
    assert a.tris.tobytes() == b.tris.tobytes()
    assert a.materials.tobytes() == b.materials.tobytes()


def test_scene_directive_splices_the_mesh(tmp_path):
    (tmp_path / "m.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nv 1 1 0\nf 1 2 3\nf 2 4 3\n")
    (tmp_path / "s.txt").write_text("E 0 0 -3\nV 0 0 0 0 1 0\nF 40\nR 64 48\nM 0.2 0.4 0.6 0.5 0 0\nT 0 0 1 1 0 1 0 1 1\nG 3\nO m.obj\n"
                                    "T 0 0 2 1 0 2 0 1 2\nL 0 3 0 0 -1 0 1 1 1 90 0 0.1\n")
    sd = SceneData.from_file(str(tmp_path / "s.txt"), obj_directive=True)
    assert len(sd.tris) == 4 and list(sd.tris["obj_id"]) == [0, 1, 2, 3]
    assert list(sd.tris["group"]) == [0, 3, 3, 4]              # mesh starts at the current G (1 x 1 grid), the next T comes after it
    m = sd.materials[sd.tris["material"]]
    assert np.allclose(m["base_color"], [0.2, 0.4, 0.6]) and np.allclose(m["roughness"], 0.5)   # faces without usemtl take the current M
    assert len(sd.lights) == 1 and (sd.width, sd.height) == (64, 48)


def test_polygons_are_triangulated_and_errors_are_loud(tmp_path):
    (tmp_path / "q.obj").write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nv -1 0.5 0\nf 1 2 3 4 5\n")
    sd = SceneData.from_obj(str(tmp_path / "q.obj"), reader=OWN)                      # own reader: a fan around the first vertex
    assert len(sd.tris) == 3
    assert np.allclose(sd.tris["v0"], [[0, 0, 0]] * 3) and np.allclose(sd.tris["v1"][:, :2], [[1, 0], [1, 1], [0, 1]])
    for reader in READERS:                                                            # any triangulation covers the polygon's area
        sd = SceneData.from_obj(str(tmp_path / "q.obj"), reader=reader)
        area = 0.5 * np.linalg.norm(np.cross(sd.tris["v1"] - sd.tris["v0"], sd.tris["v2"] - sd.tris["v0"]), axis=1).sum()
        assert len(sd.tris) == 3 and abs(area - 1.5) < 1e-6
        with pytest.raises(RuntimeError, match="cannot open OBJ"):
            SceneData.from_obj(str(tmp_path / "nope.obj"), reader=reader)
    (tmp_path / "bad.obj").write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nf 1 2 9\n")
    with pytest.raises(RuntimeError, match="out of range"):
        SceneData.from_obj(str(tmp_path / "bad.obj"), reader=OWN)
    if HAVE_TINYOBJ:                                                                  # tinyobj reports it in its own words
        with pytest.raises(RuntimeError, match="tinyobj"):
            SceneData.from_obj(str(tmp_path / "bad.obj"), reader=TINYOBJ)
    (tmp_path / "t.mtl").write_text("newmtl a\nKd 1 1 1\nmap_Kd t.png\n")
    (tmp_path / "t.png").write_bytes(b"\x89PNG....")
    (tmp_path / "t.obj").write_text("mtllib t.mtl\nv 0 0 0\nv 1 0 0\nv 1 1 0\nvt 0 0\nusemtl a\nf 1/1 2/1 3/1\n")
    with pytest.raises(RuntimeError, match="texture"):
        SceneData.from_obj(str(tmp_path / "t.obj"))


def test_spatial_groups_cover_the_requested_grid(tmp_path):
    gen.write_obj_scene(str(tmp_path), n_tris=4000, tex_size=16)
    sd = SceneData.from_obj(str(tmp_path / "mesh.obj"), group_grid=3, first_group=10)
    g = np.unique(sd.tris["group"])
    assert g.min() >= 10 and g.max() < 19 and len(g) == 9
