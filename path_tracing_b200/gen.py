"""Procedural scenes for the large configurations of BASELINE.json (there are no assets in the reference and no
network): C3 (~1 M triangles, textured diffuse / glossy materials), C4 (10 M triangles), C5 (traversal microbench).

Everything is generated from fixed seeds with numpy, as SceneData in the reference's own object model (triangles with
per-object Material, `G` groups, `L` lights) so the unmodified CPU oracle applies to any subset of it.  "Textures" are
baked per triangle (albedo = procedural texture at the triangle centroid), which is how the OBJ adapter maps `map_Kd`
as well: the reference's Material has no texture slot.
"""
import numpy as np
from . import _abi as A
from .scene import SceneData


def _value_noise(x, y, seed, octaves=4):
    """Cheap tileable value noise in [0,1] (bilinear lattice noise, summed octaves)."""
    rng = np.random.default_rng(seed)
    out = np.zeros_like(x, dtype=np.float64)
    amp, tot = 1.0, 0.0
    for o in range(octaves):
        n = 8 << o
        lat = rng.uniform(0, 1, (n + 1, n + 1))
        lat[n, :] = lat[0, :]; lat[:, n] = lat[:, 0]
        fx, fy = (x % 1.0) * n, (y % 1.0) * n
        ix, iy = np.minimum(fx.astype(np.int64), n - 1), np.minimum(fy.astype(np.int64), n - 1)
        tx, ty = fx - ix, fy - iy
        tx, ty = tx * tx * (3 - 2 * tx), ty * ty * (3 - 2 * ty)
        v = (lat[ix, iy] * (1 - tx) + lat[ix + 1, iy] * tx) * (1 - ty) + (lat[ix, iy + 1] * (1 - tx) + lat[ix + 1, iy + 1] * tx) * ty
        out += amp * v; tot += amp; amp *= 0.5
    return out / tot


def _grid_mesh(nx, ny, pos_fn):
    """(nx x ny) quads -> 2*nx*ny triangles; pos_fn(u, v) -> xyz for u, v in [0,1]."""
    u, v = np.meshgrid(np.linspace(0, 1, nx + 1), np.linspace(0, 1, ny + 1), indexing="ij")
    P = pos_fn(u, v)                                           # (nx+1, ny+1, 3)
    a, b, c, d = P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]
    t1 = np.stack([a, b, c], axis=2).reshape(-1, 3, 3)
    t2 = np.stack([a, c, d], axis=2).reshape(-1, 3, 3)
    uc = 0.5 * (u[:-1, :-1] + u[1:, 1:]).reshape(-1); vc = 0.5 * (v[:-1, :-1] + v[1:, 1:]).reshape(-1)
    tri = np.concatenate([t1, t2]); uv = np.concatenate([np.stack([uc, vc], 1)] * 2)
    return tri.astype(np.float32), uv


def terrain_scene(n_tris=1 << 20, seed=1, n_blobs=12, width=1920, height=1080, groups=64):
    """~n_tris triangles: a displaced heightfield terrain (half of them) plus `n_blobs` displaced spheres hovering over
    it, inside an open box of 5 walls.  Materials: 32 entries, rough diffuse to glossy metal, albedo from two procedural
    textures (checker x noise).  3 spot lights.  Triangles are assigned to `groups` spatial `G` groups (x-z tiles) so the
    reference's own group-AABB cull (cpu_bdpt.cpp:43-44) keeps the CPU oracle usable on subsets."""
    rng = np.random.default_rng(seed)
    n_terrain = n_tris // 2
    n = max(int(np.sqrt(n_terrain / 2)), 2)

    def terrain(u, v):
        h = 0.35 * _value_noise(u, v, seed + 11, 5) + 0.02 * _value_noise(u * 8, v * 8, seed + 12, 2)
        return np.stack([(u - 0.5) * 4.0, h - 0.6, (v - 0.5) * 4.0 + 2.0], axis=-1)

    tris, uvs, kinds = [], [], []
    t, uv = _grid_mesh(n, n, terrain)
    tris.append(t); uvs.append(uv); kinds.append(np.zeros(len(t), np.int32))
    per_blob = max((n_tris - len(t)) // max(n_blobs, 1), 8)
    m = max(int(np.sqrt(per_blob / 4)), 2)
    for bi in range(n_blobs):
        c = np.array([rng.uniform(-1.6, 1.6), rng.uniform(-0.1, 0.7), rng.uniform(0.6, 3.4)])
        r = rng.uniform(0.12, 0.3)

        def blob(u, v, c=c, r=r, bi=bi):
            th, ph = u * 2 * np.pi, v * np.pi
            d = 1.0 + 0.15 * (_value_noise(u, v, seed + 100 + bi, 3) - 0.5)
            return np.stack([c[0] + r * d * np.sin(ph) * np.cos(th), c[1] + r * d * np.cos(ph), c[2] + r * d * np.sin(ph) * np.sin(th)], axis=-1)

        t, uv = _grid_mesh(2 * m, m, blob)
        area = np.linalg.norm(np.cross(t[:, 1] - t[:, 0], t[:, 2] - t[:, 0]), axis=1)
        keep = area > 1e-12                                   # pole quads collapse to degenerate triangles
        tris.append(t[keep]); uvs.append(uv[keep]); kinds.append(np.full(int(keep.sum()), 1 + bi % 3, np.int32))
    # 5 walls (2 triangles each): floor under the terrain, back, left, right, ceiling
    X0, X1, Y0, Y1, Z0, Z1 = -2.2, 2.2, -0.8, 2.2, -0.2, 4.4
    q = lambda a, b, c, d: [np.array([a, b, c], np.float32), np.array([a, c, d], np.float32)]
    walls = q((X0, Y0, Z0), (X1, Y0, Z0), (X1, Y0, Z1), (X0, Y0, Z1)) + q((X0, Y0, Z1), (X1, Y0, Z1), (X1, Y1, Z1), (X0, Y1, Z1)) + \
        q((X0, Y0, Z0), (X0, Y0, Z1), (X0, Y1, Z1), (X0, Y1, Z0)) + q((X1, Y0, Z0), (X1, Y1, Z0), (X1, Y1, Z1), (X1, Y0, Z1)) + \
        q((X0, Y1, Z0), (X0, Y1, Z1), (X1, Y1, Z1), (X1, Y1, Z0))
    tris.append(np.stack(walls)); uvs.append(np.zeros((10, 2))); kinds.append(np.full(10, 4, np.int32))
    tri = np.concatenate(tris).astype(np.float32); uv = np.concatenate(uvs); kind = np.concatenate(kinds)
    N = len(tri)
    # materials: per class a (roughness, metallic) family; albedo quantised to 8 levels from checker x noise
    tex = 0.25 + 0.75 * _value_noise(uv[:, 0] * 4, uv[:, 1] * 4, seed + 5, 3) * (0.6 + 0.4 * (((uv[:, 0] * 16).astype(int) + (uv[:, 1] * 16).astype(int)) % 2))
    lvl = np.clip((tex * 8).astype(np.int32), 0, 7)
    fam = np.array([[1.0, 0.0], [0.35, 0.0], [0.25, 1.0], [0.6, 0.5], [1.0, 0.0]], np.float32)      # terrain, blobs x3, walls
    tint = np.array([[0.55, 0.75, 0.45], [0.9, 0.5, 0.3], [0.95, 0.85, 0.55], [0.4, 0.55, 0.9], [0.7, 0.7, 0.7]], np.float32)
    mats = np.zeros(5 * 8, A.Material)
    for k in range(5):
        for l in range(8):
            mats[k * 8 + l]["base_color"] = tint[k] * (l + 1) / 8.0
            mats[k * 8 + l]["roughness"], mats[k * 8 + l]["metallic"], mats[k * 8 + l]["eta"] = fam[k, 0], fam[k, 1], 0.0
    T = np.zeros(N, A.Tri)
    T["v0"], T["v1"], T["v2"] = tri[:, 0], tri[:, 1], tri[:, 2]
    T["obj_id"] = np.arange(N, dtype=np.int32)
    T["material"] = kind * 8 + lvl
    g = max(int(np.sqrt(groups)), 1)
    cen = tri.mean(axis=1)
    gx = np.clip(((cen[:, 0] - X0) / (X1 - X0) * g).astype(np.int32), 0, g - 1); gz = np.clip(((cen[:, 2] - Z0) / (Z1 - Z0) * g).astype(np.int32), 0, g - 1)
    T["group"] = gx * g + gz
    L = np.zeros(3, A.Light)
    L["pos"] = [[0.0, 2.0, 2.0], [-1.5, 1.8, 0.8], [1.5, 1.8, 3.2]]
    L["dir"] = [[0.0, -1.0, 0.05], [0.4, -1.0, 0.3], [-0.4, -1.0, -0.3]]
    L["illum"] = [[6.0, 6.0, 6.0], [4.0, 3.0, 2.0], [2.0, 3.0, 4.0]]
    L["cutoff"] = np.radians([70.0, 55.0, 55.0]).astype(np.float32); L["is_parallel"] = 0; L["ball_r"] = 0.05
    cam = np.zeros((), A.Camera)
    cam["eye"], cam["look_at"], cam["view_up"], cam["fov"] = [0.0, 0.9, -0.15], [0.0, -0.2, 2.2], [0.0, 1.0, 0.0], 55.0
    return SceneData(T, np.zeros(0, A.Sphere), mats, L, cam, width, height)


def soup_scene(n_tris, seed=12345, lights=2):
    """C4 / C5 style: uniform centroids in the unit box, edge length log-uniform 1e-3 .. 1e-2."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(0, 1, (n_tris, 1, 3)).astype(np.float32)
    e = (rng.normal(size=(n_tris, 3, 3)) * np.exp(rng.uniform(np.log(1e-3), np.log(1e-2), (n_tris, 1, 1)))).astype(np.float32)
    v = c + e
    T = np.zeros(n_tris, A.Tri)
    T["v0"], T["v1"], T["v2"] = v[:, 0], v[:, 1], v[:, 2]
    T["obj_id"] = np.arange(n_tris, dtype=np.int32); T["group"] = 0
    T["material"] = rng.integers(0, 4, n_tris)
    mats = np.zeros(4, A.Material)
    mats["base_color"] = [[0.8, 0.8, 0.8], [0.8, 0.3, 0.2], [0.9, 0.9, 0.6], [0.3, 0.4, 0.9]]
    mats["roughness"] = [1.0, 0.4, 0.2, 0.7]; mats["metallic"] = [0.0, 0.0, 1.0, 0.5]
    L = np.zeros(lights, A.Light)
    L["pos"] = rng.uniform(0.2, 0.8, (lights, 3)); L["dir"] = [0, -1, 0]; L["illum"] = 3.0; L["cutoff"] = np.float32(np.pi); L["ball_r"] = 0.01
    cam = np.zeros((), A.Camera)
    cam["eye"], cam["look_at"], cam["view_up"], cam["fov"] = [0.5, 0.5, -1.2], [0.5, 0.5, 0.5], [0, 1, 0], 50.0
    return SceneData(T, np.zeros(0, A.Sphere), mats, L, cam, 3840, 2160)


def random_rays(n, lo, hi, seed=7):
    """C5: origins uniform in the box, directions uniform on the sphere; second point set for any-hit segments."""
    rng = np.random.default_rng(seed)
    org = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    p2 = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    return org, d, p2


# ---- C3: the same kind of scene as an OBJ + MTL + PPM textures + scene header, for the OBJ adapter (b2_load_obj) ----
def _write_ppm(path, img):
    img = np.clip(np.asarray(img) * 255.0 + 0.5, 0, 255).astype(np.uint8)
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (img.shape[1], img.shape[0]))
        f.write(img.tobytes())


def write_obj_scene(outdir, n_tris=1 << 20, seed=1, n_blobs=12, width=1920, height=1080, tex_size=1024):
    """Writes mesh.obj, mesh.mtl, checker.ppm, noise.ppm and scene.txt (reference grammar + `O mesh.obj`) into outdir.
    ~n_tris triangles: displaced terrain + n_blobs displaced spheres + 5 walls; Kd from the two procedural textures,
    Pr in {0.2 .. 1.0}, Pm in {0, 0.9}.  Returns the path of scene.txt."""
    import os
    os.makedirs(outdir, exist_ok=True)
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.linspace(0, 1, tex_size, endpoint=False), np.linspace(0, 1, tex_size, endpoint=False), indexing="ij")
    chk = (((xx * 16).astype(int) + (yy * 16).astype(int)) % 2).astype(np.float64)
    checker = np.stack([0.35 + 0.6 * chk, 0.35 + 0.55 * chk, 0.3 + 0.5 * chk], -1)
    nz = _value_noise(xx, yy, seed + 5, 4)
    noise = np.stack([0.25 + 0.7 * nz, 0.3 + 0.65 * nz, 0.2 + 0.5 * nz], -1)
    _write_ppm(os.path.join(outdir, "checker.ppm"), checker)
    _write_ppm(os.path.join(outdir, "noise.ppm"), noise)
    mtl = [("terrain", (0.75, 0.9, 0.6), 1.0, 0.0, "noise.ppm"), ("blob_a", (0.95, 0.6, 0.4), 0.4, 0.0, "checker.ppm"),
           ("blob_b", (0.95, 0.85, 0.55), 0.2, 0.9, None), ("blob_c", (0.5, 0.65, 0.95), 0.6, 0.0, "noise.ppm"), ("wall", (0.7, 0.7, 0.7), 1.0, 0.0, None)]
    with open(os.path.join(outdir, "mesh.mtl"), "w") as f:
        for name, kd, pr, pm, tex in mtl:
            f.write(f"newmtl {name}\nKd {kd[0]} {kd[1]} {kd[2]}\nPr {pr}\nPm {pm}\nNi 1.0\nillum 2\n")
            if tex:
                f.write(f"map_Kd {tex}\n")
            f.write("\n")
    n_terrain = n_tris // 2
    n = max(int(np.sqrt(n_terrain / 2)), 2)

    def terrain(u, v):
        h = 0.35 * _value_noise(u, v, seed + 11, 5) + 0.02 * _value_noise(u * 8, v * 8, seed + 12, 2)
        return np.stack([(u - 0.5) * 4.0, h - 0.6, (v - 0.5) * 4.0 + 2.0], axis=-1)

    grids = [("terrain", n, n, terrain)]
    per_blob = max((n_tris - 2 * n * n) // max(n_blobs, 1), 8)
    m = max(int(np.sqrt(per_blob / 4)), 2)
    for bi in range(n_blobs):
        c = np.array([rng.uniform(-1.6, 1.6), rng.uniform(-0.1, 0.7), rng.uniform(0.6, 3.4)])
        r = rng.uniform(0.12, 0.3)

        def blob(u, v, c=c, r=r, bi=bi):
            th, ph = u * 2 * np.pi, v * np.pi
            d = 1.0 + 0.15 * (_value_noise(u, v, seed + 100 + bi, 3) - 0.5)
            return np.stack([c[0] + r * d * np.sin(ph) * np.cos(th), c[1] + r * d * np.cos(ph), c[2] + r * d * np.sin(ph) * np.sin(th)], axis=-1)

        grids.append((("blob_a", "blob_b", "blob_c")[bi % 3], 2 * m, m, blob))
    n_faces = 0
    with open(os.path.join(outdir, "mesh.obj"), "w") as f:
        f.write("# synthetic C3 scene (path_tracing_b200.gen.write_obj_scene)\nmtllib mesh.mtl\n")
        base = 1
        for name, nx, ny, fn in grids:
            u, v = np.meshgrid(np.linspace(0, 1, nx + 1), np.linspace(0, 1, ny + 1), indexing="ij")
            P = fn(u, v).astype(np.float32).reshape(-1, 3)
            np.savetxt(f, P, fmt="v %.9g %.9g %.9g")
            np.savetxt(f, np.stack([u.reshape(-1), v.reshape(-1)], 1).astype(np.float32), fmt="vt %.9g %.9g")
            idx = (np.arange((nx + 1) * (ny + 1)).reshape(nx + 1, ny + 1) + base)
            a, b, c, d = idx[:-1, :-1].reshape(-1), idx[1:, :-1].reshape(-1), idx[1:, 1:].reshape(-1), idx[:-1, 1:].reshape(-1)
            F = np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)])
            T = P[F - base]
            keep = np.linalg.norm(np.cross(T[:, 1] - T[:, 0], T[:, 2] - T[:, 0]), axis=1) > 1e-12   # collapsed pole triangles
            F = F[keep]
            f.write(f"usemtl {name}\n")
            np.savetxt(f, np.concatenate([F, F], 1)[:, [0, 3, 1, 4, 2, 5]], fmt="f %d/%d %d/%d %d/%d")
            base += (nx + 1) * (ny + 1)
            n_faces += len(F)
        X0, X1, Y0, Y1, Z0, Z1 = -2.2, 2.2, -0.8, 2.2, -0.2, 4.4
        quads = [((X0, Y0, Z0), (X1, Y0, Z0), (X1, Y0, Z1), (X0, Y0, Z1)), ((X0, Y0, Z1), (X1, Y0, Z1), (X1, Y1, Z1), (X0, Y1, Z1)),
                 ((X0, Y0, Z0), (X0, Y0, Z1), (X0, Y1, Z1), (X0, Y1, Z0)), ((X1, Y0, Z0), (X1, Y1, Z0), (X1, Y1, Z1), (X1, Y0, Z1)),
                 ((X0, Y1, Z0), (X0, Y1, Z1), (X1, Y1, Z1), (X1, Y1, Z0))]
        f.write("usemtl wall\n")
        for q in quads:
            for p in q:
                f.write("v %.9g %.9g %.9g\n" % p)
            f.write("f -4 -3 -2\nf -4 -2 -1\n")
            n_faces += 2
    scene = os.path.join(outdir, "scene.txt")
    with open(scene, "w") as f:
        f.write(f"// synthetic C3 scene, {n_faces} triangles\nE 0 0.9 -0.15\nV 0 -0.2 2.2  0 1 0\nF 55\nR {width} {height}\n")
        f.write("L 0 2 2  0 -1 0.05  6 6 6  70 0 0.05\nL -1.5 1.8 0.8  0.4 -1 0.3  4 3 2  55 0 0.05\nL 1.5 1.8 3.2  -0.4 -1 -0.3  2 3 4  55 0 0.05\n")
        f.write("G 0\nO mesh.obj\n")
    return scene
