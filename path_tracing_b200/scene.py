"""Host-side scene containers mirroring the reference's scene objects (include/object.h, main_cli.cpp:91-141)."""
import ctypes as C
import numpy as np
from . import _abi as A


class SceneData:
    """What the reference's parser leaves behind: objects in file order (obj_id), their group keys and
    materials, the lights, the camera and the R-line resolution."""

    def __init__(self, tris, spheres, materials, lights, camera=None, width=0, height=0):
        self.tris = np.ascontiguousarray(tris, dtype=A.Tri)
        self.spheres = np.ascontiguousarray(spheres, dtype=A.Sphere)
        self.materials = np.ascontiguousarray(materials, dtype=A.Material)
        self.lights = np.ascontiguousarray(lights, dtype=A.Light)
        self.camera = np.zeros((), A.Camera) if camera is None else np.ascontiguousarray(camera, dtype=A.Camera).reshape(())
        self.width, self.height = int(width), int(height)

    # -- the reference's text grammar, parsed by the C++ host code behind the C ABI (no GPU needed) --
    @classmethod
    def from_file(cls, path):
        L = A.lib()
        pp = C.POINTER(A.ParsedScene)()
        A.check(L.b2_parse_scene_file(str(path).encode(), C.byref(pp)))
        p = pp.contents

        def grab(addr, n, dt):
            if n == 0:
                return np.zeros(0, dt)
            buf = (C.c_char * (n * dt.itemsize)).from_address(addr)
            return np.frombuffer(buf, dtype=dt, count=n).copy()

        cam = np.zeros((), A.Camera)
        c = np.array(list(p.camera), np.float32)
        cam["eye"], cam["look_at"], cam["view_up"], cam["fov"] = c[0:3], c[3:6], c[6:9], c[9]
        out = cls(grab(p.tris, p.n_tris, A.Tri), grab(p.spheres, p.n_spheres, A.Sphere),
                  grab(p.materials, p.n_materials, A.Material), grab(p.lights, p.n_lights, A.Light), cam, p.width, p.height)
        L.b2_parsed_scene_free(pp)
        return out

    # -- flat per-object arrays (the format oracle/oracle_py.SceneDesc uses) --
    @classmethod
    def from_flat(cls, kind, group, geom, mtl, lights, cam=None, width=0, height=0):
        kind = np.asarray(kind, np.int32); group = np.asarray(group, np.int32)
        geom = np.asarray(geom, np.float32).reshape(-1, 9); mtl = np.asarray(mtl, np.float32).reshape(-1, 6)
        lights = np.asarray(lights, np.float32).reshape(-1, 12)
        umat, inv = (np.unique(mtl.view(np.uint32), axis=0, return_inverse=True) if len(mtl) else (np.zeros((0, 6), np.uint32), np.zeros(0, np.int64)))
        inv = np.asarray(inv).reshape(-1)
        mats = np.zeros(len(umat), A.Material)
        um = umat.view(np.float32).reshape(-1, 6)
        mats["base_color"], mats["roughness"], mats["metallic"], mats["eta"] = um[:, 0:3], um[:, 3], um[:, 4], um[:, 5]
        ids = np.arange(len(kind), dtype=np.int32)
        ti, si = ids[kind == 1], ids[kind == 0]
        tris = np.zeros(len(ti), A.Tri)
        tris["v0"], tris["v1"], tris["v2"] = geom[ti, 0:3], geom[ti, 3:6], geom[ti, 6:9]
        tris["obj_id"], tris["group"], tris["material"] = ti, group[ti], inv[ti]
        sph = np.zeros(len(si), A.Sphere)
        sph["center"], sph["r"] = geom[si, 0:3], geom[si, 3]
        sph["obj_id"], sph["group"], sph["material"] = si, group[si], inv[si]
        lt = np.zeros(len(lights), A.Light)
        if len(lights):
            lt["pos"], lt["dir"], lt["illum"] = lights[:, 0:3], lights[:, 3:6], lights[:, 6:9]
            lt["cutoff"], lt["is_parallel"], lt["ball_r"] = lights[:, 9], lights[:, 10].astype(np.int32), lights[:, 11]
        camera = None
        if cam is not None:
            cam = np.asarray(cam, np.float32)
            camera = np.zeros((), A.Camera)
            camera["eye"], camera["look_at"], camera["view_up"], camera["fov"] = cam[0:3], cam[3:6], cam[6:9], cam[9]
        return cls(tris, sph, mats, lt, camera, width, height)

    def to_flat(self):
        n = len(self.tris) + len(self.spheres)
        kind = np.zeros(n, np.int32); group = np.zeros(n, np.int32)
        geom = np.zeros((n, 9), np.float32); mtl = np.zeros((n, 6), np.float32)
        m = self.materials
        mflat = np.concatenate([m["base_color"].reshape(-1, 3), m["roughness"][:, None], m["metallic"][:, None], m["eta"][:, None]], axis=1) if len(m) else np.zeros((0, 6), np.float32)
        t, s = self.tris, self.spheres
        if len(t):
            i = t["obj_id"]
            kind[i] = 1; group[i] = t["group"]; mtl[i] = mflat[t["material"]]
            geom[i, 0:3], geom[i, 3:6], geom[i, 6:9] = t["v0"], t["v1"], t["v2"]
        if len(s):
            i = s["obj_id"]
            kind[i] = 0; group[i] = s["group"]; mtl[i] = mflat[s["material"]]
            geom[i, 0:3] = s["center"]; geom[i, 3] = s["r"]
        L = self.lights
        lights = np.zeros((len(L), 12), np.float32)
        if len(L):
            lights[:, 0:3], lights[:, 3:6], lights[:, 6:9] = L["pos"], L["dir"], L["illum"]
            lights[:, 9], lights[:, 10], lights[:, 11] = L["cutoff"], L["is_parallel"], L["ball_r"]
        c = self.camera
        cam = np.concatenate([c["eye"], c["look_at"], c["view_up"], [c["fov"]]]).astype(np.float32)
        return dict(kind=kind, group=group, geom=geom, mtl=mtl, lights=lights, cam=cam, W=self.width, H=self.height)
