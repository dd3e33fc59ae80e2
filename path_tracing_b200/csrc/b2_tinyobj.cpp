// b2_tinyobj.cpp -- the ONE translation unit that compiles the reference's vendored tinyobjloader
// (include/tiny_obj_loader.h, v2.0.0, MIT; found by __graft_entry__.build() under /root/reference/include or
// $B2_TINYOBJ_INCLUDE -- the header is used where it lies, nothing of it is copied into this tree) and calls
// tinyobj::ObjReader::ParseFromFile (tiny_obj_loader.h:546), the loading surface BASELINE.json's north_star names.
// The reference itself compiles the loader (src/tiny_obj_loader.cpp:1-2) but has no call site; the adapter that turns
// the result into Triangle / Material / `G` groups is b2_obj.cpp.
#define TINYOBJLOADER_IMPLEMENTATION
#include "tiny_obj_loader.h"
#include "b2_obj_mesh.hpp"

bool b2_tinyobj_available() { return true; }

bool b2_tinyobj_read(const std::string &obj_path, ObjMesh &out, std::string &err) {
    tinyobj::ObjReaderConfig cfg;              // defaults: triangulate = true, .mtl searched next to the OBJ
    cfg.vertex_color = false;
    tinyobj::ObjReader rd;
    if (!rd.ParseFromFile(obj_path, cfg)) { err = rd.Error().empty() ? std::string("parse failed") : rd.Error(); return false; }
    const tinyobj::attrib_t &at = rd.GetAttrib();
    out.pos.assign(at.vertices.begin(), at.vertices.end());
    out.uv.assign(at.texcoords.begin(), at.texcoords.end());
    const long nv = (long) out.pos.size() / 3, nt = (long) out.uv.size() / 2;
    size_t n = 0;
    for (const auto &sh : rd.GetShapes()) n += sh.mesh.num_face_vertices.size();
    out.faces.reserve(n);
    for (const auto &sh : rd.GetShapes()) {            // shapes in file order, faces in file order within a shape
        size_t off = 0;
        for (size_t f = 0; f < sh.mesh.num_face_vertices.size(); ++f) {
            const unsigned nfv = sh.mesh.num_face_vertices[f];
            if (nfv != 3) { err = "non-triangle face after triangulation"; return false; }
            ObjFace of;
            for (int k = 0; k < 3; ++k) {
                const tinyobj::index_t ix = sh.mesh.indices[off + k];
                if (ix.vertex_index < 0 || ix.vertex_index >= nv || ix.texcoord_index >= nt) { err = "face index out of range"; return false; }
                of.v[k] = ix.vertex_index; of.t[k] = ix.texcoord_index >= 0 ? ix.texcoord_index : -1;
            }
            of.m = f < sh.mesh.material_ids.size() ? sh.mesh.material_ids[f] : -1;
            out.faces.push_back(of);
            off += 3;
        }
    }
    const auto &ms = rd.GetMaterials();
    out.mtls.resize(ms.size());
    for (size_t i = 0; i < ms.size(); ++i) {
        const tinyobj::material_t &m = ms[i];
        ObjMtl &o = out.mtls[i];
        o.name = m.name; o.map_kd = m.diffuse_texname;
        o.kd[0] = (float) m.diffuse[0]; o.kd[1] = (float) m.diffuse[1]; o.kd[2] = (float) m.diffuse[2];
        o.pr = (float) m.roughness; o.pm = (float) m.metallic; o.ni = (float) m.ior; o.ns = (float) m.shininess; o.d = (float) m.dissolve;
        o.illum = m.illum;
    }
    for (ObjFace &f : out.faces) if (f.m >= (int) out.mtls.size()) f.m = -1;
    return true;
}
