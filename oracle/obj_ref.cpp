// oracle/obj_ref.cpp -- TEST INFRASTRUCTURE, not part of the product.
// Dumps what the reference's vendored tinyobjloader (include/tiny_obj_loader.h, v2.0.0, compiled from where it lies
// under /root/reference; nothing is copied) returns for an OBJ file, so that tests/test_obj.py can pin the repo's own
// OBJ/MTL reader (path_tracing_b200/csrc/b2_obj.cpp) against it.  Output: a flat text dump on stdout:
//   T <n_tris>            then per triangle: 9 position floats (hex), 6 uv floats (hex, 0 if none), has_uv, material id
//   M <n_materials>       then per material: name | Kd3 Pr Pm Ni Ns d illum | map_Kd
// usage: obj_ref file.obj
#define TINYOBJLOADER_IMPLEMENTATION
#include "tiny_obj_loader.h"
#include <cstdio>
#include <cstring>

static unsigned bits(float f) { unsigned u; memcpy(&u, &f, 4); return u; }

int main(int argc, char **argv) {
    if (argc < 2) { fprintf(stderr, "usage: obj_ref file.obj\n"); return 2; }
    tinyobj::ObjReaderConfig cfg;              // defaults: triangulate = true, mtl searched next to the OBJ
    tinyobj::ObjReader rd;
    if (!rd.ParseFromFile(argv[1], cfg)) { fprintf(stderr, "tinyobj: %s\n", rd.Error().c_str()); return 1; }
    const tinyobj::attrib_t &at = rd.GetAttrib();
    size_t n = 0;
    for (const auto &sh : rd.GetShapes()) n += sh.mesh.num_face_vertices.size();
    printf("T %zu\n", n);
    for (const auto &sh : rd.GetShapes()) {
        size_t off = 0;
        for (size_t f = 0; f < sh.mesh.num_face_vertices.size(); ++f) {
            const int nv = sh.mesh.num_face_vertices[f];
            if (nv != 3) { fprintf(stderr, "non-triangle face after triangulation\n"); return 1; }
            float uv[6] = { 0, 0, 0, 0, 0, 0 }; int has_uv = 1;
            for (int k = 0; k < 3; ++k) {
                const tinyobj::index_t ix = sh.mesh.indices[off + k];
                printf("%08x %08x %08x ", bits(at.vertices[3 * ix.vertex_index]), bits(at.vertices[3 * ix.vertex_index + 1]), bits(at.vertices[3 * ix.vertex_index + 2]));
                if (ix.texcoord_index >= 0) { uv[2 * k] = at.texcoords[2 * ix.texcoord_index]; uv[2 * k + 1] = at.texcoords[2 * ix.texcoord_index + 1]; }
                else has_uv = 0;
            }
            for (int k = 0; k < 6; ++k) printf("%08x ", bits(uv[k]));
            printf("%d %d\n", has_uv, sh.mesh.material_ids[f]);
            off += 3;
        }
    }
    const auto &ms = rd.GetMaterials();
    printf("M %zu\n", ms.size());
    for (const auto &m : ms)
        printf("%s | %08x %08x %08x %08x %08x %08x %08x %08x %d | %s\n", m.name.c_str(), bits(m.diffuse[0]), bits(m.diffuse[1]), bits(m.diffuse[2]),
               bits(m.roughness), bits(m.metallic), bits(m.ior), bits(m.shininess), bits(m.dissolve), m.illum, m.diffuse_texname.c_str());
    return 0;
}
