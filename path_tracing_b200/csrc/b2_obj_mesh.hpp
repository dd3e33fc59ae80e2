// b2_obj_mesh.hpp -- what an OBJ/MTL reader hands to the adapter of b2_obj.cpp (host only).
// Two readers fill it: the reference's vendored tinyobjloader (b2_tinyobj.cpp, the default) and the repo's own small
// reader (b2_obj.cpp: used when the loader's header was not there at build time, and pinned against it in tests/test_obj.py).
#pragma once
#include <string>
#include <vector>

struct ObjMtl {
    std::string name, map_kd;
    float kd[3] = { 0.0f, 0.0f, 0.0f };      // tinyobj::material_t starts zeroed: zero until a Kd line sets it
    float pr = 0.0f, pm = 0.0f, ni = 1.0f, ns = 1.0f, d = 1.0f;   // InitMaterial defaults (tiny_obj_loader.h:6980-7020)
    int illum = 0;
    int tex = -1;                             // filled by the adapter
};
struct ObjFace { int v[3]; int t[3]; int m; };   // vertex / texcoord indices (-1 = no texcoord), material id (-1 = none)
struct ObjMesh {
    std::vector<float> pos, uv;               // xyz / uv, flat
    std::vector<ObjFace> faces;               // triangles, file order
    std::vector<ObjMtl> mtls;
};

#define B2_OBJ_READER_AUTO 0      // tinyobjloader when it is compiled in, else the own reader
#define B2_OBJ_READER_OWN 1
#define B2_OBJ_READER_TINYOBJ 2

// b2_tinyobj.cpp (or its stub in b2_obj.cpp when the header was not found at build time)
bool b2_tinyobj_available();
bool b2_tinyobj_read(const std::string &obj_path, ObjMesh &out, std::string &err);
