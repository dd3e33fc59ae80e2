// b2_obj.cpp -- Wavefront OBJ / MTL -> the reference's object model (Triangle + Material + `G` groups), host only.
//
// The reference vendors tinyobjloader v2.0.0 (include/tiny_obj_loader.h) but never calls it (src/tiny_obj_loader.cpp
// is its only mention), so there is no adapter to mirror: this file defines one.  The file is READ
// by the reference's vendored loader -- tinyobj::ObjReader::ParseFromFile (tiny_obj_loader.h:546), compiled from where the
// header lies into b2_tinyobj.cpp -- whenever that header was there at build time (B2_OBJ_READER_AUTO / _TINYOBJ).  A
// small own reader (own_read below) stays as the fallback for builds without the header and as B2_OBJ_READER_OWN; it is
// pinned against the loader face for face and material for material on the fixtures AND on the 1 M-triangle C3 mesh
// (tests/test_obj.py).  Both fill the same ObjMesh (b2_obj_mesh.hpp); everything below the readers is shared.
//
// Adapter semantics (all new; chosen so that the UNMODIFIED CPU oracle renders the same scene):
//   * every face becomes Triangle objects in file order, obj_id = running counter; polygons with n > 3 vertices are
//     triangulated by the reader (tinyobj: its own triangulation; own reader: fan (v0, v_i, v_i+1))
//   * Material{base_color, roughness, metallic, eta}: base_color = Kd * texel(map_Kd at the triangle's centroid uv)
//     (nearest texel, v flipped; baked per triangle: the reference Material has no texture slot); roughness = Pr when
//     Pr > 0 or no Ns was given, else sqrt(2 / (Ns + 2)) clamped to [0, 1] (tinyobj's material_t cannot tell an absent
//     Pr from `Pr 0`: both are 0, and an absent Ns is its default 1); metallic = Pm; eta = Ni if illum in {4, 6, 7} or
//     d < 1, else 0 (opaque)
//   * triangles are binned into a g x g grid of spatial `G` groups over the mesh's x-z extent, g = cbrt(n / 2): the
//     reference's own group-AABB cull (cpu_bdpt.cpp:43-44) then costs g*g box tests + ~g/2 groups of n/(g*g) triangles
//     per ray, which that g minimises -- it is what keeps the brute-force CPU oracle usable on big meshes
//   * textures: binary PPM (P6, maxval <= 255) and PFM (PF); anything else is an error, not a silent white
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>
#include "../../include/b2pt.h"
#include "b2_obj_mesh.hpp"

void b2_set_error(const char *fmt, ...);

namespace {

struct Tex { int w = 0, h = 0; std::vector<float> rgb; };
typedef ObjMtl Mtl;

bool read_file(const std::string &path, std::string &out) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? (size_t) n : 0);
    size_t got = n > 0 ? fread(&out[0], 1, (size_t) n, f) : 0;
    fclose(f);
    out.resize(got);
    return true;
}
std::string dir_of(const std::string &p) { size_t k = p.find_last_of('/'); return k == std::string::npos ? std::string() : p.substr(0, k + 1); }
inline bool is_sp(char c) { return c == ' ' || c == '\t' || c == '\r'; }
inline const char *skip_sp(const char *p, const char *e) { while (p < e && is_sp(*p)) ++p; return p; }
inline const char *tok_end(const char *p, const char *e) { while (p < e && !is_sp(*p)) ++p; return p; }
// correctly rounded decimal -> float (glibc strtof), the same value fast_float produces inside tinyobj
inline bool parse_f(const char *&p, const char *e, float &v) {
    p = skip_sp(p, e);
    if (p >= e) return false;
    char buf[64];
    const char *q = tok_end(p, e);
    size_t n = (size_t) (q - p);
    if (n == 0 || n >= sizeof(buf)) return false;
    memcpy(buf, p, n); buf[n] = 0;
    char *end = nullptr;
    v = strtof(buf, &end);
    if (end == buf) return false;
    p = q;
    return true;
}
inline bool parse_i(const char *&p, const char *e, long &v) {
    bool neg = false;
    if (p < e && (*p == '-' || *p == '+')) { neg = *p == '-'; ++p; }
    if (p >= e || *p < '0' || *p > '9') return false;
    long x = 0;
    while (p < e && *p >= '0' && *p <= '9') { x = x * 10 + (*p - '0'); ++p; }
    v = neg ? -x : x;
    return true;
}

bool load_texture(const std::string &path, Tex &t) {
    std::string s;
    if (!read_file(path, s)) { b2_set_error("cannot open texture: %s", path.c_str()); return false; }
    const char *p = s.data(), *e = p + s.size();
    auto next_tok = [&](std::string &out) {
        for (;;) {
            while (p < e && (is_sp(*p) || *p == '\n')) ++p;
            if (p < e && *p == '#') { while (p < e && *p != '\n') ++p; continue; }
            break;
        }
        const char *q = p;
        while (q < e && !is_sp(*q) && *q != '\n') ++q;
        out.assign(p, q); p = q;
        return !out.empty();
    };
    std::string magic, a, b, c;
    if (!next_tok(magic) || !next_tok(a) || !next_tok(b) || !next_tok(c)) { b2_set_error("truncated texture header: %s", path.c_str()); return false; }
    t.w = atoi(a.c_str()); t.h = atoi(b.c_str());
    if (t.w <= 0 || t.h <= 0 || (long long) t.w * t.h > (1ll << 28)) { b2_set_error("bad texture size: %s", path.c_str()); return false; }
    if (p < e) ++p;                                   // the single whitespace byte after the header
    const size_t npx = (size_t) t.w * t.h;
    t.rgb.resize(3 * npx);
    if (magic == "P6") {
        const int maxv = atoi(c.c_str());
        if (maxv <= 0 || maxv > 255) { b2_set_error("PPM maxval %d not supported (8-bit only): %s", maxv, path.c_str()); return false; }
        if ((size_t) (e - p) < 3 * npx) { b2_set_error("truncated PPM: %s", path.c_str()); return false; }
        const unsigned char *u = (const unsigned char *) p;
        for (size_t i = 0; i < 3 * npx; ++i) t.rgb[i] = (float) u[i] / (float) maxv;
    } else if (magic == "PF") {
        const float scale = strtof(c.c_str(), nullptr);
        if (!(scale < 0.0f)) { b2_set_error("big-endian PFM not supported: %s", path.c_str()); return false; }
        if ((size_t) (e - p) < 12 * npx) { b2_set_error("truncated PFM: %s", path.c_str()); return false; }
        for (int y = 0; y < t.h; ++y)                 // PFM rows run bottom to top
            memcpy(&t.rgb[(size_t) 3 * t.w * y], p + (size_t) 12 * t.w * (t.h - 1 - y), (size_t) 12 * t.w);
    } else { b2_set_error("texture format '%s' not supported (binary PPM P6 / PFM PF only): %s", magic.c_str(), path.c_str()); return false; }
    return true;
}
// nearest texel, repeat wrap, OBJ's v axis points up (row 0 of the image is v = 1)
inline void sample_tex(const Tex &t, float u, float v, float out[3]) {
    u -= std::floor(u); v -= std::floor(v);
    int x = (int) (u * (float) t.w), y = (int) ((1.0f - v) * (float) t.h);
    x = x < 0 ? 0 : (x >= t.w ? t.w - 1 : x); y = y < 0 ? 0 : (y >= t.h ? t.h - 1 : y);
    const float *px = &t.rgb[3 * ((size_t) y * t.w + x)];
    out[0] = px[0]; out[1] = px[1]; out[2] = px[2];
}

bool load_mtl(const std::string &path, std::vector<Mtl> &mtls, std::unordered_map<std::string, int> &by_name) {
    std::string s;
    if (!read_file(path, s)) { b2_set_error("cannot open material library: %s", path.c_str()); return false; }
    const char *p = s.data(), *e = p + s.size();
    Mtl *cur = nullptr;
    while (p < e) {
        const char *le = (const char *) memchr(p, '\n', (size_t) (e - p));
        if (!le) le = e;
        const char *q = skip_sp(p, le);
        const char *k = tok_end(q, le);
        std::string key(q, k);
        const char *a = k;
        if (key == "newmtl") {
            a = skip_sp(a, le);
            const char *b = le; while (b > a && is_sp(b[-1])) --b;
            mtls.emplace_back(); cur = &mtls.back(); cur->name.assign(a, b);
            cur->kd[0] = cur->kd[1] = cur->kd[2] = 0.0f;     // tinyobj: material_t starts zeroed
            by_name[cur->name] = (int) mtls.size() - 1;
        } else if (cur) {
            if (key == "Kd") { parse_f(a, le, cur->kd[0]) && parse_f(a, le, cur->kd[1]) && parse_f(a, le, cur->kd[2]); }
            else if (key == "Pr") parse_f(a, le, cur->pr);
            else if (key == "Pm") parse_f(a, le, cur->pm);
            else if (key == "Ni") parse_f(a, le, cur->ni);
            else if (key == "Ns") parse_f(a, le, cur->ns);
            else if (key == "d") parse_f(a, le, cur->d);
            else if (key == "Tr") { float tr; if (parse_f(a, le, tr)) cur->d = 1.0f - tr; }
            else if (key == "illum") { float v; if (parse_f(a, le, v)) cur->illum = (int) v; }
            else if (key == "map_Kd") {                      // options (-o, -s, ...) are not supported: the file name is the last token
                const char *b = le; while (b > a && is_sp(b[-1])) --b;
                const char *n0 = b; while (n0 > a && !is_sp(n0[-1])) --n0;
                cur->map_kd.assign(n0, b);
            }
        }
        p = le < e ? le + 1 : e;
    }
    return true;
}

struct MatKey {
    float v[6];
    bool operator==(const MatKey &o) const { return memcmp(v, o.v, sizeof(v)) == 0; }
};
struct MatHash {
    size_t operator()(const MatKey &k) const {
        uint64_t h = 1469598103934665603ull;
        const unsigned char *b = (const unsigned char *) k.v;
        for (size_t i = 0; i < sizeof(k.v); ++i) { h ^= b[i]; h *= 1099511628211ull; }
        return (size_t) h;
    }
};

}  // namespace

#ifndef B2_HAVE_TINYOBJ        // built without the loader's header (no /root/reference at build time): own reader only
bool b2_tinyobj_available() { return false; }
bool b2_tinyobj_read(const std::string &, ObjMesh &, std::string &err) { err = "tinyobjloader was not compiled into this libb2pt.so"; return false; }
#endif

namespace {
// The own reader: v / vt / f / usemtl / mtllib, every index form (v, v/vt, v/vt/vn, v//vn, negative), polygons fanned.
bool own_read(const std::string &obj_path, ObjMesh &out) {
    std::string s;
    if (!read_file(obj_path, s)) { b2_set_error("cannot open OBJ file: %s", obj_path.c_str()); return false; }
    const std::string base = dir_of(obj_path);
    std::vector<float> &pos = out.pos, &uv = out.uv;
    std::vector<Mtl> &mtls = out.mtls;
    std::unordered_map<std::string, int> by_name;
    std::vector<ObjFace> &faces = out.faces;
    int cur_m = -1;
    std::vector<long> fv, ft;
    const char *p = s.data(), *e = p + s.size();
    size_t line_no = 0;
    while (p < e) {
        const char *le = (const char *) memchr(p, '\n', (size_t) (e - p));
        if (!le) le = e;
        ++line_no;
        const char *q = skip_sp(p, le);
        if (q + 1 < le && q[0] == 'v' && is_sp(q[1])) {
            const char *a = q + 1; float x, y, z;
            if (!(parse_f(a, le, x) && parse_f(a, le, y) && parse_f(a, le, z))) { b2_set_error("%s:%zu: bad vertex", obj_path.c_str(), line_no); return false; }
            pos.push_back(x); pos.push_back(y); pos.push_back(z);
        } else if (q + 2 < le && q[0] == 'v' && q[1] == 't' && is_sp(q[2])) {
            const char *a = q + 2; float u = 0.0f, v = 0.0f;
            if (!parse_f(a, le, u)) { b2_set_error("%s:%zu: bad texture coordinate", obj_path.c_str(), line_no); return false; }
            parse_f(a, le, v);
            uv.push_back(u); uv.push_back(v);
        } else if (q + 1 < le && q[0] == 'f' && is_sp(q[1])) {
            const char *a = q + 1;
            fv.clear(); ft.clear();
            const long nv = (long) pos.size() / 3, nt = (long) uv.size() / 2;
            for (;;) {
                a = skip_sp(a, le);
                if (a >= le) break;
                long vi = 0, ti = 0; bool has_t = false;
                if (!parse_i(a, le, vi)) { b2_set_error("%s:%zu: bad face index", obj_path.c_str(), line_no); return false; }
                if (a < le && *a == '/') {
                    ++a;
                    if (a < le && *a != '/' && !is_sp(*a)) { if (!parse_i(a, le, ti)) { b2_set_error("%s:%zu: bad face index", obj_path.c_str(), line_no); return false; } has_t = true; }
                    if (a < le && *a == '/') { ++a; long ni; parse_i(a, le, ni); }
                }
                vi = vi > 0 ? vi - 1 : nv + vi;             // 1-based, negative = relative to the end
                ti = has_t ? (ti > 0 ? ti - 1 : nt + ti) : -1;
                if (vi < 0 || vi >= nv || (has_t && (ti < 0 || ti >= nt))) { b2_set_error("%s:%zu: face index out of range", obj_path.c_str(), line_no); return false; }
                fv.push_back(vi); ft.push_back(ti);
            }
            if (fv.size() < 3) { b2_set_error("%s:%zu: face with fewer than 3 vertices", obj_path.c_str(), line_no); return false; }
            for (size_t k = 1; k + 1 < fv.size(); ++k)
                faces.push_back(ObjFace{ { (int) fv[0], (int) fv[k], (int) fv[k + 1] }, { (int) ft[0], (int) ft[k], (int) ft[k + 1] }, cur_m });
        } else if (le - q > 7 && memcmp(q, "usemtl", 6) == 0 && is_sp(q[6])) {
            const char *a = skip_sp(q + 6, le), *b = le; while (b > a && is_sp(b[-1])) --b;
            auto it = by_name.find(std::string(a, b));
            cur_m = it == by_name.end() ? -1 : it->second;
        } else if (le - q > 7 && memcmp(q, "mtllib", 6) == 0 && is_sp(q[6])) {
            const char *a = skip_sp(q + 6, le), *b = le; while (b > a && is_sp(b[-1])) --b;
            if (!load_mtl(base + std::string(a, b), mtls, by_name)) return false;
        }
        p = le < e ? le + 1 : e;
    }
    return true;
}
}  // namespace

// Appends the mesh to the vectors (so the scene-file parser can splice an `O file.obj` directive into a scene).
bool b2_obj_append(const std::string &obj_path, const B2ObjOptions &opt, std::vector<B2Tri> &tris, std::vector<B2Material> &mats,
                   int &obj_id, int &next_group) {
    ObjMesh mesh;
    const int reader = opt.reader;
    if (reader < B2_OBJ_READER_AUTO || reader > B2_OBJ_READER_TINYOBJ) { b2_set_error("unknown OBJ reader %d", reader); return false; }
    if (reader == B2_OBJ_READER_TINYOBJ || (reader == B2_OBJ_READER_AUTO && b2_tinyobj_available())) {
        FILE *probe = fopen(obj_path.c_str(), "rb");
        if (!probe) { b2_set_error("cannot open OBJ file: %s", obj_path.c_str()); return false; }
        fclose(probe);
        std::string err;
        if (!b2_tinyobj_read(obj_path, mesh, err)) { b2_set_error("tinyobj: %s (%s)", err.c_str(), obj_path.c_str()); return false; }
    } else if (!own_read(obj_path, mesh)) return false;
    const std::string base = dir_of(obj_path);
    std::vector<float> &pos = mesh.pos, &uv = mesh.uv;
    std::vector<Mtl> &mtls = mesh.mtls;
    std::vector<ObjFace> &faces = mesh.faces;
    std::vector<Tex> texs;
    std::unordered_map<std::string, int> tex_by_name;
    for (Mtl &m : mtls)
        if (!m.map_kd.empty()) {
            auto it = tex_by_name.find(m.map_kd);
            if (it == tex_by_name.end()) {
                texs.emplace_back();
                if (!load_texture(base + m.map_kd, texs.back())) return false;
                it = tex_by_name.emplace(m.map_kd, (int) texs.size() - 1).first;
            }
            m.tex = it->second;
        }
    // spatial groups over the x-z extent
    const size_t n = faces.size();
    int g = opt.group_grid > 0 ? opt.group_grid : (int) std::lround(std::cbrt((double) n / 2.0));
    g = g < 1 ? 1 : (g > 256 ? 256 : g);
    float mn[3] = { 1e30f, 1e30f, 1e30f }, mx[3] = { -1e30f, -1e30f, -1e30f };
    for (size_t i = 0; i < pos.size(); ++i) { const int a = (int) (i % 3); mn[a] = std::fmin(mn[a], pos[i]); mx[a] = std::fmax(mx[a], pos[i]); }
    const float ex = mx[0] - mn[0] > 0.0f ? mx[0] - mn[0] : 1.0f, ez = mx[2] - mn[2] > 0.0f ? mx[2] - mn[2] : 1.0f;
    std::unordered_map<MatKey, int, MatHash> mat_index;
    // materials already in the table keep their indices; identical values are shared
    for (size_t i = 0; i < mats.size(); ++i) {
        MatKey k{ { mats[i].base_color[0], mats[i].base_color[1], mats[i].base_color[2], mats[i].roughness, mats[i].metallic, mats[i].eta } };
        mat_index.emplace(k, (int) i);
    }
    const int group0 = next_group;
    tris.reserve(tris.size() + n);
    for (size_t i = 0; i < n; ++i) {
        const ObjFace &f = faces[i];
        B2Tri t; memset(&t, 0, sizeof(t));
        memcpy(t.v0, &pos[3 * (size_t) f.v[0]], 12); memcpy(t.v1, &pos[3 * (size_t) f.v[1]], 12); memcpy(t.v2, &pos[3 * (size_t) f.v[2]], 12);
        MatKey k;
        if (f.m >= 0) {
            const Mtl &m = mtls[(size_t) f.m];
            float tx[3] = { 1.0f, 1.0f, 1.0f };
            if (m.tex >= 0 && f.t[0] >= 0 && f.t[1] >= 0 && f.t[2] >= 0) {
                const float u = (uv[2 * (size_t) f.t[0]] + uv[2 * (size_t) f.t[1]] + uv[2 * (size_t) f.t[2]]) / 3.0f;
                const float v = (uv[2 * (size_t) f.t[0] + 1] + uv[2 * (size_t) f.t[1] + 1] + uv[2 * (size_t) f.t[2] + 1]) / 3.0f;
                sample_tex(texs[(size_t) m.tex], u, v, tx);
            }
            float rough = (m.pr > 0.0f || m.ns == 1.0f) ? m.pr : std::sqrt(2.0f / (m.ns + 2.0f));
            rough = rough < 0.0f ? 0.0f : (rough > 1.0f ? 1.0f : rough);
            const bool transparent = m.illum == 4 || m.illum == 6 || m.illum == 7 || m.d < 1.0f;
            k = MatKey{ { m.kd[0] * tx[0], m.kd[1] * tx[1], m.kd[2] * tx[2], rough, m.pm, transparent ? m.ni : 0.0f } };
        } else {
            k = MatKey{ { opt.default_color[0], opt.default_color[1], opt.default_color[2], opt.default_roughness, 0.0f, 0.0f } };
        }
        auto it = mat_index.find(k);
        if (it == mat_index.end()) {
            B2Material bm; memset(&bm, 0, sizeof(bm));
            bm.base_color[0] = k.v[0]; bm.base_color[1] = k.v[1]; bm.base_color[2] = k.v[2]; bm.roughness = k.v[3]; bm.metallic = k.v[4]; bm.eta = k.v[5];
            mats.push_back(bm);
            it = mat_index.emplace(k, (int) mats.size() - 1).first;
        }
        t.material = it->second;
        t.obj_id = obj_id++;
        const float cx = (t.v0[0] + t.v1[0] + t.v2[0]) / 3.0f, cz = (t.v0[2] + t.v1[2] + t.v2[2]) / 3.0f;
        int gx = (int) ((cx - mn[0]) / ex * (float) g), gz = (int) ((cz - mn[2]) / ez * (float) g);
        gx = gx < 0 ? 0 : (gx >= g ? g - 1 : gx); gz = gz < 0 ? 0 : (gz >= g ? g - 1 : gz);
        t.group = group0 + gx * g + gz;
        tris.push_back(t);
    }
    next_group = group0 + g * g;
    return true;
}

static B2ParsedScene *pack_parsed(const std::vector<B2Tri> &tris, const std::vector<B2Material> &mats) {
    B2ParsedScene *P = (B2ParsedScene *) calloc(1, sizeof(B2ParsedScene));
    if (!P) return nullptr;
    auto dup = [](const void *src, size_t bytes) -> void * { void *p = malloc(bytes ? bytes : 1); if (p && bytes) memcpy(p, src, bytes); return p; };
    P->n_tris = tris.size(); P->n_materials = mats.size();
    P->tris = (B2Tri *) dup(tris.data(), tris.size() * sizeof(B2Tri));
    P->materials = (B2Material *) dup(mats.data(), mats.size() * sizeof(B2Material));
    P->spheres = (B2Sphere *) dup(nullptr, 0); P->lights = (B2Light *) dup(nullptr, 0);
    return P;
}

extern "C" void b2_obj_default_options(B2ObjOptions *opt) {
    if (!opt) return;
    memset(opt, 0, sizeof(*opt));
    opt->default_color[0] = opt->default_color[1] = opt->default_color[2] = 0.8f;
    opt->default_roughness = 1.0f;
}

extern "C" int b2_obj_has_tinyobj(void) { return b2_tinyobj_available() ? 1 : 0; }

extern "C" int b2_load_obj(const char *obj_path, const B2ObjOptions *opt, B2ParsedScene **out) {
    if (!obj_path || !out) { b2_set_error("NULL argument"); return B2_ERR_INVALID; }
    *out = nullptr;
    B2ObjOptions o;
    if (opt) o = *opt; else b2_obj_default_options(&o);
    std::vector<B2Tri> tris; std::vector<B2Material> mats;
    int obj_id = o.first_obj_id, next_group = o.first_group;
    if (!b2_obj_append(obj_path, o, tris, mats, obj_id, next_group)) return B2_ERR_IO;
    B2ParsedScene *P = pack_parsed(tris, mats);
    if (!P) { b2_set_error("out of memory"); return B2_ERR_OOM; }
    *out = P;
    return B2_OK;
}
