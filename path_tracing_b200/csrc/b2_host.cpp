// b2_host.cpp -- host-only entry points of include/b2pt.h: the scene-text grammar and the film writers.
// These mirror the reference's surface around the hot path (they do no rendering work):
//   b2_parse_scene_file  <- the `while(input >> t)` loop of src/main_cli.cpp:99-141
//   b2_film_to_rgb8      <- clamp -> pow(1/2.2) -> *255 truncated, src/main_cli.cpp:225-243
//   b2_write_image       <- cv::imwrite (OpenCV is not available here): PPM (P6), PNG (zlib) and PFM writers
#include <zlib.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>
#include "../../include/b2pt.h"

void b2_set_error(const char *fmt, ...);

namespace {
struct V3 { float x, y, z; };
// formatted extraction exactly as the reference's operator>>(istream&, vec3&) (object.cpp:11-14)
std::istream &operator>>(std::istream &is, V3 &v) { is >> v.x >> v.y >> v.z; return is; }
void put3(float *d, const V3 &v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }
}  // namespace

bool b2_obj_append(const std::string &obj_path, const B2ObjOptions &opt, std::vector<B2Tri> &tris, std::vector<B2Material> &mats,
                   int &obj_id, int &next_group);

extern "C" int b2_parse_scene_file(const char *path, B2ParsedScene **out) { return b2_parse_scene_file_ex(path, 0, out); }

extern "C" int b2_parse_scene_file_ex(const char *path, int flags, B2ParsedScene **out) {
    if (!path || !out) { b2_set_error("NULL argument"); return B2_ERR_INVALID; }
    *out = nullptr;
    std::ifstream input(path);
    if (!input.is_open()) { b2_set_error("cannot open input file: %s", path); return B2_ERR_IO; }   // main_cli.cpp:86-89
    std::vector<B2Tri> tris; std::vector<B2Sphere> sph; std::vector<B2Material> mats; std::vector<B2Light> lights;
    B2Camera cam; memset(&cam, 0, sizeof(cam));
    int W = 0, H = 0;
    B2Material mtl; memset(&mtl, 0, sizeof(mtl));     // `Material mtl;` default-initialised members (object.h:28-33)
    int cur_mat = -1;                                  // index of `mtl` in the table, -1 = not stored yet
    int group_id = 0, obj_id = 0;
    auto mat_index = [&]() { if (cur_mat < 0) { mats.push_back(mtl); cur_mat = (int) mats.size() - 1; } return cur_mat; };
    char t;
    while (input >> t) {
        if (t == 'E') { V3 v{ cam.eye[0], cam.eye[1], cam.eye[2] }; input >> v; put3(cam.eye, v); }
        else if (t == 'V') {
            V3 a{ cam.look_at[0], cam.look_at[1], cam.look_at[2] }, b{ cam.view_up[0], cam.view_up[1], cam.view_up[2] };
            input >> a >> b; put3(cam.look_at, a); put3(cam.view_up, b);
        }
        else if (t == 'F') { input >> cam.fov; }
        else if (t == 'R') { input >> W >> H; }
        else if (t == 'S') {
            B2Sphere s; memset(&s, 0, sizeof(s));
            V3 c{ 0, 0, 0 }; input >> c >> s.r; put3(s.center, c);
            s.material = mat_index(); s.obj_id = obj_id++; s.group = group_id;
            sph.push_back(s);
        }
        else if (t == 'T') {
            B2Tri tr; memset(&tr, 0, sizeof(tr));
            V3 a{ 0, 0, 0 }, b{ 0, 0, 0 }, c{ 0, 0, 0 }; input >> a >> b >> c; put3(tr.v0, a); put3(tr.v1, b); put3(tr.v2, c);
            tr.material = mat_index(); tr.obj_id = obj_id++; tr.group = group_id;
            tris.push_back(tr);
        }
        else if (t == 'M') {
            V3 c{ mtl.base_color[0], mtl.base_color[1], mtl.base_color[2] };
            input >> c >> mtl.roughness >> mtl.metallic >> mtl.eta; put3(mtl.base_color, c);
            cur_mat = -1;
        }
        else if (t == 'G') { input >> group_id; }
        else if (t == 'O' && (flags & B2_PARSE_OBJ_DIRECTIVE)) {      // new directive, not part of the reference grammar
            std::string rel; input >> rel;
            std::string p(path); const size_t k = p.find_last_of('/');
            const std::string full = (rel.empty() || rel[0] == '/' || k == std::string::npos) ? rel : p.substr(0, k + 1) + rel;
            B2ObjOptions o; b2_obj_default_options(&o);
            o.default_color[0] = mtl.base_color[0]; o.default_color[1] = mtl.base_color[1]; o.default_color[2] = mtl.base_color[2];
            o.default_roughness = mtl.roughness;
            if (!b2_obj_append(full, o, tris, mats, obj_id, group_id)) return B2_ERR_IO;
        }
        else if (t == '/') {
            input >> t;
            if (t == '/') { std::string trash; std::getline(input, trash); continue; }
        }
        else if (t == 'L') {
            B2Light L; memset(&L, 0, sizeof(L));
            V3 p{ 0, 0, 0 }, d{ 0, 0, 0 }, il{ 0, 0, 0 }; float cutoff_deg = 0.0f;
            input >> p >> d >> il >> cutoff_deg;
            put3(L.pos, p); put3(L.dir, d); put3(L.illum, il);
            L.cutoff = cutoff_deg * 0.01745329251994329576923690768489f;      // glm::radians, main_cli.cpp:135
            input >> L.is_parallel >> L.ball_r;
            lights.push_back(L);
        }
    }
    B2ParsedScene *P = (B2ParsedScene *) calloc(1, sizeof(B2ParsedScene));
    if (!P) { b2_set_error("out of memory"); return B2_ERR_OOM; }
    auto dup = [](const void *src, size_t bytes) -> void * { void *p = malloc(bytes ? bytes : 1); if (p && bytes) memcpy(p, src, bytes); return p; };
    P->n_tris = tris.size(); P->n_spheres = sph.size(); P->n_materials = mats.size(); P->n_lights = lights.size();
    P->tris = (B2Tri *) dup(tris.data(), tris.size() * sizeof(B2Tri));
    P->spheres = (B2Sphere *) dup(sph.data(), sph.size() * sizeof(B2Sphere));
    P->materials = (B2Material *) dup(mats.data(), mats.size() * sizeof(B2Material));
    P->lights = (B2Light *) dup(lights.data(), lights.size() * sizeof(B2Light));
    P->camera = cam; P->width = W; P->height = H;
    *out = P;
    return B2_OK;
}

extern "C" void b2_parsed_scene_free(B2ParsedScene *p) {
    if (!p) return;
    free(p->tris); free(p->spheres); free(p->materials); free(p->lights); free(p);
}

extern "C" int b2_film_to_rgb8(const float *film, int32_t W, int32_t H, uint8_t *rgb8) {
    if (!film || !rgb8 || W <= 0 || H <= 0) { b2_set_error("invalid argument"); return B2_ERR_INVALID; }
    const size_t n = (size_t) 3 * W * H;
    for (size_t i = 0; i < n; ++i) {
        // std::pow(std::max(0.0f, std::min(x, 1.0f)), 1.0f / 2.2f) * 255 truncated; NaN -> 0 through the max()
        const float x = film[i];
        const float m = (1.0f < x) ? 1.0f : x;
        const float c = (0.0f < m) ? m : 0.0f;
        rgb8[i] = (uint8_t) (std::pow(c, 1.0f / 2.2f) * 255.0f);
    }
    return B2_OK;
}

// the viewer's convergence measure between two displayed frames (main.cpp:502-531): sqrt(sum (cur - last)^2) / 255
extern "C" int b2_frame_rms(const uint8_t *cur, const uint8_t *last, size_t n, float *out) {
    if (!cur || !last || !out) { b2_set_error("NULL argument"); return B2_ERR_INVALID; }
    float acc = 0.0f;                                   // float accumulation and powf-free squares of small integers: exact per term
    for (size_t i = 0; i < n; ++i) { const float d = (float) cur[i] - (float) last[i]; acc += d * d; }
    *out = std::sqrt(acc) / 255.0f;
    return B2_OK;
}

namespace {
void be32(std::vector<uint8_t> &v, uint32_t x) { v.push_back(x >> 24); v.push_back(x >> 16); v.push_back(x >> 8); v.push_back(x); }
void png_chunk(std::vector<uint8_t> &out, const char *type, const uint8_t *data, size_t n) {
    be32(out, (uint32_t) n);
    const size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    if (n) out.insert(out.end(), data, data + n);
    be32(out, (uint32_t) crc32(0L, out.data() + start, (uInt) (n + 4)));
}
bool ends_with(const std::string &s, const char *suf) {
    const size_t n = strlen(suf);
    if (s.size() < n) return false;
    for (size_t i = 0; i < n; ++i) if (tolower((unsigned char) s[s.size() - n + i]) != suf[i]) return false;
    return true;
}
}  // namespace

extern "C" int b2_write_image(const char *path, const float *film, int32_t W, int32_t H) {
    if (!path || !film || W <= 0 || H <= 0) { b2_set_error("invalid argument"); return B2_ERR_INVALID; }
    const std::string p(path);
    FILE *f = nullptr;
    if (ends_with(p, ".pfm")) {
        f = fopen(path, "wb");
        if (!f) { b2_set_error("cannot open %s for writing", path); return B2_ERR_IO; }
        fprintf(f, "PF\n%d %d\n-1.0\n", W, H);                         // little endian, rows bottom-to-top
        for (int y = H - 1; y >= 0; --y) fwrite(film + (size_t) 3 * W * y, sizeof(float), (size_t) 3 * W, f);
        fclose(f);
        return B2_OK;
    }
    std::vector<uint8_t> rgb((size_t) 3 * W * H);
    int rc = b2_film_to_rgb8(film, W, H, rgb.data());
    if (rc != B2_OK) return rc;
    if (ends_with(p, ".ppm")) {
        f = fopen(path, "wb");
        if (!f) { b2_set_error("cannot open %s for writing", path); return B2_ERR_IO; }
        fprintf(f, "P6\n%d %d\n255\n", W, H);
        fwrite(rgb.data(), 1, rgb.size(), f);
        fclose(f);
        return B2_OK;
    }
    if (ends_with(p, ".png")) {
        std::vector<uint8_t> raw((size_t) H * (1 + (size_t) 3 * W));
        for (int y = 0; y < H; ++y) {
            raw[(size_t) y * (1 + 3 * W)] = 0;                          // filter type None
            memcpy(&raw[(size_t) y * (1 + 3 * W) + 1], &rgb[(size_t) 3 * W * y], (size_t) 3 * W);
        }
        uLongf zlen = compressBound((uLong) raw.size());
        std::vector<uint8_t> z(zlen);
        if (compress2(z.data(), &zlen, raw.data(), (uLong) raw.size(), 6) != Z_OK) { b2_set_error("zlib compress failed"); return B2_ERR_IO; }
        std::vector<uint8_t> out = { 0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a };
        std::vector<uint8_t> ihdr;
        be32(ihdr, (uint32_t) W); be32(ihdr, (uint32_t) H);
        ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);   // 8-bit RGB
        png_chunk(out, "IHDR", ihdr.data(), ihdr.size());
        png_chunk(out, "IDAT", z.data(), zlen);
        png_chunk(out, "IEND", nullptr, 0);
        f = fopen(path, "wb");
        if (!f) { b2_set_error("cannot open %s for writing", path); return B2_ERR_IO; }
        fwrite(out.data(), 1, out.size(), f);
        fclose(f);
        return B2_OK;
    }
    b2_set_error("unsupported image extension (use .png, .ppm or .pfm): %s", path);
    return B2_ERR_INVALID;
}
