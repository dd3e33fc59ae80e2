// b2_host_scene.hpp -- host-side flattening of the reference's scene objects into the upload format.
// Plays the role of move_data_to_cuda_bdpt (src/bdpt_cu_helper.cpp:13-68) for the new path; shared by
// libb2pt.so (b2_api.cu) and the test-only host build of the device headers (tests/hostsim).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <thread>
#include <vector>
#include "../../include/b2pt.h"
#include "b2_bvh_build.cuh"

// The upload records of the triangles: one uninitialised block (every record is written by flatten_scene), obtained through
// a hook so that libb2pt.so can hand out page-locked memory it keeps across scenes (b2_api.cu: a fresh 50 MB malloc costs
// ~13 000 first-touch page faults, 70 ms for the 1 M-triangle C3 scene whatever the number of threads, and a pageable upload
// is staged by the driver); the test-only host build (tests/hostsim) uses malloc / free.
struct TriBuf {
    typedef void *(*AllocFn)(size_t bytes);
    typedef void (*FreeFn)(void *);
    static AllocFn &alloc_hook() { static AllocFn f = nullptr; return f; }
    static FreeFn &free_hook() { static FreeFn f = nullptr; return f; }
    BuildTri *p = nullptr; size_t n = 0; FreeFn owner_free = nullptr; bool hooked = false;
    TriBuf() = default;
    TriBuf(const TriBuf &) = delete;
    TriBuf &operator=(const TriBuf &) = delete;
    ~TriBuf() { release(); }
    void release() { if (p) { if (hooked) owner_free(p); else free(p); } p = nullptr; n = 0; hooked = false; }
    void resize(size_t count) {                            // contents are NOT kept
        release();
        if (count == 0) return;
        if (alloc_hook() && free_hook()) { p = (BuildTri *) alloc_hook()(count * sizeof(BuildTri)); hooked = p != nullptr; owner_free = free_hook(); }
        if (!p) p = (BuildTri *) malloc(count * sizeof(BuildTri));
        n = p ? count : 0;
    }
    size_t size() const { return n; }
    BuildTri *data() { return p; }
    const BuildTri *data() const { return p; }
    BuildTri &operator[](size_t i) { return p[i]; }
    const BuildTri &operator[](size_t i) const { return p[i]; }
};
struct HostScene {
    TriBuf tris;
    size_t n_tris = 0;       // = tris.size() while the records are held (b2_api.cu drops them after the upload)
    std::vector<float4> sph, lights, mats, groups;
    float3 scene_min, scene_max;
    float max_abs;      // largest |coordinate| / extent of the triangle set (box padding scale)
    int n_sph = 0, n_lights = 0, n_mats = 0, n_groups = 0;
};

inline float4 hs_f4(float x, float y, float z, float w) { return make_float4(x, y, z, w); }
inline float hs_bits(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

// material and light tables (4 + 2 float4 each)
inline void flatten_tables(const B2Material *mats, size_t n_mats, const B2Light *lights, size_t n_lights, HostScene &hs) {
    hs.mats.clear();
    for (size_t i = 0; i < n_mats; ++i) {
        hs.mats.push_back(hs_f4(mats[i].base_color[0], mats[i].base_color[1], mats[i].base_color[2], mats[i].roughness));
        hs.mats.push_back(hs_f4(mats[i].metallic, mats[i].eta, 0.f, 0.f));
    }
    hs.n_mats = (int) n_mats;
    hs.lights.clear();
    for (size_t i = 0; i < n_lights; ++i) {
        const B2Light &L = lights[i];
        float len = std::sqrt(L.dir[0] * L.dir[0] + L.dir[1] * L.dir[1] + L.dir[2] * L.dir[2]);   // geometric.cuh:97-98
        hs.lights.push_back(hs_f4(L.pos[0], L.pos[1], L.pos[2], L.ball_r));
        hs.lights.push_back(hs_f4(L.dir[0], L.dir[1], L.dir[2], L.cutoff));
        hs.lights.push_back(hs_f4(L.illum[0], L.illum[1], L.illum[2], L.is_parallel ? 1.0f : 0.0f));
        hs.lights.push_back(hs_f4(L.dir[0] / len, L.dir[1] / len, L.dir[2] / len, cosf(L.cutoff)));
    }
    hs.n_lights = (int) n_lights;
}

// Triangle-only scenes: the same group boxes, visiting ranks and upload records as the general path below, in two passes over
// the triangles split over a few threads.  The general path keeps a group pointer and a sort key per object (32 bytes each) and
// zero-fills the upload array first: 82 ms for the 1 M-triangle C3 scene on one core, most of the 75 ms bench.py's end-to-end
// step spends outside the render.  Returns false without *err when object ids are not ascending inside some group (the general
// path then sorts them); false with *err on a bad material index.
inline bool flatten_triangles_fast(const B2Tri *tris, size_t n_tris, const B2Material *mats, size_t n_mats, HostScene &hs, const char **err) {
    struct G { float mn[3] = { 99999.f, 99999.f, 99999.f }, mx[3] = { -99999.f, -99999.f, -99999.f }; size_t count = 0; int first_id = 0, last_id = 0; };
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    int T = (int) std::min<size_t>(std::min(8u, hw), std::max<size_t>(1, n_tris >> 16));
    if (const char *e = getenv("B2_FLATTEN_THREADS")) T = std::max(1, std::min(64, atoi(e)));
    std::vector<std::map<int, G>> part(T);
    std::vector<int> bad(T, 0);                           // 1 = material index out of range, 2 = ids not ascending
    auto chunk = [&](int c, size_t &b, size_t &e) { b = n_tris * (size_t) c / (size_t) T; e = n_tris * (size_t) (c + 1) / (size_t) T; };
    auto pass1 = [&](int c) {
        size_t b, e; chunk(c, b, e);
        std::map<int, G> &gm = part[c];
        int last_key = 0; G *last = nullptr;
        for (size_t i = b; i < e; ++i) {
            const B2Tri &t = tris[i];
            if (t.material < 0 || (size_t) t.material >= n_mats) { bad[c] = 1; return; }
            if (!last || t.group != last_key) { last_key = t.group; last = &gm[last_key]; }
            G &g = *last;
            if (g.count == 0) g.first_id = t.obj_id; else if (t.obj_id < g.last_id) bad[c] = 2;
            g.last_id = t.obj_id; g.count++;
            const float *vs[3] = { t.v0, t.v1, t.v2 };
            for (int k = 0; k < 3; ++k)
                for (int a = 0; a < 3; ++a) { g.mn[a] = std::min(g.mn[a], vs[k][a]); g.mx[a] = std::max(g.mx[a], vs[k][a]); }
        }
    };
    auto run = [&](auto &&fn) {
        std::vector<std::thread> th;
        int started = 1;                                   // chunks [1, started) run on their own threads
        try { for (; started < T; ++started) th.emplace_back(fn, started); } catch (...) {}   // no more threads: the rest runs here
        fn(0);
        for (int c = started; c < T; ++c) fn(c);
        for (auto &t : th) t.join();
    };
    run(pass1);
    for (int c = 0; c < T; ++c) if (bad[c] == 1) { *err = "triangle material index out of range"; return false; }
    for (int c = 0; c < T; ++c) if (bad[c] == 2) return false;
    // merge in chunk order: bounds, counts, ascending ids across chunk borders; the rank a chunk's first triangle of a group gets
    std::map<int, G> gm;
    std::vector<std::map<int, size_t>> before(T);          // chunk, group key -> triangles of that group in earlier chunks
    for (int c = 0; c < T; ++c)
        for (auto &kv : part[c]) {
            G &g = gm[kv.first]; const G &p = kv.second;
            if (g.count > 0 && p.first_id < g.last_id) return false;
            if (g.count == 0) g.first_id = p.first_id;
            before[c][kv.first] = g.count;
            g.last_id = p.last_id; g.count += p.count;
            for (int a = 0; a < 3; ++a) { g.mn[a] = std::min(g.mn[a], p.mn[a]); g.mx[a] = std::max(g.mx[a], p.mx[a]); }
        }
    float smn[3] = { 1e9f, 1e9f, 1e9f }, smx[3] = { -1e9f, -1e9f, -1e9f };
    hs.groups.clear();
    std::map<int, std::pair<uint32_t, size_t>> slot;        // group key -> (dense index, first rank)
    {
        size_t acc = 0; uint32_t dense = 0;
        for (auto &kv : gm) {
            slot[kv.first] = std::make_pair(dense++, acc); acc += kv.second.count;
            G g = kv.second;
            for (int a = 0; a < 3; ++a) { smn[a] = std::min(smn[a], g.mn[a]); smx[a] = std::max(smx[a], g.mx[a]); }
            for (int a = 0; a < 3; ++a)                    // object.cpp:108-111 applied to its fixed point, as in the general path
                for (int it = 0; it < 64 && g.mx[a] - g.mn[a] < 1e-6f; ++it) { g.mn[a] -= 0.5f * 1e-6f; g.mx[a] += 0.5f * 1e-6f; }
            hs.groups.push_back(hs_f4(g.mn[0], g.mn[1], g.mn[2], 0.f));
            hs.groups.push_back(hs_f4(g.mx[0], g.mx[1], g.mx[2], 0.f));
        }
    }
    hs.n_groups = (int) gm.size();
    hs.scene_min = make_float3(smn[0], smn[1], smn[2]);
    hs.scene_max = make_float3(smx[0], smx[1], smx[2]);
    hs.tris.resize(n_tris); hs.n_tris = n_tris;            // uninitialised: every record is written below
    if (n_tris && !hs.tris.data()) { *err = "out of host memory"; return false; }
    std::vector<float> cmax(T, 0.f);
    auto pass2 = [&](int c) {
        size_t b, e; chunk(c, b, e);
        struct Cur { uint32_t dense; size_t next; };
        std::map<int, Cur> cur;
        for (auto &kv : before[c]) { const auto &sl = slot[kv.first]; cur[kv.first] = Cur{ sl.first, sl.second + kv.second }; }
        int last_key = 0; Cur *last = nullptr;
        float max_abs = 0.f;
        for (size_t i = b; i < e; ++i) {
            const B2Tri &src = tris[i];
            if (!last || src.group != last_key) { last_key = src.group; last = &cur[last_key]; }
            BuildTri &t = hs.tris[i];
            memcpy(t.v, src.v0, 12); memcpy(t.v + 3, src.v1, 12); memcpy(t.v + 6, src.v2, 12);
            for (int k = 0; k < 9; ++k) max_abs = std::max(max_abs, std::fabs(t.v[k]));
            t.order = (uint32_t) last->next++;
            t.matflag = (uint32_t) src.material | (mats[src.material].eta > 0.0f ? B2_MAT_TRANSPARENT : 0u);
            t.group = last->dense;
            t.obj_id = src.obj_id;
        }
        cmax[c] = max_abs;
    };
    run(pass2);
    hs.max_abs = *std::max_element(cmax.begin(), cmax.end());
    return true;
}

inline bool flatten_scene(const B2Tri *tris, size_t n_tris, const B2Sphere *sph, size_t n_sph, const B2Material *mats, size_t n_mats,
                          const B2Light *lights, size_t n_lights, HostScene &hs, const char **err) {
    *err = nullptr;
    const char *force_general = getenv("B2_FLATTEN_GENERAL");                        // tests: the general path on a scene the fast one would take
    if (n_sph == 0 && n_tris >= (1u << 15) && !(force_general && atoi(force_general) != 0)) {            // large triangle-only scenes (OBJ meshes, generated scenes): several threads, no per-object tables
        if (flatten_triangles_fast(tris, n_tris, mats, n_mats, hs, err)) {
            hs.sph.clear(); hs.n_sph = 0;
            flatten_tables(mats, n_mats, lights, n_lights, hs);
            return true;
        }
        if (*err) return false;                            // (otherwise: object ids not ascending inside a group -> the general path)
    }
    // groups: std::map<int,AABB> in key order; bounds grown as AABB::add_obj does (object.cpp:123-147).
    // The per-object work is O(1): the group of consecutive objects rarely changes, so the map is only consulted when it
    // does, and every object remembers a pointer-stable slot (10 M triangles: 4.0 s -> 0.5 s).
    struct G { float mn[3] = { 99999.f, 99999.f, 99999.f }, mx[3] = { -99999.f, -99999.f, -99999.f }; int dense = -1; size_t count = 0; };
    std::map<int, G> gm;
    std::vector<G *> tri_g(n_tris), sph_g(n_sph);
    {
        int last_key = 0; G *last = nullptr;
        for (size_t i = 0; i < n_tris; ++i) {
            if (tris[i].material < 0 || (size_t) tris[i].material >= n_mats) { *err = "triangle material index out of range"; return false; }
            if (!last || tris[i].group != last_key) { last_key = tris[i].group; last = &gm[last_key]; }
            G &g = *last; tri_g[i] = last; g.count++;
            const float *vs[3] = { tris[i].v0, tris[i].v1, tris[i].v2 };
            for (int k = 0; k < 3; ++k)
                for (int a = 0; a < 3; ++a) { g.mn[a] = std::min(g.mn[a], vs[k][a]); g.mx[a] = std::max(g.mx[a], vs[k][a]); }
        }
        for (size_t i = 0; i < n_sph; ++i) {
            if (sph[i].material < 0 || (size_t) sph[i].material >= n_mats) { *err = "sphere material index out of range"; return false; }
            if (!last || sph[i].group != last_key) { last_key = sph[i].group; last = &gm[last_key]; }
            G &g = *last; sph_g[i] = last; g.count++;
            for (int a = 0; a < 3; ++a) {
                float p = sph[i].center[a] + sph[i].r, m = sph[i].center[a] - sph[i].r;
                g.mn[a] = std::min({ g.mn[a], p, m }); g.mx[a] = std::max({ g.mx[a], p, m });
            }
        }
    }
    // scene bounds for parallel lights: union of the group boxes as run_cpu_bdpt sees them (cpu_bdpt.cpp:181-187)
    float smn[3] = { 1e9f, 1e9f, 1e9f }, smx[3] = { -1e9f, -1e9f, -1e9f };
    hs.groups.clear();
    std::vector<size_t> g_start;                       // first rank of every group, in key order
    {
        size_t acc = 0; int dense = 0;
        for (auto &kv : gm) {
            kv.second.dense = dense++;
            g_start.push_back(acc); acc += kv.second.count;
            G g = kv.second;
            for (int a = 0; a < 3; ++a) { smn[a] = std::min(smn[a], g.mn[a]); smx[a] = std::max(smx[a], g.mx[a]); }
            // AABB::intersectAABB widens degenerate extents in place on every call (object.cpp:108-111); apply to the fixed point
            for (int a = 0; a < 3; ++a)
                for (int it = 0; it < 64 && g.mx[a] - g.mn[a] < 1e-6f; ++it) { g.mn[a] -= 0.5f * 1e-6f; g.mx[a] += 0.5f * 1e-6f; }
            hs.groups.push_back(hs_f4(g.mn[0], g.mn[1], g.mn[2], 0.f));
            hs.groups.push_back(hs_f4(g.mx[0], g.mx[1], g.mx[2], 0.f));
        }
    }
    hs.n_groups = (int) gm.size();
    hs.scene_min = make_float3(smn[0], smn[1], smn[2]);
    hs.scene_max = make_float3(smx[0], smx[1], smx[2]);
    // reference visiting order: group key, then insertion order (= obj_id order).  Bucket by group (stable), then sort a
    // bucket by obj_id only if it is not already ascending (it is, for every scene the parser or the OBJ adapter makes).
    struct Key { int obj_id, kind; size_t idx; };
    std::vector<Key> keys(n_tris + n_sph);
    {
        std::vector<size_t> fill(g_start);
        for (size_t i = 0; i < n_tris; ++i) keys[fill[(size_t) tri_g[i]->dense]++] = Key{ tris[i].obj_id, 1, i };
        for (size_t i = 0; i < n_sph; ++i) keys[fill[(size_t) sph_g[i]->dense]++] = Key{ sph[i].obj_id, 0, i };
        // the bucket pass visited triangles before spheres: a group holding both kinds needs the merge by obj_id
        for (size_t gi = 0; gi < g_start.size(); ++gi) {
            const size_t b0 = g_start[gi], b1 = gi + 1 < g_start.size() ? g_start[gi + 1] : keys.size();
            bool sorted = true;
            for (size_t r = b0 + 1; r < b1 && sorted; ++r) sorted = keys[r - 1].obj_id <= keys[r].obj_id;
            if (!sorted) std::stable_sort(keys.begin() + (long) b0, keys.begin() + (long) b1, [](const Key &a, const Key &b) { return a.obj_id < b.obj_id; });
        }
    }
    std::vector<uint32_t> tri_order(n_tris), sph_order(n_sph);
    for (size_t r = 0; r < keys.size(); ++r) (keys[r].kind ? tri_order[keys[r].idx] : sph_order[keys[r].idx]) = (uint32_t) r;

    hs.tris.resize(n_tris); hs.n_tris = n_tris;
    if (n_tris && !hs.tris.data()) { *err = "out of host memory"; return false; }
    float max_abs = 0.f;
    for (size_t i = 0; i < n_tris; ++i) {
        BuildTri &t = hs.tris[i];
        memcpy(t.v, tris[i].v0, 12); memcpy(t.v + 3, tris[i].v1, 12); memcpy(t.v + 6, tris[i].v2, 12);
        for (int k = 0; k < 9; ++k) max_abs = std::max(max_abs, std::fabs(t.v[k]));
        t.order = tri_order[i];
        t.matflag = (uint32_t) tris[i].material | (mats[tris[i].material].eta > 0.0f ? B2_MAT_TRANSPARENT : 0u);
        t.group = (uint32_t) tri_g[i]->dense;
        t.obj_id = tris[i].obj_id;
    }
    hs.max_abs = max_abs;
    hs.sph.clear();
    for (size_t i = 0; i < n_sph; ++i) {
        uint32_t mf = (uint32_t) sph[i].material | (mats[sph[i].material].eta > 0.0f ? B2_MAT_TRANSPARENT : 0u);
        hs.sph.push_back(hs_f4(sph[i].center[0], sph[i].center[1], sph[i].center[2], sph[i].r));
        hs.sph.push_back(hs_f4(hs_bits(sph_order[i]), hs_bits(mf), hs_bits((uint32_t) sph_g[i]->dense), hs_bits((uint32_t) sph[i].obj_id)));
    }
    hs.n_sph = (int) n_sph;
    flatten_tables(mats, n_mats, lights, n_lights, hs);
    return true;
}
