// tests/hostsim/hostsim_bvh.hpp -- TEST-ONLY sequential host driver for the BVH build bodies
// (path_tracing_b200/csrc/b2_bvh_build.cuh).  Mirrors the level loop of b2_bvh_build.cu so the
// build/traversal logic can be debugged without a GPU.  Never linked into libb2pt.so.
#pragma once
#include <algorithm>
#include <numeric>
#include <vector>
#include "../../path_tracing_b200/csrc/b2_host_scene.hpp"
#include "../../path_tracing_b200/csrc/b2_bvh.cuh"

struct HostBvh {
    std::vector<uint4> nodes;
    std::vector<float4> tri;
    std::vector<int32_t> tri_obj;
    int depth = 0;
    float sah_cost = 0.0f;      // cost of the chosen wide tree / root area (C_node = 1)
};

inline void host_build_bvh(const HostScene &hs, int max_leaf, float pad_rel, HostBvh &out, int builder = B2_BUILDER_PLOC, int ploc_radius = 2) {
    const int n = (int) hs.tris.size();
    out.nodes.clear(); out.tri.assign(3 * (size_t) n, make_float4(0, 0, 0, 0)); out.tri_obj.assign(n, 0); out.depth = 0;
    if (n == 0) return;
    float ext = std::max({ hs.scene_max.x - hs.scene_min.x, hs.scene_max.y - hs.scene_min.y, hs.scene_max.z - hs.scene_min.z });
    float pad = pad_rel * std::max(hs.max_abs, ext) + 1e-30f;
    std::vector<float> tlo(3 * n), thi(3 * n);
    float cmn[3] = { 3e38f, 3e38f, 3e38f }, cmx[3] = { -3e38f, -3e38f, -3e38f };
    for (int i = 0; i < n; ++i) {
        tri_bounds(hs.tris[i], pad, &tlo[3 * i], &thi[3 * i]);
        for (int a = 0; a < 3; ++a) { float c = 0.5f * (tlo[3 * i + a] + thi[3 * i + a]); cmn[a] = std::min(cmn[a], c); cmx[a] = std::max(cmx[a], c); }
    }
    float3 cmin = make_float3(cmn[0], cmn[1], cmn[2]);
    float3 cinv = make_float3(2097152.0f / std::max(cmx[0] - cmn[0], 1e-30f), 2097152.0f / std::max(cmx[1] - cmn[1], 1e-30f), 2097152.0f / std::max(cmx[2] - cmn[2], 1e-30f));
    std::vector<uint64_t> keys(n); std::vector<int> sorted(n);
    for (int i = 0; i < n; ++i) keys[i] = morton63(&tlo[3 * i], &thi[3 * i], cmin, cinv);
    std::iota(sorted.begin(), sorted.end(), 0);
    std::stable_sort(sorted.begin(), sorted.end(), [&](int a, int b) { return keys[a] < keys[b]; });
    std::vector<uint64_t> skeys(n);
    for (int i = 0; i < n; ++i) skeys[i] = keys[sorted[i]];
    std::vector<int> left(n), right(n), parent(n), first(n), last(n), leaf_parent(n, -1), count(2 * (size_t) n, 0);
    std::vector<float> lo(3 * (2 * (size_t) n)), hi(3 * (2 * (size_t) n));
    std::vector<uint32_t> visit(n, 0);
    std::vector<float> sah(16 * (size_t) n); std::vector<unsigned char> dec(16 * (size_t) n);
    BinTree bt{ n, skeys.data(), sorted.data(), left.data(), right.data(), parent.data(), first.data(), last.data(), leaf_parent.data(), lo.data(), hi.data(), visit.data(), nullptr,
                sah.data(), dec.data(), max_leaf };
    out.sah_cost = 0.0f;
    if (n < 2) {
        for (int a = 0; a < 3; ++a) { lo[a] = tlo[a]; hi[a] = thi[a]; }   // node id (n-1)+0 = 0
    } else if (builder == B2_BUILDER_LBVH) {
        for (int i = 0; i < n - 1; ++i) karras_node(bt, i);
        for (int j = 0; j < n; ++j) refit_from_leaf(bt, tlo.data(), thi.data(), j);
    } else {                                                                // the pass loop of b2_bvh_build.cu, one "thread" at a time
        bt.count = count.data();
        std::vector<int> cid(n), nxt, nn;
        for (int j = 0; j < n; ++j) { ploc_leaf(bt, tlo.data(), thi.data(), j); cid[j] = (n - 1) + j; }
        int created = 0;
        while (cid.size() > 1) {
            const int m = (int) cid.size();
            nn.assign(m, -1);
            for (int i = 0; i < m; ++i) nn[i] = ploc_nearest(bt, cid.data(), m, i, ploc_radius);
            nxt.clear();
            int merged = 0;
            for (int i = 0; i < m; ++i) {
                const int r = ploc_role(nn.data(), i);
                if (r == 2) continue;
                if (r == 1) { const int node = n - 2 - (created + merged); ploc_merge(bt, cid[i], cid[nn[i]], node); nxt.push_back(node); merged++; }
                else nxt.push_back(cid[i]);
            }
            created += merged;
            cid.swap(nxt);
        }
    }
    if (n >= 2) out.sah_cost = sah[0] / std::max(bin_area(bt, 0), 1e-30f);
    std::vector<int> perm(n, -1);
    std::vector<CollapseItem> items(1), next;
    items[0].bin_node = 0;     // root: internal node 0 (or the only leaf when n == 1)
    uint32_t level_start = 0, tris_done = 0;
    while (!items.empty()) {
        const int m = (int) items.size();
        std::vector<CollapseOut> outs(m); std::vector<int> nint(m), ntri(m);
        for (int k = 0; k < m; ++k) collapse_pick(bt, items[k].bin_node, max_leaf, outs[k], nint[k], ntri[k]);
        std::vector<int> ioff(m), toff(m);
        int itot = 0, ttot = 0;
        for (int k = 0; k < m; ++k) { ioff[k] = itot; itot += nint[k]; toff[k] = ttot; ttot += ntri[k]; }
        const uint32_t next_start = level_start + m;
        out.nodes.resize((size_t) (next_start) * B2_NODE_UINT4);
        next.assign(itot, CollapseItem{ -1 });
        for (int k = 0; k < m; ++k)
            collapse_emit(bt, items[k].bin_node, outs[k], max_leaf, next_start + ioff[k], tris_done + toff[k],
                          &out.nodes[(size_t) (level_start + k) * B2_NODE_UINT4], next.data(), ioff[k], perm.data());
        level_start = next_start; tris_done += ttot;
        items.swap(next);
        out.depth++;
    }
    for (int s = 0; s < n; ++s) pack_tri(hs.tris.data(), sorted.data(), perm.data(), s, out.tri.data(), out.tri_obj.data());
}
