// b2_bvh_build.cuh -- per-item bodies of the GPU BVH build (the __global__ wrappers and the level loop
// live in b2_bvh_build.cu; tests/hostsim runs the same bodies sequentially on the host).
//
// Pipeline: padded triangle boxes + centroid bounds -> 63-bit Morton keys -> radix sort -> binary tree ->
// level-synchronous collapse into 8-wide nodes with 8-bit quantised child boxes (b2_scene.cuh) and leaf-contiguous
// triangle reordering.  Two binary-tree builders over the Morton order:
//   PLOC (default)  parallel locally-ordered clustering (Meister & Bittner 2018): every cluster looks `radius` positions
//                   left and right along the Morton curve for the neighbour whose union has the smallest surface area;
//                   mutual nearest neighbours merge; repeat until one cluster is left.  Agglomerative, area-driven: the
//                   tree quality of a surface-area-heuristic build at a few dozen passes over a shrinking array.
//   LBVH            Karras (HPG 2012) binary radix tree + bottom-up refit: splits by Morton bits only (A/B baseline).
// The collapse is SAH-optimal for the given binary tree (Ylitie, Karras, Laine 2017, section 3.1): while the binary tree is
// built bottom-up, every node n gets the table C(n, i), i = 1..7 = the smallest cost of representing its subtree with at most
// i child slots of a wide node (a slot is a leaf of <= max_leaf triangles, cost area * count * B2_SAH_C_PRIM, or a wide node,
// cost area * B2_SAH_C_NODE + the best distribution of its two children over 8 slots), plus the decisions that reach it;
// the level-synchronous collapse then only follows the decisions.  Round 1 opened the largest-area child greedily: on the
// 1 M-triangle scene the bottom wide nodes held 6.4 of 24 possible triangles (tools/bvh_quality.py, profiles/r02_bvh.md).
// Every step is deterministic (no atomics decide an ordering), so all GPUs build identical trees.
#pragma once
#include "b2_scene.cuh"

struct BuildTri { float v[9]; uint32_t order, matflag, group; int32_t obj_id; };   // upload format, 52 B

struct BinTree {             // binary radix tree over n sorted triangles: internal nodes [0,n-2], leaf j = (n-1)+j
    int n;
    const uint64_t *keys;    // sorted Morton keys
    const int *sorted;       // sorted position -> input triangle index
    int *left, *right, *parent, *first, *last;   // internal nodes
    int *leaf_parent;
    float *lo, *hi;          // 3 floats per node, 2n-1 nodes
    uint32_t *visit;         // refit arrival counters (internal nodes)
    int *count;              // PLOC: leaves under each node (2n-1 entries; a subtree is NOT a contiguous range of sorted
                             // leaves there); nullptr for the radix tree, whose nodes cover [first, last]
    float *sah;              // 8 per node: [i] = C(node, i) for i = 1..7, [0] = cost of the node as a wide node (C_internal)
    unsigned char *dec;      // 8 per node: [1] = 1 if C(node,1) is the leaf, [i] (i = 2..7) = slots given to the left child, 0 = "i-1
                             // slots are as good"; [0] = slots of the left child when the node becomes a wide node
    int max_leaf;            // most triangles in a leaf slot (1..3)
};
#ifndef B2_SAH_C_NODE
#define B2_SAH_C_NODE 1.0f   // one wide-node test (80 B, ~260 instructions)
#endif
#ifndef B2_SAH_C_PRIM
#define B2_SAH_C_PRIM 0.3f   // one triangle test (48 B, 40 - 75 instructions)
#endif

B2_HD void tri_bounds(const BuildTri &t, float pad, float lo[3], float hi[3]) {
    for (int a = 0; a < 3; ++a) {
        float x0 = t.v[a], x1 = t.v[3 + a], x2 = t.v[6 + a];
        lo[a] = fminf(x0, fminf(x1, x2)) - pad;
        hi[a] = fmaxf(x0, fmaxf(x1, x2)) + pad;
    }
}
B2_HD uint64_t expand21(uint32_t v) {   // spread the low 21 bits to every third bit
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}
B2_HD uint64_t morton63(const float lo[3], const float hi[3], float3 cmin, float3 cinv) {
    float cx = 0.5f * (lo[0] + hi[0]), cy = 0.5f * (lo[1] + hi[1]), cz = 0.5f * (lo[2] + hi[2]);
    float fx = fminf(fmaxf((cx - cmin.x) * cinv.x, 0.0f), 2097151.0f);
    float fy = fminf(fmaxf((cy - cmin.y) * cinv.y, 0.0f), 2097151.0f);
    float fz = fminf(fmaxf((cz - cmin.z) * cinv.z, 0.0f), 2097151.0f);
    return (expand21((uint32_t) fx) << 2) | (expand21((uint32_t) fy) << 1) | expand21((uint32_t) fz);
}

B2_HD float bin_area(const BinTree &bt, int node) {
    float dx = bt.hi[3 * node] - bt.lo[3 * node], dy = bt.hi[3 * node + 1] - bt.lo[3 * node + 1], dz = bt.hi[3 * node + 2] - bt.lo[3 * node + 2];
    return dx * dy + dy * dz + dz * dx;
}
// ---- SAH-optimal collapse: cost tables (bottom-up, fused into the binary-tree construction) ----
B2_HD float sah_ld(const float *p) {
#if defined(__CUDA_ARCH__)
    return __ldcg(p);
#else
    return *p;
#endif
}
B2_HD void sah_leaf(const BinTree &bt, int node) {            // a single triangle: one leaf slot whatever the budget
    if (!bt.sah) return;
    const float c = bin_area(bt, node) * B2_SAH_C_PRIM;
    for (int i = 0; i < 8; ++i) { bt.sah[8 * (size_t) node + i] = c; bt.dec[8 * (size_t) node + i] = 0; }
    bt.dec[8 * (size_t) node + 1] = 1;
}
B2_HD void sah_combine(const BinTree &bt, int node, int l, int r, int count) {
    if (!bt.sah) return;
    // the children's tables may have been written by other threads of this launch (radix-tree refit): L2 loads, like the boxes
    float lc[8], rc[8];
    for (int i = 1; i < 8; ++i) { lc[i] = sah_ld(bt.sah + 8 * (size_t) l + i); rc[i] = sah_ld(bt.sah + 8 * (size_t) r + i); }
    float *c = bt.sah + 8 * (size_t) node; unsigned char *d = bt.dec + 8 * (size_t) node;
    const float area = bin_area(bt, node);
    float dist[9]; unsigned char dk[9];
    for (int j = 2; j <= 8; ++j) {                               // best distribution of j slots over the two children
        float best = 3.0e38f; int bk = 1;
        for (int k = (j - 7 > 1 ? j - 7 : 1); k <= j - 1 && k <= 7; ++k) { const float v = lc[k] + rc[j - k]; if (v < best) { best = v; bk = k; } }
        dist[j] = best; dk[j] = (unsigned char) bk;
    }
    const float c_int = dist[8] + area * B2_SAH_C_NODE;
    const float c_leaf = count <= bt.max_leaf ? area * (float) count * B2_SAH_C_PRIM : 3.0e38f;
    c[0] = c_int; d[0] = dk[8];
    if (c_leaf <= c_int) { c[1] = c_leaf; d[1] = 1; } else { c[1] = c_int; d[1] = 0; }
    for (int i = 2; i <= 7; ++i) {
        if (dist[i] < c[i - 1]) { c[i] = dist[i]; d[i] = dk[i]; } else { c[i] = c[i - 1]; d[i] = 0; }
    }
}

B2_HD int karras_delta(const uint64_t *keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + clz32((uint32_t) i ^ (uint32_t) j);
    return clz64(a ^ b);
}
B2_HD void karras_node(const BinTree &bt, int i) {
    const int n = bt.n;
    const uint64_t *k = bt.keys;
    int d = (karras_delta(k, n, i, i + 1) - karras_delta(k, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = karras_delta(k, n, i, i - d);
    int lmax = 2;
    while (karras_delta(k, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (karras_delta(k, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = karras_delta(k, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (karras_delta(k, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + (d < 0 ? d : 0);
    int lo = i < j ? i : j, hi = i < j ? j : i;
    int L = (lo == gamma) ? (n - 1) + gamma : gamma;
    int R = (hi == gamma + 1) ? (n - 1) + gamma + 1 : gamma + 1;
    bt.left[i] = L; bt.right[i] = R; bt.first[i] = lo; bt.last[i] = hi;
    if (L >= n - 1) bt.leaf_parent[L - (n - 1)] = i; else bt.parent[L] = i;
    if (R >= n - 1) bt.leaf_parent[R - (n - 1)] = i; else bt.parent[R] = i;
    if (i == 0) bt.parent[0] = -1;
}
// Leaf j writes its box, then climbs: the second arrival at each internal node merges the children.
B2_HD void refit_from_leaf(const BinTree &bt, const float *tri_lo, const float *tri_hi, int j) {
    const int n = bt.n;
    int src = bt.sorted[j];
    int node = (n - 1) + j;
    for (int a = 0; a < 3; ++a) { bt.lo[3 * node + a] = tri_lo[3 * src + a]; bt.hi[3 * node + a] = tri_hi[3 * src + a]; }
    sah_leaf(bt, node);
    int p = bt.leaf_parent[j];
    while (p >= 0) {
#if defined(__CUDA_ARCH__)
        __threadfence();
        if (atomicAdd(&bt.visit[p], 1u) == 0u) return;
        __threadfence();
#else
        if (bt.visit[p]++ == 0u) return;
#endif
        int L = bt.left[p], R = bt.right[p];
        for (int a = 0; a < 3; ++a) {
#if defined(__CUDA_ARCH__)
            float l0 = __ldcg(&bt.lo[3 * L + a]), l1 = __ldcg(&bt.lo[3 * R + a]), h0 = __ldcg(&bt.hi[3 * L + a]), h1 = __ldcg(&bt.hi[3 * R + a]);
#else
            float l0 = bt.lo[3 * L + a], l1 = bt.lo[3 * R + a], h0 = bt.hi[3 * L + a], h1 = bt.hi[3 * R + a];
#endif
            bt.lo[3 * p + a] = fminf(l0, l1);
            bt.hi[3 * p + a] = fmaxf(h0, h1);
        }
        sah_combine(bt, p, L, R, bt.last[p] - bt.first[p] + 1);
        p = bt.parent[p];
    }
}

B2_HD int bin_count(const BinTree &bt, int node) { return node >= bt.n - 1 ? 1 : (bt.count ? bt.count[node] : bt.last[node] - bt.first[node] + 1); }
// sorted positions of the (at most 3) leaves under a node that becomes a leaf child of a wide node
B2_HD int bin_leaves(const BinTree &bt, int node, int out[3]) {
    if (!bt.count) {
        const int f = node >= bt.n - 1 ? node - (bt.n - 1) : bt.first[node], c = bin_count(bt, node);
        for (int j = 0; j < c && j < 3; ++j) out[j] = f + j;
        return c;
    }
    int stack[4], sp = 0, k = 0;
    stack[sp++] = node;
    while (sp > 0) {
        const int x = stack[--sp];
        if (x >= bt.n - 1) { if (k < 3) out[k] = x - (bt.n - 1); k++; }
        else { stack[sp++] = bt.right[x]; stack[sp++] = bt.left[x]; }
    }
    return k;
}

// ---- PLOC ----
// Clusters live in an array `cid` (node ids) in Morton order.  Search key = (surface area of the union box, i ^ j): the
// second component is symmetric in (i, j) and unique among the candidates of one cluster, so the pair that minimises the
// key globally is always mutual (the loop makes progress) and equal boxes pair up as (0,1), (2,3), ... instead of chaining.
B2_HD float ploc_union_area(const BinTree &bt, int a, int b) {
    const float dx = fmaxf(bt.hi[3 * a], bt.hi[3 * b]) - fminf(bt.lo[3 * a], bt.lo[3 * b]);
    const float dy = fmaxf(bt.hi[3 * a + 1], bt.hi[3 * b + 1]) - fminf(bt.lo[3 * a + 1], bt.lo[3 * b + 1]);
    const float dz = fmaxf(bt.hi[3 * a + 2], bt.hi[3 * b + 2]) - fminf(bt.lo[3 * a + 2], bt.lo[3 * b + 2]);
    return dx * dy + dy * dz + dz * dx;
}
B2_HD int ploc_nearest(const BinTree &bt, const int *cid, int m, int i, int radius) {
    const int a = cid[i];
    const float alx = bt.lo[3 * a], aly = bt.lo[3 * a + 1], alz = bt.lo[3 * a + 2], ahx = bt.hi[3 * a], ahy = bt.hi[3 * a + 1], ahz = bt.hi[3 * a + 2];
    int best = -1; float best_area = 3.0e38f; uint32_t best_x = 0xffffffffu;
    const int j0 = i - radius > 0 ? i - radius : 0, j1 = i + radius < m - 1 ? i + radius : m - 1;
    for (int j = j0; j <= j1; ++j) {
        if (j == i) continue;
        const int b = cid[j];
        const float dx = fmaxf(ahx, bt.hi[3 * b]) - fminf(alx, bt.lo[3 * b]);
        const float dy = fmaxf(ahy, bt.hi[3 * b + 1]) - fminf(aly, bt.lo[3 * b + 1]);
        const float dz = fmaxf(ahz, bt.hi[3 * b + 2]) - fminf(alz, bt.lo[3 * b + 2]);
        const float area = dx * dy + dy * dz + dz * dx;
        const uint32_t x = (uint32_t) i ^ (uint32_t) j;
        if (area < best_area || (area == best_area && x < best_x)) { best_area = area; best_x = x; best = j; }
    }
    return best;
}
// 1 = cluster i is the left partner of a merging pair, 2 = the right partner (it disappears), 0 = it stays
B2_HD int ploc_role(const int *nn, int i) { const int j = nn[i]; return (j >= 0 && nn[j] == i) ? (i < j ? 1 : 2) : 0; }
B2_HD void ploc_merge(const BinTree &bt, int l, int r, int node) {
    bt.left[node] = l; bt.right[node] = r;
    for (int a = 0; a < 3; ++a) { bt.lo[3 * node + a] = fminf(bt.lo[3 * l + a], bt.lo[3 * r + a]); bt.hi[3 * node + a] = fmaxf(bt.hi[3 * l + a], bt.hi[3 * r + a]); }
    bt.count[node] = bt.count[l] + bt.count[r];
    sah_combine(bt, node, l, r, bt.count[node]);
}
B2_HD void ploc_leaf(const BinTree &bt, const float *tri_lo, const float *tri_hi, int j) {
    const int src = bt.sorted[j], node = (bt.n - 1) + j;
    for (int a = 0; a < 3; ++a) { bt.lo[3 * node + a] = tri_lo[3 * src + a]; bt.hi[3 * node + a] = tri_hi[3 * src + a]; }
    bt.count[node] = 1;
    sah_leaf(bt, node);
}

B2_HD uint32_t pack4(const uint32_t *b) { return b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24); }

struct CollapseItem { int bin_node; };          // one per wide node of the current level
struct CollapseOut { int kids[8]; int n_kids; uint32_t leaf_mask; };      // leaf_mask bit k: kids[k] is a leaf slot

// Pass 1: the (up to 8) children of the wide node rooted at binary node X, read off the decisions of the cost tables:
// X's best split of 8 slots over its two children, then every child either keeps its budget, gives a slot back (a smaller
// budget is as good) or splits it again, until every part has one slot -- a leaf or the root of a wide node of the next level.
B2_HD void collapse_pick(const BinTree &bt, int X, int max_leaf, CollapseOut &out, int &n_internal, int &n_tris) {
    int kids[8], nk = 0; uint32_t leaf_mask = 0u;
    if (X >= bt.n - 1 || !bt.sah) {
        if (X >= bt.n - 1 || bin_count(bt, X) <= max_leaf) { kids[nk] = X; leaf_mask |= 1u << nk; nk++; }      // the whole tree is one leaf
        else {                                                  // no cost tables (not used by the builders): greedy by area
            kids[nk++] = bt.left[X]; kids[nk++] = bt.right[X];
            while (nk < 8) {
                int best = -1; float best_a = -1.0f;
                for (int k = 0; k < nk; ++k) { if (bin_count(bt, kids[k]) <= max_leaf) continue; const float a = bin_area(bt, kids[k]); if (a > best_a) { best_a = a; best = k; } }
                if (best < 0) break;
                const int nd = kids[best]; kids[best] = bt.left[nd]; kids[nk++] = bt.right[nd];
            }
            for (int k = 0; k < nk; ++k) if (bin_count(bt, kids[k]) <= max_leaf) leaf_mask |= 1u << k;
        }
    } else {
        int st_node[8], st_budget[8], sp = 0;
        const int k0 = bt.dec[8 * (size_t) X];
        st_node[sp] = bt.right[X]; st_budget[sp++] = 8 - k0;
        st_node[sp] = bt.left[X]; st_budget[sp++] = k0;
        while (sp > 0) {
            const int nd = st_node[--sp]; int i = st_budget[sp];
            if (nd >= bt.n - 1) { kids[nk] = nd; leaf_mask |= 1u << nk; nk++; continue; }
            const unsigned char *d = bt.dec + 8 * (size_t) nd;
            while (i >= 2 && d[i] == 0) --i;
            if (i <= 1) { kids[nk] = nd; if (d[1]) leaf_mask |= 1u << nk; nk++; continue; }
            st_node[sp] = bt.right[nd]; st_budget[sp++] = i - d[i];
            st_node[sp] = bt.left[nd]; st_budget[sp++] = d[i];
        }
    }
    n_internal = 0; n_tris = 0;
    for (int k = 0; k < 8; ++k) {
        out.kids[k] = k < nk ? kids[k] : -1;
        if (k < nk) { if ((leaf_mask >> k) & 1u) n_tris += bin_count(bt, kids[k]); else n_internal++; }
    }
    out.n_kids = nk; out.leaf_mask = leaf_mask;
}

// Pass 2: slot assignment, quantisation, node write-out, next-level work items and the leaf-order permutation.
// child_base / tri_base come from exclusive scans over the level.  perm[tri slot] = sorted position.
B2_HD void collapse_emit(const BinTree &bt, int X, const CollapseOut &in, int max_leaf, uint32_t child_base, uint32_t tri_base,
                         uint4 *node_out, CollapseItem *next_items, int next_local_base, int *perm) {
    const int nk = in.n_kids;
    // node box = union of the kids (== box of X except for the single-leaf root)
    float nlo[3] = { 1e30f, 1e30f, 1e30f }, nhi[3] = { -1e30f, -1e30f, -1e30f };
    for (int k = 0; k < nk; ++k)
        for (int a = 0; a < 3; ++a) { nlo[a] = fminf(nlo[a], bt.lo[3 * in.kids[k] + a]); nhi[a] = fmaxf(nhi[a], bt.hi[3 * in.kids[k] + a]); }
    // greedy slot assignment: repeatedly take the (kid, slot) pair with the largest projection of the kid's
    // centroid offset on the slot's corner direction
    int slot_of[8]; bool slot_used[8], kid_done[8];
    for (int k = 0; k < 8; ++k) { slot_used[k] = false; kid_done[k] = false; slot_of[k] = -1; }
    float ncx = 0.5f * (nlo[0] + nhi[0]), ncy = 0.5f * (nlo[1] + nhi[1]), ncz = 0.5f * (nlo[2] + nhi[2]);
    for (int it = 0; it < nk; ++it) {
        int bk = -1, bs = -1; float bc = -3.0e38f;
        for (int k = 0; k < nk; ++k) {
            if (kid_done[k]) continue;
            int nd = in.kids[k];
            float dx = 0.5f * (bt.lo[3 * nd] + bt.hi[3 * nd]) - ncx, dy = 0.5f * (bt.lo[3 * nd + 1] + bt.hi[3 * nd + 1]) - ncy, dz = 0.5f * (bt.lo[3 * nd + 2] + bt.hi[3 * nd + 2]) - ncz;
            for (int s = 0; s < 8; ++s) {
                if (slot_used[s]) continue;
                float c = ((s & 1) ? dx : -dx) + ((s & 2) ? dy : -dy) + ((s & 4) ? dz : -dz);
                if (c > bc) { bc = c; bk = k; bs = s; }
            }
        }
        kid_done[bk] = true; slot_used[bs] = true; slot_of[bk] = bs;
    }
    int kid_at[8];
    for (int s = 0; s < 8; ++s) kid_at[s] = -1;
    for (int k = 0; k < nk; ++k) kid_at[slot_of[k]] = k;
    // quantisation grid: scale = 2^e >= extent / 255 per axis
    uint32_t ebits[3]; float scale[3];
    for (int a = 0; a < 3; ++a) {
        float ext = nhi[a] - nlo[a];
        uint32_t e = 1u;
        if (ext > 0.0f) {
            uint32_t b = f2u(ext / 255.0f);
            e = (b >> 23) & 0xffu;
            if (b & 0x7fffffu) e += 1u;          // round the scale up to a power of two
            if (e < 1u) e = 1u;
            if (e > 254u) e = 254u;
        }
        // guard against rounding in (hi-lo)/scale: grow until the far plane fits in 8 bits
        for (;;) {
            float sc = u2f(e << 23);
            if (ceilf((nhi[a] - nlo[a]) / sc) <= 255.0f || e >= 254u) break;
            e++;
        }
        ebits[a] = e; scale[a] = u2f(e << 23);
    }
    uint32_t imask = 0, vmask = 0, q[6][8];
    int n_int = 0, tri_off = 0;
    for (int s = 0; s < 8; ++s) {
        int k = kid_at[s];
        if (k < 0) {                                 // empty slot: an inverted box (lo 255, hi 0) that no ray can hit
            for (int a = 0; a < 3; ++a) { q[a][s] = 255u; q[3 + a][s] = 0u; }
            continue;
        }
        int nd = in.kids[k];
        for (int a = 0; a < 3; ++a) {
            float ql = floorf((bt.lo[3 * nd + a] - nlo[a]) / scale[a]);
            float qh = ceilf((bt.hi[3 * nd + a] - nlo[a]) / scale[a]);
            // make sure the dequantised planes really enclose the child (fp rounding of the divisions)
            while (ql > 0.0f && nlo[a] + ql * scale[a] > bt.lo[3 * nd + a]) ql -= 1.0f;
            while (qh < 255.0f && nlo[a] + qh * scale[a] < bt.hi[3 * nd + a]) qh += 1.0f;
            q[a][s] = (uint32_t) fminf(fmaxf(ql, 0.0f), 255.0f);
            q[3 + a][s] = (uint32_t) fminf(fmaxf(qh, 0.0f), 255.0f);
        }
        int c = bin_count(bt, nd);
        if (!((in.leaf_mask >> k) & 1u)) {
            imask |= 1u << s;
            n_int++;
        } else {
            vmask |= ((1u << c) - 1u) << (3 * s);    // leaf slot s owns bits 3s .. 3s+2; its triangles sit compactly in slot order
            int lv[3];
            bin_leaves(bt, nd, lv);
            for (int j = 0; j < c; ++j) perm[tri_base + tri_off + j] = lv[j];
            tri_off += c;
        }
    }
    // internal children are numbered in slot order: child_base + popc(imask below slot)
    for (int s = 0; s < 8; ++s)
        if (imask & (1u << s)) next_items[next_local_base + popc32(imask & ((1u << s) - 1u))].bin_node = in.kids[kid_at[s]];
    node_out[0] = make_uint4(f2u(nlo[0]), f2u(nlo[1]), f2u(nlo[2]), ebits[0] | (ebits[1] << 8) | (ebits[2] << 16) | (imask << 24));
    node_out[1] = make_uint4(child_base, tri_base, vmask, 0u);
    node_out[2] = make_uint4(pack4(q[0]), pack4(q[0] + 4), pack4(q[1]), pack4(q[1] + 4));
    node_out[3] = make_uint4(pack4(q[2]), pack4(q[2] + 4), pack4(q[3]), pack4(q[3] + 4));
    node_out[4] = make_uint4(pack4(q[4]), pack4(q[4] + 4), pack4(q[5]), pack4(q[5] + 4));
}

// Final gather into leaf order: tri slot -> packed 3 x float4 + obj id.
B2_HD void pack_tri(const BuildTri *in, const int *sorted, const int *perm, int slot, float4 *tri_out, int32_t *obj_out) {
    const BuildTri &t = in[sorted[perm[slot]]];
    float3 v0 = mk3(t.v[0], t.v[1], t.v[2]);
    float3 e1 = x3_sub(mk3(t.v[3], t.v[4], t.v[5]), v0);     // object.cpp:74: e1 = v1 - v0, e2 = v2 - v0
    float3 e2 = x3_sub(mk3(t.v[6], t.v[7], t.v[8]), v0);
    tri_out[3 * slot] = mk4(v0, u2f(t.order));
    tri_out[3 * slot + 1] = mk4(e1, u2f(t.matflag));
    tri_out[3 * slot + 2] = mk4(e2, u2f(t.group));
    obj_out[slot] = t.obj_id;
}
