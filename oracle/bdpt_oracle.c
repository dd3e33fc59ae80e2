/* oracle/bdpt_oracle.c -- TEST INFRASTRUCTURE ONLY (see bdpt_oracle.h).
 *
 * Plain-C restatement of the reference CPU BDPT path.  Every function cites the reference
 * file:line it follows (paths relative to /root/reference).  Operation order is kept
 * exactly (fp32, no contraction) so that results are bit-identical to the reference built
 * with g++ -O2 on baseline x86-64; tests/test_oracle_pin.py proves that against oracle/_ref.
 * Parity: pinned against the compiled reference (no golden vectors exist upstream).
 */
#define _POSIX_C_SOURCE 200809L
#include "bdpt_oracle.h"
#include <math.h>
#include <omp.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define OR_EPSILON 1e-4f                      /* geometric.cuh:6 */
#define OR_PI 3.14159265358979323846f         /* geometric.cuh:7 / object.h:8 */

typedef struct { float x, y, z; } v3;

/* ---- float3 operators, geometric.cuh:90-107 (division-based normalize) ---- */
static inline v3 mk3(float x, float y, float z) { v3 r = { x, y, z }; return r; }
static inline v3 f3_add(v3 a, v3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 f3_sub(v3 a, v3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 f3_muls(v3 a, float b) { return mk3(a.x * b, a.y * b, a.z * b); }
static inline v3 f3_mul(v3 a, v3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 f3_divs(v3 a, float b) { return mk3(a.x / b, a.y / b, a.z / b); }
static inline float f3_dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 f3_cross(v3 a, v3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
static inline float f3_length(v3 a) { return sqrtf(f3_dot(a, a)); }
static inline v3 f3_normalize(v3 a) { return f3_divs(a, f3_length(a)); }
static inline v3 f3_reflect(v3 I, v3 N) { return f3_sub(I, f3_muls(f3_muls(N, 2.0f), f3_dot(N, I))); } /* :99 */

/* ---- glm scalar semantics (oracle/compat/glm/glm.hpp; upstream func_geometric.inl) ---- */
static inline float glm_dot(v3 a, v3 b) { v3 t = f3_mul(a, b); return t.x + t.y + t.z; }
static inline v3 glm_cross(v3 x, v3 y) { return mk3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
static inline v3 glm_normalize(v3 v) { return f3_muls(v, 1.0f / sqrtf(glm_dot(v, v))); }
static inline v3 ld3(const float *p) { return mk3(p[0], p[1], p[2]); }
static inline void st3(float *p, v3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }

/* =====================================================================================
 * RNG
 * ===================================================================================== */
/* std::mt19937 (MT19937, Matsumoto & Nishimura) as libstdc++ instantiates it. */
typedef struct { uint32_t mt[624]; int idx; } Mt19937;
static void mt_seed(Mt19937 *m, uint32_t seed) {
    m->mt[0] = seed;
    for (int i = 1; i < 624; ++i) m->mt[i] = 1812433253u * (m->mt[i - 1] ^ (m->mt[i - 1] >> 30)) + (uint32_t) i;
    m->idx = 624;
}
static uint32_t mt_next(Mt19937 *m) {
    if (m->idx >= 624) {
        for (int i = 0; i < 624; ++i) {
            uint32_t y = (m->mt[i] & 0x80000000u) | (m->mt[(i + 1) % 624] & 0x7fffffffu);
            m->mt[i] = m->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        m->idx = 0;
    }
    uint32_t y = m->mt[m->idx++];
    y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
    return y;
}
/* libstdc++ std::uniform_real_distribution<float>(0,1)(mt19937): generate_canonical<float,24> draws ONE
 * 32-bit word: float(word) / 2^32 (rounded to nearest), and 1.0f is replaced by nextafterf(1,0). */
static float mt_unit_float(Mt19937 *m) {
    float sum = (float) mt_next(m) * 1.0f;
    float ret = sum / 4294967296.0f;
    if (ret >= 1.0f) ret = nextafterf(1.0f, 0.0f);
    return ret * (1.0f - 0.0f) + 0.0f;
}
void or_mt19937_floats(uint32_t seed, int n, float *out) {
    Mt19937 m; mt_seed(&m, seed);
    for (int i = 0; i < n; ++i) out[i] = mt_unit_float(&m);
}

/* Philox4x32-10 (Salmon et al., SC'11; Random123 constants). */
void or_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t) 0xD2511F53u * c0, p1 = (uint64_t) 0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t) (p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t) p1;
        uint32_t n2 = (uint32_t) (p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t) p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
float or_u32_to_unit_float(uint32_t x) { return (float) (x >> 8) * 5.9604644775390625e-8f; /* 2^-24 */ }

/* Draw schedule (shared with the GPU, path_tracing_b200/csrc/b2_rng.cuh):
 *   ctr = (id0, id1, block, kind), key = (seed_lo, seed_hi)
 *   kind 1 = light subpath (id0 = light-path index, id1 = 0); kind 2 = eye subpath (id0 = pixel index, id1 = sample index)
 *   block 0 lanes 0,1     : emission (r1,r2 | u1,u2; cpu_bdpt.cpp:236,248) or pixel jitter (:350-351)
 *   block 1+k lanes 0,1,2 : u_rr,u1,u2 of the k-th bsdf_sample call of the subpath, delta bounces included (:297,:448) */
typedef struct {
    int mode;
    Mt19937 *mt;
    uint32_t key[2], id0, id1, kind, bounce;
} Rng;
static void rng_begin(Rng *r, uint32_t kind, uint32_t id0, uint32_t id1) { r->kind = kind; r->id0 = id0; r->id1 = id1; r->bounce = 0; }
static void rng_draw2(Rng *r, float *a, float *b) {
    if (r->mode == OR_RNG_MT19937) { *a = mt_unit_float(r->mt); *b = mt_unit_float(r->mt); return; }
    uint32_t c[4] = { r->id0, r->id1, 0u, r->kind }, o[4];
    or_philox4x32_10(c, r->key, o);
    *a = or_u32_to_unit_float(o[0]); *b = or_u32_to_unit_float(o[1]);
}
static void rng_draw3(Rng *r, float *a, float *b, float *c3) {
    if (r->mode == OR_RNG_MT19937) { *a = mt_unit_float(r->mt); *b = mt_unit_float(r->mt); *c3 = mt_unit_float(r->mt); return; }
    uint32_t c[4] = { r->id0, r->id1, 1u + r->bounce, r->kind }, o[4];
    r->bounce++;
    or_philox4x32_10(c, r->key, o);
    *a = or_u32_to_unit_float(o[0]); *b = or_u32_to_unit_float(o[1]); *c3 = or_u32_to_unit_float(o[2]);
}

/* =====================================================================================
 * Scene (object.h:28-110, object.cpp:123-147, main_cli.cpp:99-141)
 * ===================================================================================== */
typedef struct { v3 base_color; float roughness, metallic, eta; } Mtl; /* CudaMaterial, geometric.cuh:21-27 (type unused) */
typedef struct { int kind, obj_id; Mtl mtl; v3 v0, v1, v2; float r; } Obj; /* sphere: v0 = centre, scale = (1,1,1) */
typedef struct { int key; v3 mn, mx; int n, cap; int *objs; } Group;    /* AABB, object.h:94-102 */
typedef struct { v3 pos, dir, illum; float cutoff; int is_parallel; float ball_r; } Light; /* CudaLight, geometric.cuh:73-78 */
struct OrScene { int n_obj; Obj *objs; int n_groups; Group *groups; int n_lights; Light *lights; };

static float min3f(float a, float b, float c) { float m = a; if (b < m) m = b; if (c < m) m = c; return m; } /* std::min({..}) */
static float max3f(float a, float b, float c) { float m = a; if (m < b) m = b; if (m < c) m = c; return m; } /* std::max({..}) */

/* AABB::add_obj, object.cpp:123-147 */
static void group_add(Group *g, const Obj *o, int index) {
    if (o->kind == 0) {
        v3 c = o->v0; float r = o->r;
        g->mn.x = min3f(g->mn.x, c.x + r, c.x - r); g->mn.y = min3f(g->mn.y, c.y + r, c.y - r); g->mn.z = min3f(g->mn.z, c.z + r, c.z - r);
        g->mx.x = max3f(g->mx.x, c.x + r, c.x - r); g->mx.y = max3f(g->mx.y, c.y + r, c.y - r); g->mx.z = max3f(g->mx.z, c.z + r, c.z - r);
    } else {
        const v3 *vs[3] = { &o->v0, &o->v1, &o->v2 };
        for (int i = 0; i < 3; ++i) {
            if (vs[i]->x < g->mn.x) g->mn.x = vs[i]->x;
            if (vs[i]->y < g->mn.y) g->mn.y = vs[i]->y;
            if (vs[i]->z < g->mn.z) g->mn.z = vs[i]->z;
            if (g->mx.x < vs[i]->x) g->mx.x = vs[i]->x;
            if (g->mx.y < vs[i]->y) g->mx.y = vs[i]->y;
            if (g->mx.z < vs[i]->z) g->mx.z = vs[i]->z;
        }
    }
    if (g->n == g->cap) { g->cap = g->cap ? 2 * g->cap : 16; g->objs = (int *) realloc(g->objs, sizeof(int) * g->cap); }
    g->objs[g->n++] = index;
}

OrScene *or_scene_create(int n_obj, const int *kind, const int *group, const float *geom, const float *mtl,
                         int n_lights, const float *lights) {
    OrScene *s = (OrScene *) calloc(1, sizeof(OrScene));
    s->n_obj = n_obj;
    s->objs = (Obj *) calloc(n_obj > 0 ? n_obj : 1, sizeof(Obj));
    for (int i = 0; i < n_obj; ++i) {
        Obj *o = &s->objs[i];
        const float *g = geom + 9 * i, *m = mtl + 6 * i;
        o->kind = kind[i]; o->obj_id = i;
        o->mtl.base_color = ld3(m); o->mtl.roughness = m[3]; o->mtl.metallic = m[4]; o->mtl.eta = m[5];
        o->v0 = ld3(g);
        if (kind[i] == 0) o->r = g[3]; else { o->v1 = ld3(g + 3); o->v2 = ld3(g + 6); }
        /* std::map<int,AABB>: find-or-insert keeping key order */
        int gi = 0;
        while (gi < s->n_groups && s->groups[gi].key < group[i]) ++gi;
        if (gi == s->n_groups || s->groups[gi].key != group[i]) {
            s->groups = (Group *) realloc(s->groups, sizeof(Group) * (s->n_groups + 1));
            memmove(&s->groups[gi + 1], &s->groups[gi], sizeof(Group) * (s->n_groups - gi));
            Group *ng = &s->groups[gi];
            memset(ng, 0, sizeof(Group));
            ng->key = group[i];
            ng->mn = mk3(99999.f, 99999.f, 99999.f); ng->mx = mk3(-99999.f, -99999.f, -99999.f); /* object.h:96-97 */
            s->n_groups++;
        }
        group_add(&s->groups[gi], o, i);
    }
    s->n_lights = n_lights;
    s->lights = (Light *) calloc(n_lights > 0 ? n_lights : 1, sizeof(Light));
    for (int i = 0; i < n_lights; ++i) {
        const float *l = lights + 12 * i;
        Light *L = &s->lights[i];
        L->pos = ld3(l); L->dir = ld3(l + 3); L->illum = ld3(l + 6);
        L->cutoff = l[9]; L->is_parallel = (int) l[10]; L->ball_r = l[11];
    }
    return s;
}
void or_scene_destroy(OrScene *s) {
    if (!s) return;
    for (int i = 0; i < s->n_groups; ++i) free(s->groups[i].objs);
    free(s->groups); free(s->objs); free(s->lights); free(s);
}
int or_group_bounds(OrScene *s, float *out, int cap) {
    for (int i = 0; i < s->n_groups && i < cap; ++i) {
        float *o = out + 7 * i;
        o[0] = (float) s->groups[i].key; st3(o + 1, s->groups[i].mn); st3(o + 4, s->groups[i].mx);
    }
    return s->n_groups;
}

/* =====================================================================================
 * Scene grammar parser, main_cli.cpp:99-141.  `input >> char` skips whitespace and takes one
 * character; characters matching no rule are dropped one at a time; a failed numeric
 * extraction puts the stream in the fail state, which ends the `while(input >> t)` loop.
 * ===================================================================================== */
typedef struct { const char *p, *end; int fail; } Stream;
static void skip_ws(Stream *s) { while (s->p < s->end && (*s->p == ' ' || *s->p == '\t' || *s->p == '\n' || *s->p == '\r' || *s->p == '\v' || *s->p == '\f')) s->p++; }
static int get_char(Stream *s, char *c) { if (s->fail) return 0; skip_ws(s); if (s->p >= s->end) { s->fail = 1; return 0; } *c = *s->p++; return 1; }
static float get_float(Stream *s) {
    if (s->fail) return 0.0f;
    skip_ws(s);
    char *e; float v = strtof(s->p, &e);
    if (e == s->p || e > s->end) { s->fail = 1; return 0.0f; }
    s->p = e; return v;
}
static int get_int(Stream *s) {
    if (s->fail) return 0;
    skip_ws(s);
    char *e; long v = strtol(s->p, &e, 10);
    if (e == s->p) { s->fail = 1; return 0; }
    s->p = e; return (int) v;
}
OrParsed *or_parse_scene_file(const char *path) {
    FILE *f = fopen(path, "rb");
    if (!f) return NULL;
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    char *buf = (char *) malloc(sz + 1);
    if (fread(buf, 1, sz, f) != (size_t) sz) { fclose(f); free(buf); return NULL; }
    fclose(f); buf[sz] = 0;
    Stream st = { buf, buf + sz, 0 };
    OrParsed *P = (OrParsed *) calloc(1, sizeof(OrParsed));
    int cap_o = 0, cap_l = 0;
    float mtl[6] = { 0, 0, 0, 0, 0, 0 };
    int group_id = 0;
    char t;
    while (get_char(&st, &t)) {
        if (t == 'E') { for (int i = 0; i < 3; ++i) P->cam[i] = get_float(&st); }
        else if (t == 'V') { for (int i = 3; i < 9; ++i) P->cam[i] = get_float(&st); }
        else if (t == 'F') { P->cam[9] = get_float(&st); }
        else if (t == 'R') { P->W = get_int(&st); P->H = get_int(&st); }
        else if (t == 'S' || t == 'T') {
            if (P->n_obj == cap_o) {
                cap_o = cap_o ? 2 * cap_o : 64;
                P->kind = (int *) realloc(P->kind, sizeof(int) * cap_o); P->group = (int *) realloc(P->group, sizeof(int) * cap_o);
                P->geom = (float *) realloc(P->geom, sizeof(float) * 9 * cap_o); P->mtl = (float *) realloc(P->mtl, sizeof(float) * 6 * cap_o);
            }
            float *g = P->geom + 9 * P->n_obj;
            memset(g, 0, 9 * sizeof(float));
            int nf = (t == 'S') ? 4 : 9;
            for (int i = 0; i < nf; ++i) g[i] = get_float(&st);
            P->kind[P->n_obj] = (t == 'S') ? 0 : 1;
            P->group[P->n_obj] = group_id;
            memcpy(P->mtl + 6 * P->n_obj, mtl, sizeof(mtl));
            P->n_obj++;
        }
        else if (t == 'M') { for (int i = 0; i < 6; ++i) mtl[i] = get_float(&st); }
        else if (t == 'G') { group_id = get_int(&st); }
        else if (t == '/') {
            if (get_char(&st, &t) && t == '/') { while (st.p < st.end && *st.p != '\n') st.p++; if (st.p < st.end) st.p++; continue; }
        }
        else if (t == 'L') {
            if (P->n_lights == cap_l) { cap_l = cap_l ? 2 * cap_l : 8; P->lights = (float *) realloc(P->lights, sizeof(float) * 12 * cap_l); }
            float *l = P->lights + 12 * P->n_lights;
            for (int i = 0; i < 9; ++i) l[i] = get_float(&st);
            float cutoff_deg = get_float(&st);
            l[9] = cutoff_deg * 0.01745329251994329576923690768489f; /* glm::radians, main_cli.cpp:135 */
            l[10] = (float) get_int(&st);
            l[11] = get_float(&st);
            P->n_lights++;
        }
    }
    free(buf);
    return P;
}
void or_parsed_free(OrParsed *p) { if (!p) return; free(p->kind); free(p->group); free(p->geom); free(p->mtl); free(p->lights); free(p); }

/* =====================================================================================
 * Intersection (object.cpp:16-121, geometric.cuh:240-259, cpu_bdpt.cpp:30-107)
 * ===================================================================================== */
/* Triangle::check_intersect, object.cpp:72-95 */
static int tri_check_intersect(const Obj *o, v3 rp, v3 rv, float *t, float tMin, float tMax) {
    v3 e1 = f3_sub(o->v1, o->v0), e2 = f3_sub(o->v2, o->v0);
    v3 pvec = glm_cross(rv, e2);
    float det = glm_dot(e1, pvec);
    if ((double) fabsf(det) < 1e-6) return 0;          /* ESP is the double literal 1e-6, object.h:7 */
    float invDet = 1.0f / det;
    v3 tvec = f3_sub(rp, o->v0);
    float u = glm_dot(tvec, pvec) * invDet;
    if (u < 0.0f || u > 1.0f) return 0;
    v3 qvec = glm_cross(tvec, e1);
    float v = glm_dot(rv, qvec) * invDet;
    if (v < 0.0f || (u + v) > 1.0f) return 0;
    float tt = glm_dot(e2, qvec) * invDet;
    *t = tt;                                            /* the reference writes t before the range check */
    if (tt < tMin || tt > tMax) return 0;
    return 1;
}
/* Sphere::check_intersect, object.cpp:16-56 (scale = (1,1,1); spherical u,v are dead for BDPT) */
static int sph_check_intersect(const Obj *o, v3 rp, v3 rv, float *t, float tMin, float tMax) {
    v3 D = glm_normalize(rv);
    v3 OC = f3_sub(rp, o->v0);
    const float Sx = 1.0f, Sy = 1.0f, Sz = 1.0f;
    float a = Sx * D.x * D.x + Sy * D.y * D.y + Sz * D.z * D.z;
    float b = 2.0f * (Sx * D.x * OC.x + Sy * D.y * OC.y + Sz * D.z * OC.z);
    float c = Sx * OC.x * OC.x + Sy * OC.y * OC.y + Sz * OC.z * OC.z - o->r * o->r;
    if (c <= 1e-6f) return 0;                           /* origin on/inside the sphere: refused, :26-29 */
    float disc = b * b - 4.0f * a * c;
    if (disc < 0.0f) return 0;
    float sdisc = sqrtf(0.0f < disc ? disc : 0.0f);     /* std::max(0.0f, disc) */
    float t0 = (-b - sdisc) / (2.0f * a);
    float t1 = (-b + sdisc) / (2.0f * a);
    if (t0 > t1) { float tmp = t0; t0 = t1; t1 = tmp; }
    float tc = (t0 >= tMin) ? t0 : t1;
    if (tc < tMin || tc > tMax) return 0;
    *t = tc;
    return 1;
}
/* AABB::intersectAABB, object.cpp:104-121 -- including the in-place widening of degenerate extents */
static int group_intersect(Group *g, v3 rp, v3 rv, float tmin, float tmax) {
    float *mn = &g->mn.x, *mx = &g->mx.x;
    const float *p = &rp.x, *d = &rv.x;
    for (int a = 0; a < 3; ++a) {
        if (mx[a] - mn[a] < 1e-6f) { mn[a] -= 0.5f * 1e-6f; mx[a] += 0.5f * 1e-6f; }
        float invD = 1.0f / d[a];
        float t0 = (mn[a] - p[a]) * invD;
        float t1 = (mx[a] - p[a]) * invD;
        if (invD < 0) { float tmp = t0; t0 = t1; t1 = tmp; }
        tmin = t0 > tmin ? t0 : tmin;
        tmax = t1 < tmax ? t1 : tmax;
        if (tmax <= tmin) return 0;
    }
    return 1;
}
/* intersect_sphere (light balls), geometric.cuh:240-259 */
static int light_ball_intersect(v3 ro, v3 rd, v3 center, float r, float *t, float max_dist) {
    v3 oc = f3_sub(ro, center);
    float b = f3_dot(oc, rd);
    float c = f3_dot(oc, oc) - r * r;
    float h = b * b - c;
    if (h < 0.0f) return 0;
    h = sqrtf(h);
    float th = -b - h;
    if (th > OR_EPSILON && th < max_dist) { *t = th; return 1; }
    th = -b + h;
    if (th > OR_EPSILON && th < max_dist) { *t = th; return 1; }
    return 0;
}

typedef struct { int hit; float t; v3 pos, normal; Mtl mtl; int is_light; int id; } Hit; /* CudaHit, geometric.cuh:44-51 (+id) */

/* cpu_find_closest_hit, cpu_bdpt.cpp:30-80 */
static Hit find_closest_hit(OrScene *s, v3 ro, v3 rd) {
    Hit best; memset(&best, 0, sizeof(best));
    best.hit = 0; best.t = 1e20f; best.id = -1;
    for (int gi = 0; gi < s->n_groups; ++gi) {
        Group *g = &s->groups[gi];
        if (!group_intersect(g, ro, rd, 1e-4f, best.t)) continue;
        for (int k = 0; k < g->n; ++k) {
            const Obj *o = &s->objs[g->objs[k]];
            float t;
            int h = (o->kind == 0) ? sph_check_intersect(o, ro, rd, &t, 1e-4f, best.t) : tri_check_intersect(o, ro, rd, &t, 1e-4f, best.t);
            if (h) {
                best.hit = 1; best.t = t; best.mtl = o->mtl; best.id = o->obj_id;
                v3 P = f3_add(ro, f3_muls(rd, t));                     /* ray.point + ray.vec * t */
                v3 N;
                if (o->kind == 0) N = glm_normalize(mk3(1.0f * (P.x - o->v0.x), 1.0f * (P.y - o->v0.y), 1.0f * (P.z - o->v0.z))); /* object.cpp:58-66 */
                else N = glm_normalize(glm_cross(f3_sub(o->v1, o->v0), f3_sub(o->v2, o->v0)));                                   /* object.cpp:68-70 */
                if (glm_dot(N, rd) > 0.0f) N = mk3(-N.x, -N.y, -N.z);
                best.pos = P; best.normal = N; best.is_light = 0;
            }
        }
    }
    for (int k = 0; k < s->n_lights; ++k) {
        const Light *L = &s->lights[k];
        float t;
        if (light_ball_intersect(ro, rd, L->pos, L->ball_r, &t, best.t)) {
            best.hit = 1; best.t = t; best.id = -2 - k;
            best.mtl.base_color = L->illum; best.mtl.eta = 0.0f; best.mtl.roughness = 1.0f; best.mtl.metallic = 0.0f;
            best.pos = f3_add(ro, f3_muls(rd, t));
            best.normal = f3_normalize(f3_sub(best.pos, L->pos));
            best.is_light = 1;
            if (f3_dot(best.normal, rd) > 0.0f) best.normal = f3_muls(best.normal, -1.0f);
        }
    }
    return best;
}
/* cpu_check_visibility, cpu_bdpt.cpp:82-107: 1 = unoccluded.  Objects with eta > 0 never block; light balls never block. */
static float check_visibility(OrScene *s, v3 p1, v3 p2) {
    v3 diff = f3_sub(p2, p1);
    float dist = f3_length(diff);
    v3 dir = f3_divs(diff, dist);
    float max_dist = dist - 1e-3f;
    for (int gi = 0; gi < s->n_groups; ++gi) {
        Group *g = &s->groups[gi];
        float tMin = 1e-3f, tMax = max_dist;
        if (!group_intersect(g, p1, dir, tMin, tMax)) continue;
        for (int k = 0; k < g->n; ++k) {
            const Obj *o = &s->objs[g->objs[k]];
            float t;
            int h = (o->kind == 0) ? sph_check_intersect(o, p1, dir, &t, tMin, tMax) : tri_check_intersect(o, p1, dir, &t, tMin, tMax);
            if (h && o->mtl.eta <= 0.0f) return 0.0f;
        }
    }
    return 1.0f;
}

void or_closest_hit(OrScene *s, int n, const float *org, const float *dir, int32_t *id_out, float *t_out,
                    float *pos_out, float *nrm_out, float *mtl_out) {
    for (int i = 0; i < n; ++i) {
        Hit h = find_closest_hit(s, ld3(org + 3 * i), ld3(dir + 3 * i));
        id_out[i] = h.id;
        t_out[i] = h.hit ? h.t : 1e20f;
        if (pos_out) st3(pos_out + 3 * i, h.hit ? h.pos : mk3(0, 0, 0));
        if (nrm_out) st3(nrm_out + 3 * i, h.hit ? h.normal : mk3(0, 0, 0));
        if (mtl_out) {
            float *m = mtl_out + 6 * i;
            if (h.hit) { st3(m, h.mtl.base_color); m[3] = h.mtl.roughness; m[4] = h.mtl.metallic; m[5] = h.mtl.eta; }
            else memset(m, 0, 6 * sizeof(float));
        }
    }
}
void or_visibility(OrScene *s, int n, const float *p1, const float *p2, float *vis_out) {
    for (int i = 0; i < n; ++i) vis_out[i] = check_visibility(s, ld3(p1 + 3 * i), ld3(p2 + 3 * i));
}

/* =====================================================================================
 * BSDF (geometric.cuh:119-221, 419-562)
 * ===================================================================================== */
static void build_local_frame(v3 N, v3 *T, v3 *B) {                     /* :119-123 */
    if (fabsf(N.z) < 0.999f) *T = f3_normalize(f3_cross(mk3(0, 0, 1), N));
    else *T = f3_normalize(f3_cross(mk3(0, 1, 0), N));
    *B = f3_cross(N, *T);
}
static v3 world_to_local(v3 v, v3 T, v3 B, v3 N) { return mk3(f3_dot(v, T), f3_dot(v, B), f3_dot(v, N)); } /* :124 */
static v3 local_to_world(v3 v, v3 T, v3 B, v3 N) {                      /* :127-133 */
    return mk3(T.x * v.x + B.x * v.y + N.x * v.z, T.y * v.x + B.y * v.y + N.y * v.z, T.z * v.x + B.z * v.y + N.z * v.z);
}
static float Cos2Theta(v3 w) { return w.z * w.z; }
static float AbsCosTheta(v3 w) { return fabsf(w.z); }
static float Sin2Theta(v3 w) { return fmaxf(0.0f, 1.0f - Cos2Theta(w)); }
static float SinTheta(v3 w) { return sqrtf(Sin2Theta(w)); }
static float TanTheta(v3 w) { return SinTheta(w) / (w.z + 1e-7f); }                       /* :141 */
static float Tan2Theta(v3 w) { return Sin2Theta(w) / (Cos2Theta(w) + 1e-7f); }            /* :142 */

static float FrDielectric(float cosThetaI, float etaI, float etaT) {   /* :145-160 */
    cosThetaI = fmaxf(-1.0f, fminf(1.0f, cosThetaI));
    int entering = cosThetaI > 0.0f;
    if (!entering) { float tmp = etaI; etaI = etaT; etaT = tmp; cosThetaI = fabsf(cosThetaI); }
    float sinThetaI = sqrtf(fmaxf(0.0f, 1.0f - cosThetaI * cosThetaI));
    float sinThetaT = etaI / etaT * sinThetaI;
    if (sinThetaT >= 1.0f) return 1.0f;
    float cosThetaT = sqrtf(fmaxf(0.0f, 1.0f - sinThetaT * sinThetaT));
    float Rparl = ((etaT * cosThetaI) - (etaI * cosThetaT)) / ((etaT * cosThetaI) + (etaI * cosThetaT));
    float Rperp = ((etaI * cosThetaI) - (etaT * cosThetaT)) / ((etaI * cosThetaI) + (etaT * cosThetaT));
    return (Rparl * Rparl + Rperp * Rperp) / 2.0f;
}
static v3 FrSchlick(float cosThetaI, v3 R0) {                           /* :163-167 */
    float c = fmaxf(0.0f, 1.0f - cosThetaI);
    float c5 = c * c * c * c * c;
    return f3_add(R0, f3_muls(f3_sub(mk3(1.0f, 1.0f, 1.0f), R0), c5));
}
static float RoughnessToAlpha(float roughness) { float x = fmaxf(roughness, 1e-3f); return x * x; } /* :173 */
static float TrowbridgeReitzD(v3 wh, float alpha) {                     /* :178-185 */
    float tan2Theta = Tan2Theta(wh);
    if (isinf(tan2Theta)) return 0.0f;
    float cos4Theta = Cos2Theta(wh) * Cos2Theta(wh);
    float e = (cos4Theta * (alpha * alpha + tan2Theta * tan2Theta));
    if (e < 1e-12f) return 0.0f;
    return (alpha * alpha) / (OR_PI * e);
}
static float TrowbridgeReitzLambda(v3 w, float alpha) {                 /* :187-192 */
    float absTanTheta = fabsf(TanTheta(w));
    if (isinf(absTanTheta)) return 0.0f;
    float alpha2Tan2Theta = (alpha * absTanTheta) * (alpha * absTanTheta);
    return (-1.0f + sqrtf(1.0f + alpha2Tan2Theta)) / 2.0f;
}
static float TrowbridgeReitzG(v3 wo, v3 wi, float alpha) {              /* :194-197 */
    return 1.0f / (1.0f + TrowbridgeReitzLambda(wo, alpha) + TrowbridgeReitzLambda(wi, alpha));
}
static v3 SampleTRVisibleNormal(v3 wo, float alpha, float u1, float u2) { /* :200-221 */
    v3 V = f3_normalize(mk3(alpha * wo.x, alpha * wo.y, wo.z));
    v3 T1 = (V.z < 0.9999f) ? f3_normalize(f3_cross(mk3(0, 0, 1), V)) : mk3(1, 0, 0);
    v3 T2 = f3_cross(V, T1);
    float r = sqrtf(u1);
    float phi = 2.0f * OR_PI * u2;
    float t1 = r * cosf(phi);
    float t2 = r * sinf(phi);
    float s = 0.5f * (1.0f + V.z);
    t2 = (1.0f - s) * sqrtf(fmaxf(0.0f, 1.0f - t1 * t1)) + s * t2;
    v3 Nh = f3_add(f3_add(f3_muls(T1, t1), f3_muls(T2, t2)), f3_muls(V, sqrtf(fmaxf(0.0f, 1.0f - t1 * t1 - t2 * t2))));
    return f3_normalize(mk3(alpha * Nh.x, alpha * Nh.y, fmaxf(0.0f, Nh.z)));
}

static v3 bsdf_evaluate(const Mtl *mtl, v3 wo_w, v3 wi_w, v3 N) {       /* :419-456 */
    v3 T, B;
    build_local_frame(N, &T, &B);
    v3 wo = world_to_local(wo_w, T, B, N);
    v3 wi = world_to_local(wi_w, T, B, N);
    if (wo.z == 0.0f || wi.z == 0.0f) return mk3(0, 0, 0);
    if (mtl->eta > 0.0f && mtl->roughness < 0.001f) return mk3(0, 0, 0);
    float alpha = RoughnessToAlpha(mtl->roughness);
    v3 wh_vec = f3_add(wo, wi);
    if (f3_length(wh_vec) < 1e-6f) return mk3(0, 0, 0);
    v3 wh = f3_normalize(wh_vec);
    if (wh.z < 0.0f) wh = f3_muls(wh, -1.0f);
    v3 diffuse = f3_muls(f3_divs(mtl->base_color, OR_PI), (1.0f - mtl->metallic));
    if (wo.z * wi.z < 0.0f) diffuse = mk3(0, 0, 0);
    float D = TrowbridgeReitzD(wh, alpha);
    float G = TrowbridgeReitzG(wo, wi, alpha);
    v3 F;
    if (mtl->metallic > 0.0f) F = FrSchlick(AbsCosTheta(wo), mtl->base_color);
    else { float fr = FrDielectric(f3_dot(wo, wh), 1.0f, mtl->eta); F = mk3(fr, fr, fr); }
    v3 specular = f3_divs(f3_muls(f3_muls(F, D), G), fmaxf(4.0f * AbsCosTheta(wo) * AbsCosTheta(wi), 1e-4f));
    if (wo.z * wi.z > 0.0f) return f3_add(diffuse, specular);
    return diffuse;
}
static float bsdf_pdf(const Mtl *mtl, v3 wo_w, v3 wi_w, v3 N) {         /* :458-484 */
    v3 T, B;
    build_local_frame(N, &T, &B);
    v3 wo = world_to_local(wo_w, T, B, N);
    v3 wi = world_to_local(wi_w, T, B, N);
    if (wo.z * wi.z <= 0.0f) return 0.0f;
    if (mtl->eta > 0.0f && mtl->roughness < 0.001f) return 0.0f;
    float alpha = RoughnessToAlpha(mtl->roughness);
    v3 wh_vec = f3_add(wo, wi);
    if (f3_length(wh_vec) < 1e-6f) return 0.0f;
    v3 wh = f3_normalize(wh_vec);
    if (wh.z < 0.0f) wh = f3_muls(wh, -1.0f);
    float pdf_diffuse = AbsCosTheta(wi) / OR_PI;
    float G1 = 1.0f / (1.0f + TrowbridgeReitzLambda(wo, alpha));
    float pdf_wh = TrowbridgeReitzD(wh, alpha) * G1 * fmaxf(0.0f, f3_dot(wo, wh)) / AbsCosTheta(wo);
    float pdf_specular = pdf_wh / (4.0f * f3_dot(wo, wh) + 1e-7f);
    float spec_weight = mtl->metallic > 0.0f ? 1.0f : 0.5f;
    float diff_weight = 1.0f - spec_weight;
    return diff_weight * pdf_diffuse + spec_weight * pdf_specular;
}
/* Outputs the reference leaves untouched on early returns keep the caller's values (as in C++ by-reference). */
static void bsdf_sample(const Mtl *mtl, v3 wo_w, v3 N, float u_rr, float u1, float u2, float current_eta,
                        v3 *wi_w, v3 *bsdf_val, float *pdf, int *is_delta, float *new_eta) { /* :486-562 */
    *is_delta = 0;
    *new_eta = current_eta;
    v3 T, B;
    build_local_frame(N, &T, &B);
    v3 wo = world_to_local(wo_w, T, B, N);
    v3 wi;
    if (mtl->eta > 0.0f && mtl->roughness < 0.001f && mtl->metallic < 0.01f) {
        *is_delta = 1;
        float F = FrDielectric(wo.z, current_eta, mtl->eta);
        if (u_rr < F) {
            wi = mk3(-wo.x, -wo.y, wo.z);
            *pdf = F;
            *bsdf_val = f3_divs(mk3(F, F, F), AbsCosTheta(wi));
        } else {
            float eta = wo.z > 0.0f ? (current_eta / mtl->eta) : (mtl->eta / current_eta);
            float sin2ThetaI = fmaxf(0.0f, 1.0f - Cos2Theta(wo));
            float sin2ThetaT = eta * eta * sin2ThetaI;
            if (sin2ThetaT >= 1.0f) { *pdf = 0.0f; return; }
            float cosThetaT = sqrtf(1.0f - sin2ThetaT);
            if (wo.z > 0.0f) cosThetaT = -cosThetaT;
            wi = mk3(-eta * wo.x, -eta * wo.y, cosThetaT);
            *new_eta = (wo.z > 0.0f) ? mtl->eta : 1.0f;
            *pdf = 1.0f - F;
            *bsdf_val = f3_divs(f3_muls(mtl->base_color, (1.0f - F)), AbsCosTheta(wi));
        }
        *wi_w = local_to_world(wi, T, B, N);
        return;
    }
    if (mtl->metallic > 0.99f && mtl->roughness < 0.001f) {
        *is_delta = 1;
        wi = mk3(-wo.x, -wo.y, wo.z);
        *pdf = 1.0f;
        *bsdf_val = f3_divs(FrSchlick(AbsCosTheta(wo), mtl->base_color), AbsCosTheta(wi));
        *wi_w = local_to_world(wi, T, B, N);
        return;
    }
    float alpha = RoughnessToAlpha(mtl->roughness);
    float spec_weight = mtl->metallic > 0.0f ? 1.0f : 0.5f;
    if (u_rr < spec_weight) {
        v3 wh = SampleTRVisibleNormal(wo.z > 0 ? wo : f3_muls(wo, -1.0f), alpha, u1, u2);
        if (wo.z < 0.0f) wh = f3_muls(wh, -1.0f);
        wi = f3_reflect(f3_muls(wo, -1.0f), wh);
        if (wo.z * wi.z <= 0.0f) { *pdf = 0.0f; return; }
    } else {
        float r = sqrtf(u1);
        float phi = 2.0f * OR_PI * u2;
        wi = mk3(r * cosf(phi), r * sinf(phi), sqrtf(fmaxf(0.0f, 1.0f - u1)));
        if (wo.z < 0.0f) wi.z *= -1.0f;
    }
    *wi_w = local_to_world(wi, T, B, N);
    *pdf = bsdf_pdf(mtl, wo_w, *wi_w, N);
    *bsdf_val = bsdf_evaluate(mtl, wo_w, *wi_w, N);
}

static Mtl ld_mtl(const float *m) { Mtl r; r.base_color = ld3(m); r.roughness = m[3]; r.metallic = m[4]; r.eta = m[5]; return r; }
void or_bsdf_evaluate(int n, const float *mtl, const float *wo, const float *wi, const float *N, float *out) {
    for (int i = 0; i < n; ++i) { Mtl m = ld_mtl(mtl + 6 * i); st3(out + 3 * i, bsdf_evaluate(&m, ld3(wo + 3 * i), ld3(wi + 3 * i), ld3(N + 3 * i))); }
}
void or_bsdf_pdf(int n, const float *mtl, const float *wo, const float *wi, const float *N, float *out) {
    for (int i = 0; i < n; ++i) { Mtl m = ld_mtl(mtl + 6 * i); out[i] = bsdf_pdf(&m, ld3(wo + 3 * i), ld3(wi + 3 * i), ld3(N + 3 * i)); }
}
void or_bsdf_sample(int n, const float *mtl, const float *wo, const float *N, const float *u, const float *cur_eta,
                    float *wi_out, float *f_out, float *pdf_out, int32_t *delta_out, float *eta_out) {
    for (int i = 0; i < n; ++i) {
        Mtl m = ld_mtl(mtl + 6 * i);
        v3 wi = mk3(0, 0, 0), f = mk3(0, 0, 0);
        float pdf = 0.0f, ne = 0.0f; int isd = 0;
        bsdf_sample(&m, ld3(wo + 3 * i), ld3(N + 3 * i), u[3 * i], u[3 * i + 1], u[3 * i + 2], cur_eta[i], &wi, &f, &pdf, &isd, &ne);
        st3(wi_out + 3 * i, wi); st3(f_out + 3 * i, f); pdf_out[i] = pdf; delta_out[i] = isd; eta_out[i] = ne;
    }
}

/* =====================================================================================
 * Subpath vertices and MIS (bdpt_cu.cuh:6-27, cpu_bdpt.cpp:112-167)
 * ===================================================================================== */
typedef struct { v3 pos, normal, throughput; Mtl mtl; int is_light_source, is_parallel; float source_cutoff, pdf_fwd, pdf_rev; } LightVertex;
typedef struct { v3 pos, normal, throughput; Mtl mtl; float pdf_fwd, pdf_rev; } EyeVertex;

static int is_valid_color(v3 c) {                                        /* cpu_bdpt.cpp:12-16 */
    return !(isnan(c.x) || isnan(c.y) || isnan(c.z) || isinf(c.x) || isinf(c.y) || isinf(c.z) || c.x < 0.0f || c.y < 0.0f || c.z < 0.0f);
}
static v3 clamp_radiance(v3 c, float max_val) {                          /* cpu_bdpt.cpp:18-25 */
    float max_channel = max3f(c.x, c.y, c.z);
    if (max_channel > max_val) { float scale = max_val / max_channel; return mk3(c.x * scale, c.y * scale, c.z * scale); }
    return c;
}

static float calculate_mis_weight(const EyeVertex *eye_path, int s_idx, const LightVertex *light_path, int t_idx,
                                  v3 dir_e_to_l, float dist2, v3 camera_pos) { /* cpu_bdpt.cpp:112-167 */
    if (s_idx < 0 || t_idx < 0) return 0.0f;
    const EyeVertex *ev = &eye_path[s_idx];
    const LightVertex *lv = &light_path[t_idx];
    v3 ns = f3_normalize(ev->normal);
    v3 nt = f3_normalize(lv->normal);
    float cos_s = fmaxf(0.0f, f3_dot(ns, dir_e_to_l));
    float cos_t = fmaxf(0.0f, f3_dot(nt, f3_muls(dir_e_to_l, -1.0f)));
    if (cos_s <= 0.0f || cos_t <= 0.0f || dist2 < 1e-6f) return 0.0f;
    v3 wo_s = (s_idx == 0) ? f3_normalize(f3_sub(camera_pos, ev->pos)) : f3_normalize(f3_sub(eye_path[s_idx - 1].pos, ev->pos));
    v3 wo_t = (t_idx == 0) ? f3_normalize(lv->normal) : f3_normalize(f3_sub(light_path[t_idx - 1].pos, lv->pos));
    float pdf_omega_s = fmaxf(bsdf_pdf(&ev->mtl, wo_s, dir_e_to_l, ns), 1e-6f);
    float pdf_omega_t = fmaxf(bsdf_pdf(&lv->mtl, wo_t, f3_muls(dir_e_to_l, -1.0f), nt), 1e-6f);
    float pdf_s_to_t = pdf_omega_s * cos_t / dist2;
    float pdf_t_to_s = pdf_omega_t * cos_s / dist2;
    float sum_ratios = 1.0f;
    float current_ratio = 1.0f;
    float prev_pdf_rev = pdf_t_to_s;
    for (int i = s_idx; i > 0; --i) {
        if (eye_path[i].mtl.eta > 0.0f) break;
        current_ratio *= prev_pdf_rev / fmaxf(eye_path[i].pdf_fwd, 1e-8f);
        sum_ratios += current_ratio;
        prev_pdf_rev = eye_path[i].pdf_rev;
    }
    current_ratio = 1.0f;
    prev_pdf_rev = pdf_s_to_t;
    for (int i = t_idx; i > 0; --i) {
        if (light_path[i].is_light_source) {
            current_ratio *= prev_pdf_rev / fmaxf(light_path[i].pdf_fwd, 1e-8f);
            sum_ratios += current_ratio;
            break;
        }
        if (light_path[i].mtl.eta > 0.0f) break;
        current_ratio *= prev_pdf_rev / fmaxf(light_path[i].pdf_fwd, 1e-8f);
        sum_ratios += current_ratio;
        prev_pdf_rev = light_path[i].pdf_rev;
    }
    if (isnan(sum_ratios) || isinf(sum_ratios) || sum_ratios <= 0.0f) return 0.0f;
    return 1.0f / sum_ratios;
}

float or_mis_weight(const float *eye, int n_eye, int s_idx, const float *lit, int n_lit, int t_idx,
                    const float *d_vec, float dist2, const float *cam_pos) {
    EyeVertex *E = (EyeVertex *) calloc(n_eye > 0 ? n_eye : 1, sizeof(EyeVertex));
    LightVertex *L = (LightVertex *) calloc(n_lit > 0 ? n_lit : 1, sizeof(LightVertex));
    for (int i = 0; i < n_eye; ++i) {
        const float *p = eye + 17 * i;
        E[i].pos = ld3(p); E[i].normal = ld3(p + 3); E[i].throughput = ld3(p + 6); E[i].mtl = ld_mtl(p + 9);
        E[i].pdf_fwd = p[15]; E[i].pdf_rev = p[16];
    }
    for (int i = 0; i < n_lit; ++i) {
        const float *p = lit + 20 * i;
        L[i].pos = ld3(p); L[i].normal = ld3(p + 3); L[i].throughput = ld3(p + 6); L[i].mtl = ld_mtl(p + 9);
        L[i].is_light_source = p[15] != 0.0f; L[i].is_parallel = p[16] != 0.0f; L[i].source_cutoff = p[17];
        L[i].pdf_fwd = p[18]; L[i].pdf_rev = p[19];
    }
    float w = calculate_mis_weight(E, s_idx, L, t_idx, ld3(d_vec), dist2, ld3(cam_pos));
    free(E); free(L);
    return w;
}

/* =====================================================================================
 * Light-vertex pool, cpu_bdpt.cpp:181-187, 202-325
 * ===================================================================================== */
typedef struct { uint64_t closest, shadow, path_samples, eye_samples, delta_cap; } Counters;

static void scene_bounds(OrScene *s, v3 *mn, v3 *mx) {                   /* :181-187 (glm::min/max keep the first on ties/NaN) */
    v3 c_min = mk3(1e9f, 1e9f, 1e9f), c_max = mk3(-1e9f, -1e9f, -1e9f);
    for (int i = 0; i < s->n_groups; ++i) {
        const Group *g = &s->groups[i];
        c_min = mk3(g->mn.x < c_min.x ? g->mn.x : c_min.x, g->mn.y < c_min.y ? g->mn.y : c_min.y, g->mn.z < c_min.z ? g->mn.z : c_min.z);
        c_max = mk3(c_max.x < g->mx.x ? g->mx.x : c_max.x, c_max.y < g->mx.y ? g->mx.y : c_max.y, c_max.z < g->mx.z ? g->mx.z : c_max.z);
    }
    *mn = c_min; *mx = c_max;
}

static void trace_light_path(OrScene *s, int idx, int light_depth, int spl, v3 min_bound, v3 max_bound, Rng *rng,
                             int max_delta, LightVertex *light_vertices, Counters *cnt) { /* :218-325 body */
    int total_lights = s->n_lights;
    int light_idx = idx % total_lights;
    Light light = s->lights[light_idx];
    int path_base_idx = idx * light_depth;
    v3 lightray_point, lightray_dir;
    float ray_refract = 1.0f;
    rng_begin(rng, 1u, (uint32_t) idx, 0u);
    if (light.is_parallel) {                                             /* :227-240 */
        lightray_dir = f3_normalize(light.dir);
        v3 scene_center = f3_muls(f3_add(min_bound, max_bound), 0.5f);
        float scene_radius = f3_length(f3_sub(max_bound, min_bound)) * 0.5f;
        v3 w = lightray_dir;
        v3 u_vec = (fabsf(w.x) > 0.9f) ? mk3(0, 1, 0) : mk3(1, 0, 0);
        v3 v_vec = f3_normalize(f3_cross(w, u_vec));
        u_vec = f3_normalize(f3_cross(v_vec, w));
        float r1, r2; rng_draw2(rng, &r1, &r2);
        float offset_u = (r1 - 0.5f) * scene_radius * 2.0f;
        float offset_v = (r2 - 0.5f) * scene_radius * 2.0f;
        lightray_point = f3_add(f3_add(f3_sub(scene_center, f3_muls(lightray_dir, (scene_radius * 2.0f))), f3_muls(u_vec, offset_u)), f3_muls(v_vec, offset_v));
    } else {                                                             /* :241-254 */
        lightray_point = light.pos;
        v3 w = f3_normalize(light.dir);
        v3 u_vec = (fabsf(w.x) > 0.9f) ? mk3(0, 1, 0) : mk3(1, 0, 0);
        v3 v_vec = f3_normalize(f3_cross(w, u_vec));
        u_vec = f3_normalize(f3_cross(v_vec, w));
        float u1, u2; rng_draw2(rng, &u1, &u2);
        float theta = acosf(1.0f - u1 * (1.0f - cosf(light.cutoff)));
        float phi = 2.0f * OR_PI * u2;
        v3 local_dir = mk3(sinf(theta) * cosf(phi), sinf(theta) * sinf(phi), cosf(theta));
        lightray_dir = f3_normalize(f3_add(f3_add(f3_muls(u_vec, local_dir.x), f3_muls(v_vec, local_dir.y)), f3_muls(w, local_dir.z)));
        lightray_point = f3_add(lightray_point, f3_muls(lightray_dir, light.ball_r));
    }
    v3 throughput = f3_divs(light.illum, fmaxf((float) spl, 1.0f));      /* :256 */
    LightVertex *vertex0 = &light_vertices[path_base_idx];
    vertex0->pos = lightray_point; vertex0->normal = lightray_dir; vertex0->throughput = throughput;
    vertex0->is_light_source = 1; vertex0->source_cutoff = light.cutoff; vertex0->is_parallel = light.is_parallel;
    v3 last_normal = lightray_dir, last_pos = lightray_point;
    float last_pdf_omega = 1.0f / OR_PI;
    int n_delta = 0;
    v3 wi = mk3(0, 0, 0), bsdf_val = mk3(0, 0, 0);                        /* uninitialised in the reference (:293) */
    for (int depth = 1; depth < light_depth; depth++) {                  /* :271-324 */
        LightVertex *vertex = &light_vertices[path_base_idx + depth];
        vertex->throughput = mk3(0, 0, 0);
        Hit hit = find_closest_hit(s, lightray_point, lightray_dir);
        cnt->closest++;
        if (!hit.hit) break;
        if (hit.is_light) {
            vertex->pos = hit.pos; vertex->normal = hit.normal; vertex->throughput = throughput;
            vertex->mtl = hit.mtl; vertex->is_light_source = 1; vertex->source_cutoff = 0.0f; vertex->is_parallel = 0;
            break;
        }
        if (f3_length(throughput) < 1e-4f) break;
        float dist2 = f3_dot(f3_sub(hit.pos, last_pos), f3_sub(hit.pos, last_pos));
        if (dist2 < 1e-6f) break;
        float cos_at_hit = fabsf(f3_dot(hit.normal, f3_muls(lightray_dir, -1.0f)));
        float cos_at_prev = fabsf(f3_dot(last_normal, lightray_dir));
        float pdf_fwd = last_pdf_omega * cos_at_hit / dist2;
        v3 wo = f3_muls(lightray_dir, -1.0f);
        float pdf_omega = 0.0f, new_eta = 0.0f; int is_delta = 0;
        float u_rr, u1, u2; rng_draw3(rng, &u_rr, &u1, &u2);
        bsdf_sample(&hit.mtl, wo, hit.normal, u_rr, u1, u2, ray_refract, &wi, &bsdf_val, &pdf_omega, &is_delta, &new_eta);
        if (pdf_omega <= 0.0f && !is_delta) break;
        if (is_delta) {
            throughput = f3_mul(throughput, bsdf_val);
            lightray_dir = wi; ray_refract = new_eta;
            lightray_point = f3_add(hit.pos, f3_muls(hit.normal, (f3_dot(wi, hit.normal) < 0.0f ? -OR_EPSILON : OR_EPSILON)));
            if (max_delta > 0 && ++n_delta >= max_delta) { cnt->delta_cap++; break; }
            depth--;
            continue;
        }
        vertex->pos = hit.pos; vertex->normal = hit.normal; vertex->throughput = throughput;
        vertex->mtl = hit.mtl; vertex->is_light_source = 0;
        float pdf_rev_omega = bsdf_pdf(&hit.mtl, wi, wo, hit.normal);
        float pdf_rev = pdf_rev_omega * cos_at_prev / dist2;
        vertex->pdf_fwd = pdf_fwd; vertex->pdf_rev = pdf_rev;
        throughput = f3_divs(f3_muls(f3_mul(throughput, bsdf_val), fabsf(f3_dot(hit.normal, wi))), pdf_omega);
        if (!is_valid_color(throughput)) break;
        lightray_dir = wi;
        lightray_point = f3_add(hit.pos, f3_muls(hit.normal, OR_EPSILON));
        last_pdf_omega = pdf_omega; last_normal = hit.normal; last_pos = hit.pos;
    }
}

static void build_light_pool(OrScene *s, int light_depth, int spl, const OrRenderOpts *opts, LightVertex *pool, Counters *total) {
    v3 min_bound, max_bound;
    scene_bounds(s, &min_bound, &max_bound);
    int total_light_paths = s->n_lights * spl;
    int nthreads = (opts && opts->nthreads > 0) ? opts->nthreads : omp_get_max_threads();
    int mode = opts ? opts->rng_mode : OR_RNG_MT19937;
#pragma omp parallel num_threads(nthreads)
    {
        Mt19937 mt; mt_seed(&mt, (uint32_t) (1337 + omp_get_thread_num()));  /* :213 */
        Rng rng; memset(&rng, 0, sizeof(rng));
        rng.mode = mode; rng.mt = &mt;
        if (opts) { rng.key[0] = (uint32_t) opts->seed; rng.key[1] = (uint32_t) (opts->seed >> 32); }
        Counters c; memset(&c, 0, sizeof(c));
#pragma omp for schedule(dynamic, 64)
        for (int idx = 0; idx < total_light_paths; ++idx)
            trace_light_path(s, idx, light_depth, spl, min_bound, max_bound, &rng, opts ? opts->max_delta : 0, pool, &c);
#pragma omp critical
        { total->closest += c.closest; total->delta_cap += c.delta_cap; }
    }
}

static void pool_to_flat(const LightVertex *v, float *p) {
    st3(p, v->pos); st3(p + 3, v->normal); st3(p + 6, v->throughput);
    st3(p + 9, v->mtl.base_color); p[12] = v->mtl.roughness; p[13] = v->mtl.metallic; p[14] = v->mtl.eta;
    p[15] = (float) v->is_light_source; p[16] = (float) v->is_parallel; p[17] = v->source_cutoff; p[18] = v->pdf_fwd; p[19] = v->pdf_rev;
}
void or_light_pool(OrScene *s, int light_depth, int spl, const OrRenderOpts *opts, float *pool20) {
    int n = s->n_lights * spl * light_depth;
    LightVertex *pool = (LightVertex *) calloc(n > 0 ? n : 1, sizeof(LightVertex));
    Counters c; memset(&c, 0, sizeof(c));
    if (s->n_lights > 0) build_light_pool(s, light_depth, spl, opts, pool, &c);
    for (int i = 0; i < n; ++i) pool_to_flat(&pool[i], pool20 + 20 * i);
    free(pool);
}

/* =====================================================================================
 * Whole render, cpu_bdpt.cpp:173-488
 * ===================================================================================== */
static double now_s(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }

void or_render(OrScene *s, const float *cam, int eye_depth, int light_depth, int W, int H, int spp, int spl,
               const OrRenderOpts *opts, float *image, OrStats *stats) {
    if (stats) memset(stats, 0, sizeof(*stats));
    int total_lights = s->n_lights;
    if (total_lights == 0) return;                                       /* :178 (film left untouched) */

    /* camera basis, :190-200 (glm arithmetic) */
    v3 eye = ld3(cam), look_at = ld3(cam + 3), view_up = ld3(cam + 6);
    float fov = cam[9];
    float aspect = (float) W / (float) H;
    float theta = fov * OR_PI / 180.0f;
    float half_height = tanf(theta / 2.0f);
    float half_width = aspect * half_height;
    v3 cw = glm_normalize(f3_sub(eye, look_at));
    v3 cu = glm_normalize(glm_cross(view_up, cw));
    v3 cv = glm_cross(cw, cu);
    v3 cUL = f3_sub(f3_add(f3_sub(eye, f3_muls(cu, half_width)), f3_muls(cv, half_height)), cw);
    v3 cdx = f3_divs(f3_muls(cu, 2.0f * half_width), (float) W);
    v3 cdy = f3_divs(f3_muls(cv, -2.0f * half_height), (float) H);
    v3 cam_eye = eye;

    int total_light_paths = total_lights * spl;
    int pool_n = total_light_paths * light_depth;
    LightVertex *light_vertices = (LightVertex *) calloc(pool_n > 0 ? pool_n : 1, sizeof(LightVertex)); /* value-initialised, :203 */
    Counters total; memset(&total, 0, sizeof(total));
    int nthreads = (opts && opts->nthreads > 0) ? opts->nthreads : omp_get_max_threads();
    int mode = opts ? opts->rng_mode : OR_RNG_MT19937;
    int max_delta = opts ? opts->max_delta : 0;
    int sample_begin = (opts && mode == OR_RNG_PHILOX) ? opts->sample_begin : 0;

    double t0 = now_s();
    build_light_pool(s, light_depth, spl, opts, light_vertices, &total);

#pragma omp parallel num_threads(nthreads)
    {
        Mt19937 mt; mt_seed(&mt, (uint32_t) (9999 + omp_get_thread_num()));  /* :338 */
        Rng rng; memset(&rng, 0, sizeof(rng));
        rng.mode = mode; rng.mt = &mt;
        if (opts) { rng.key[0] = (uint32_t) opts->seed; rng.key[1] = (uint32_t) (opts->seed >> 32); }
        EyeVertex *local_eye_path = (EyeVertex *) calloc(eye_depth > 0 ? eye_depth : 1, sizeof(EyeVertex));
        Counters c; memset(&c, 0, sizeof(c));
#pragma omp for schedule(dynamic, 16)
        for (int idx = 0; idx < W * H; ++idx) {
            int px = idx % W, py = idx / W;
            v3 pixel_accum_color = mk3(0, 0, 0);
            for (int sidx = 0; sidx < spp; ++sidx) {
                c.eye_samples++;
                rng_begin(&rng, 2u, (uint32_t) idx, (uint32_t) (sample_begin + sidx));
                float jx, jy; rng_draw2(&rng, &jx, &jy);
                float pixel_x = (float) px + jx;
                float pixel_y = (float) py + jy;
                v3 pixel_pos = f3_add(f3_add(cUL, f3_muls(cdx, pixel_x)), f3_muls(cdy, pixel_y));
                v3 eyeray_point = cam_eye;
                v3 eyeray_dir = f3_normalize(f3_sub(pixel_pos, cam_eye));
                float eyeray_refract = 1.0f;
                v3 throughput = mk3(1.0f, 1.0f, 1.0f);
                v3 last_normal = eyeray_dir, last_pos = cam_eye;
                float last_pdf_omega = 1.0f;
                v3 final_color = mk3(0, 0, 0);
                int n_delta = 0;
                v3 wi = mk3(0, 0, 0), bsdf_val = mk3(0, 0, 0);

                for (int depth = 0; depth < eye_depth; depth++) {        /* :365-473 */
                    EyeVertex *vertex = &local_eye_path[depth];
                    vertex->throughput = mk3(0, 0, 0);
                    Hit hit = find_closest_hit(s, eyeray_point, eyeray_dir);
                    c.closest++;
                    if (!hit.hit) break;
                    if (hit.is_light && depth == 0) { final_color = f3_add(final_color, hit.mtl.base_color); c.path_samples++; break; }
                    float pdf_fwd = 1.0f;
                    if (depth > 0) {
                        float dist2 = f3_dot(f3_sub(hit.pos, last_pos), f3_sub(hit.pos, last_pos));
                        float cos_at_hit = fabsf(f3_dot(hit.normal, f3_muls(eyeray_dir, -1.0f)));
                        pdf_fwd = last_pdf_omega * cos_at_hit / fmaxf(dist2, 1e-6f);
                    }
                    vertex->pos = hit.pos; vertex->normal = hit.normal; vertex->throughput = throughput;
                    vertex->mtl = hit.mtl; vertex->pdf_fwd = 0.0f; vertex->pdf_rev = 1.0f;

                    /* connection loop, :388-440 */
                    v3 total_L = mk3(0, 0, 0);
                    for (int light_idx = 0; light_idx < total_light_paths * light_depth; light_idx++) {
                        const LightVertex *lv = &light_vertices[light_idx];
                        if (f3_length(lv->throughput) < 1e-6f) continue;
                        v3 d_vec = f3_sub(lv->pos, vertex->pos);
                        float dist2 = f3_dot(d_vec, d_vec);
                        if (dist2 < 1e-6f) continue;
                        float dist = sqrtf(dist2);
                        v3 wi_c = f3_divs(d_vec, dist);
                        float cosE = fmaxf(0.0f, f3_dot(vertex->normal, wi_c));
                        float cosL = fmaxf(0.0f, f3_dot(lv->normal, f3_muls(wi_c, -1.0f)));
                        if (cosE <= 0.0f || cosL <= 0.0f) continue;
                        int current_t_idx = light_idx % light_depth;
                        if (lv->is_light_source && lv->source_cutoff > 0.0f && !lv->is_parallel) {
                            int real_light_idx = (light_idx / light_depth) % total_lights;
                            v3 light_dir = f3_normalize(s->lights[real_light_idx].dir);
                            if (f3_dot(light_dir, f3_muls(wi_c, -1.0f)) < cosf(lv->source_cutoff)) continue;
                        }
                        v3 wo_e = f3_muls(eyeray_dir, -1.0f);
                        v3 fE = bsdf_evaluate(&vertex->mtl, wo_e, wi_c, vertex->normal);
                        v3 fL = mk3(1.0f, 1.0f, 1.0f);
                        if (!lv->is_light_source && current_t_idx > 0) {
                            v3 prev_lv_pos = light_vertices[light_idx - 1].pos;
                            v3 wo_l = f3_normalize(f3_sub(prev_lv_pos, lv->pos));
                            fL = bsdf_evaluate(&lv->mtl, wo_l, f3_muls(wi_c, -1.0f), lv->normal);
                        }
                        if ((fE.x <= 0.0f && fE.y <= 0.0f && fE.z <= 0.0f) || (fL.x <= 0.0f && fL.y <= 0.0f && fL.z <= 0.0f)) continue;
                        float vis = check_visibility(s, f3_add(vertex->pos, f3_muls(vertex->normal, OR_EPSILON)),
                                                     f3_add(lv->pos, f3_muls(lv->normal, OR_EPSILON)));
                        c.shadow++;
                        if (vis <= 0.0f) continue;
                        v3 transmittance = mk3(vis, vis, vis);
                        float G = (cosE * cosL) / fmaxf(dist2, 1e-4f);
                        const LightVertex *base = &light_vertices[(light_idx / light_depth) * light_depth];
                        float mis_w = calculate_mis_weight(local_eye_path, depth, base, current_t_idx, d_vec, dist2, cam_eye);
                        c.path_samples++;
                        v3 contrib = f3_muls(f3_mul(f3_mul(f3_mul(f3_muls(f3_mul(vertex->throughput, fE), G), fL), lv->throughput), transmittance), mis_w);
                        if (is_valid_color(contrib)) { contrib = clamp_radiance(contrib, 15.0f); total_L = f3_add(total_L, contrib); }
                    }
                    final_color = f3_add(final_color, total_L);

                    /* extend the eye subpath, :442-472 */
                    v3 wo = f3_muls(eyeray_dir, -1.0f);
                    float pdf_omega = 0.0f, new_eta = 0.0f; int is_delta = 0;
                    float u_rr, u1, u2; rng_draw3(&rng, &u_rr, &u1, &u2);
                    bsdf_sample(&hit.mtl, wo, hit.normal, u_rr, u1, u2, eyeray_refract, &wi, &bsdf_val, &pdf_omega, &is_delta, &new_eta);
                    if (pdf_omega <= 0.0f && !is_delta) break;
                    if (is_delta) {
                        throughput = f3_mul(throughput, bsdf_val);
                        eyeray_dir = wi; eyeray_refract = new_eta;
                        eyeray_point = f3_add(hit.pos, f3_muls(hit.normal, (f3_dot(wi, hit.normal) < 0.0f ? -OR_EPSILON : OR_EPSILON)));
                        last_pos = hit.pos; last_normal = hit.normal; last_pdf_omega = 1.0f;
                        if (max_delta > 0 && ++n_delta >= max_delta) { c.delta_cap++; break; }
                        depth--; continue;
                    }
                    float pdf_rev_omega = bsdf_pdf(&hit.mtl, wi, wo, hit.normal);
                    float dist2 = f3_dot(f3_sub(hit.pos, last_pos), f3_sub(hit.pos, last_pos));
                    float cos_at_prev = fabsf(f3_dot(last_normal, eyeray_dir));
                    vertex->pdf_fwd = pdf_fwd;
                    vertex->pdf_rev = pdf_rev_omega * cos_at_prev / fmaxf(dist2, 1e-6f);
                    throughput = f3_divs(f3_muls(f3_mul(throughput, bsdf_val), fabsf(f3_dot(hit.normal, wi))), pdf_omega);
                    if (!is_valid_color(throughput)) break;
                    eyeray_dir = wi;
                    eyeray_point = f3_add(hit.pos, f3_muls(hit.normal, OR_EPSILON));
                    last_pdf_omega = pdf_omega; last_normal = hit.normal; last_pos = hit.pos;
                }
                if (!is_valid_color(final_color)) final_color = mk3(0, 0, 0);
                pixel_accum_color = f3_add(pixel_accum_color, final_color);
            }
            st3(image + 3 * (size_t) idx, f3_divs(pixel_accum_color, (float) spp)); /* :479 */
        }
        free(local_eye_path);
#pragma omp critical
        {
            total.closest += c.closest; total.shadow += c.shadow; total.path_samples += c.path_samples;
            total.eye_samples += c.eye_samples; total.delta_cap += c.delta_cap;
        }
    }
    double t1 = now_s();
    free(light_vertices);
    if (stats) {
        stats->eye_samples = total.eye_samples; stats->closest_rays = total.closest; stats->shadow_rays = total.shadow;
        stats->path_samples = total.path_samples; stats->delta_cap_hits = total.delta_cap; stats->seconds = t1 - t0;
    }
}

/* =====================================================================================
 * The reference's two other integrators, pt and ppm (src/pt_cu.cu, src/ppm_cu.cu), restated on the CPU.
 *
 * Parity status of THIS section: UNPINNED on the CPU.  The reference has no CPU implementation of these
 * integrators and seeds curand from the clock (pt_cu.cu:279, ppm_cu.cu:357), so nothing here can be compared bit for
 * bit with it.  The restatement follows the kernels line by line on the Philox streams of the GPU path and is pinned
 * STATISTICALLY against the reference's own kernels run on the GPU box (oracle/_ref/libref_cuda.so,
 * tests/test_gpu_pt_ppm.py).  Intersection follows the reference's GPU-side helpers, which differ from cpu_bdpt.cpp's:
 * intersect_triangle / intersect_sphere for every object (geometric.cuh:240-291), first hit wins ties in the order
 * spheres, light balls, triangles (:327-390), shadow rays are blocked by every object -- check_visibility (:293-325)
 * reads mtl_old.refract, which the pt / ppm helpers leave uninitialised (pt_cu_helper.cpp:25,41: the assignment is
 * commented out); it is taken as zero-initialised here, i.e. opaque.
 * Draw schedule (shared with path_tracing_b200/csrc/b2_pt.cuh):
 *   eye / pt paths   kind 2: (pixel, sample) block 0 = jitter, block 1+k = u_rr,u1,u2 of the k-th bsdf_sample call
 *   pt NEE           kind 4: (pixel, sample) block 1+depth lanes 0,1,2 = light choice, point on the light ball
 *   photons          kind 5: (photon, 0) block 0 = emission, block 1+k = k-th bsdf_sample call
 * ===================================================================================== */
static void philox_unit4(const uint32_t key[2], uint32_t id0, uint32_t id1, uint32_t block, uint32_t kind, float u[4]) {
    uint32_t c[4] = { id0, id1, block, kind }, o[4];
    or_philox4x32_10(c, key, o);
    for (int i = 0; i < 4; ++i) u[i] = or_u32_to_unit_float(o[i]);
}
/* intersect_triangle, geometric.cuh:261-291 */
static int gpu_intersect_triangle(v3 ro, v3 rd, const Obj *o, float *t, float max_dist) {
    v3 e1 = f3_sub(o->v1, o->v0), e2 = f3_sub(o->v2, o->v0);
    v3 h = f3_cross(rd, e2);
    float a = f3_dot(e1, h);
    if (a > -1e-6f && a < 1e-6f) return 0;
    float f = 1.0f / a;
    v3 s = f3_sub(ro, o->v0);
    float u = f * f3_dot(s, h);
    if (u < 0.0f || u > 1.0f) return 0;
    v3 q = f3_cross(s, e1);
    float v = f * f3_dot(rd, q);
    if (v < 0.0f || u + v > 1.0f) return 0;
    float th = f * f3_dot(e2, q);
    if (th > OR_EPSILON && th < max_dist) { *t = th; return 1; }
    return 0;
}
/* find_closest_hit, geometric.cuh:327-390; the helpers fill the sphere / triangle arrays in group-map order (pt_cu_helper.cpp:17-55) */
static Hit gpu_find_closest_hit(OrScene *s, v3 ro, v3 rd) {
    Hit best; memset(&best, 0, sizeof(best));
    best.hit = 0; best.t = 1e20f; best.id = -1;
    float t;
    for (int gi = 0; gi < s->n_groups; ++gi) for (int k = 0; k < s->groups[gi].n; ++k) {
        const Obj *o = &s->objs[s->groups[gi].objs[k]];
        if (o->kind != 0) continue;
        if (light_ball_intersect(ro, rd, o->v0, o->r, &t, 1e20f) && t < best.t) {
            best.hit = 1; best.t = t; best.mtl = o->mtl; best.id = o->obj_id; best.is_light = 0;
            best.pos = f3_add(ro, f3_muls(rd, t));
            best.normal = f3_normalize(f3_sub(best.pos, o->v0));
            if (f3_dot(best.normal, rd) > 0.0f) best.normal = f3_muls(best.normal, -1.0f);
        }
    }
    for (int k = 0; k < s->n_lights; ++k) {
        const Light *L = &s->lights[k];
        if (light_ball_intersect(ro, rd, L->pos, L->ball_r, &t, 1e20f) && t < best.t) {
            best.hit = 1; best.t = t; best.id = -2 - k; best.is_light = 1;
            best.mtl.base_color = L->illum;                                  /* :359; the other fields keep the previous best's */
            best.pos = f3_add(ro, f3_muls(rd, t));
            best.normal = f3_normalize(f3_sub(best.pos, L->pos));
            if (f3_dot(best.normal, rd) > 0.0f) best.normal = f3_muls(best.normal, -1.0f);
        }
    }
    for (int gi = 0; gi < s->n_groups; ++gi) for (int k = 0; k < s->groups[gi].n; ++k) {
        const Obj *o = &s->objs[s->groups[gi].objs[k]];
        if (o->kind == 0) continue;
        if (gpu_intersect_triangle(ro, rd, o, &t, 1e20f) && t < best.t) {
            best.hit = 1; best.t = t; best.mtl = o->mtl; best.id = o->obj_id; best.is_light = 0;
            best.pos = f3_add(ro, f3_muls(rd, t));
            best.normal = f3_normalize(f3_cross(f3_sub(o->v1, o->v0), f3_sub(o->v2, o->v0)));
            if (f3_dot(best.normal, rd) > 0.0f) best.normal = f3_muls(best.normal, -1.0f);
        }
    }
    return best;
}
/* check_visibility, geometric.cuh:293-325, with mtl_old zero-initialised: 1 = unoccluded */
static int gpu_check_visibility(OrScene *s, v3 p1, v3 p2) {
    v3 diff = f3_sub(p2, p1);
    float dist = f3_length(diff);
    v3 dir = f3_divs(diff, dist);
    float max_d = dist - 1e-3f, min_d = 1e-3f, t;
    for (int i = 0; i < s->n_obj; ++i) {
        const Obj *o = &s->objs[i];
        int h = (o->kind == 0) ? light_ball_intersect(p1, dir, o->v0, o->r, &t, max_d) : gpu_intersect_triangle(p1, dir, o, &t, max_d);
        if (h && t > min_d) return 0;
    }
    return 1;
}
static void camera_frame(const float *cam, int W, int H, v3 *cam_eye, v3 *cUL, v3 *cdx, v3 *cdy) {   /* cpu_bdpt.cpp:190-200, as or_render */
    v3 eye = ld3(cam), look_at = ld3(cam + 3), view_up = ld3(cam + 6);
    float aspect = (float) W / (float) H;
    float theta = cam[9] * OR_PI / 180.0f;
    float half_height = tanf(theta / 2.0f);
    float half_width = aspect * half_height;
    v3 cw = glm_normalize(f3_sub(eye, look_at));
    v3 cu = glm_normalize(glm_cross(view_up, cw));
    v3 cv = glm_cross(cw, cu);
    *cUL = f3_sub(f3_add(f3_sub(eye, f3_muls(cu, half_width)), f3_muls(cv, half_height)), cw);
    *cdx = f3_divs(f3_muls(cu, 2.0f * half_width), (float) W);
    *cdy = f3_divs(f3_muls(cv, -2.0f * half_height), (float) H);
    *cam_eye = eye;
}
static int gathers_light(const Mtl *m) { return m->eta <= 0.0f && (m->metallic < 0.99f || m->roughness > 0.01f); }   /* pt_cu.cu:126, ppm_cu.cu:221 */

/* cuda_path_trace_kernel, pt_cu.cu:20-252.  image = mean over spp. */
void or_render_pt(OrScene *s, const float *cam, int max_depth, int W, int H, int spp, const OrRenderOpts *opts, float *image, OrStats *stats) {
    if (stats) memset(stats, 0, sizeof(*stats));
    v3 cam_eye, cUL, cdx, cdy;
    camera_frame(cam, W, H, &cam_eye, &cUL, &cdx, &cdy);
    const int num_lights = s->n_lights;
    const int nthreads = (opts && opts->nthreads > 0) ? opts->nthreads : omp_get_max_threads();
    const int max_delta = opts ? opts->max_delta : 0;
    const int sample_begin = opts ? opts->sample_begin : 0;
    uint32_t key[2] = { 0u, 0u };
    if (opts) { key[0] = (uint32_t) opts->seed; key[1] = (uint32_t) (opts->seed >> 32); }
    Counters total; memset(&total, 0, sizeof(total));
    double t0 = now_s();
#pragma omp parallel num_threads(nthreads)
    {
        Counters c; memset(&c, 0, sizeof(c));
#pragma omp for schedule(dynamic, 16)
        for (int idx = 0; idx < W * H; ++idx) {
            int px = idx % W, py = idx / W;
            v3 color_sum = mk3(0, 0, 0);
            for (int sidx = 0; sidx < spp; ++sidx) {
                c.eye_samples++;
                const uint32_t sg = (uint32_t) (sample_begin + sidx);
                float u[4];
                philox_unit4(key, (uint32_t) idx, sg, 0u, 2u, u);
                v3 final_color = mk3(0, 0, 0);
                float pixel_x = (float) px + u[0], pixel_y = (float) py + u[1];
                v3 eyeray_point = cam_eye;
                v3 pixel_pos = f3_add(f3_add(cUL, f3_muls(cdx, pixel_x)), f3_muls(cdy, pixel_y));
                v3 eyeray_dir = f3_normalize(f3_sub(pixel_pos, eyeray_point));
                float eyeray_refract = 1.0f;
                v3 throughput = mk3(1.0f, 1.0f, 1.0f);
                int last_is_delta = 1, n_delta = 0;
                for (int depth = 0; depth < max_depth; ++depth) {
                    Hit hit = gpu_find_closest_hit(s, eyeray_point, eyeray_dir);
                    c.closest++;
                    if (!hit.hit) break;
                    v3 wo = f3_muls(eyeray_dir, -1.0f);
                    const uint32_t bounce = (uint32_t) (depth + n_delta);
                    if (hit.is_light) {                                       /* :58-122 */
                        v3 emission = hit.mtl.base_color;
                        float area = 1.0f, cone_ratio = 1.0f; int valid_light = 0;
                        for (int i = 0; i < num_lights; ++i) {
                            const Light *L = &s->lights[i];
                            v3 center_to_hit = f3_sub(hit.pos, L->pos);
                            if (fabsf(f3_length(center_to_hit) - L->ball_r) < 1e-2f) {
                                valid_light = 1;
                                area = 4.0f * OR_PI * L->ball_r * L->ball_r;
                                if (L->cutoff > 0.0f && !L->is_parallel) {
                                    cone_ratio = (1.0f - cosf(L->cutoff)) / 2.0f;
                                    v3 main_dir = f3_normalize(L->dir);
                                    if (depth == 0) cone_ratio = 1.f;
                                    else if (f3_dot(main_dir, f3_normalize(center_to_hit)) < cosf(L->cutoff)) cone_ratio = 0.0f;
                                }
                                break;
                            }
                        }
                        if (valid_light && cone_ratio > 0.0f) emission = f3_divs(emission, area * cone_ratio);
                        else emission = mk3(0, 0, 0);
                        if ((emission.x > 0.0f || emission.y > 0.0f || emission.z > 0.0f) && last_is_delta) {
                            v3 contrib = f3_mul(throughput, emission);
                            if (is_valid_color(contrib)) { final_color = f3_add(final_color, clamp_radiance(contrib, 15.0f)); c.path_samples++; }
                        }                                                     /* the !last_is_delta branch never adds: pdf_light_dir = 0, :105-119 */
                        break;
                    }
                    if (gathers_light(&hit.mtl) && num_lights > 0) {          /* :125-208 */
                        float un[4];
                        philox_unit4(key, (uint32_t) idx, sg, 1u + (uint32_t) depth, 4u, un);   /* keyed by the vertex depth: NEE vertices are never delta bounces */
                        int l_idx = (int) (un[0] * num_lights); if (l_idx > num_lights - 1) l_idx = num_lights - 1;
                        const Light *light = &s->lights[l_idx];
                        if (light->is_parallel) {
                            v3 light_dir = f3_normalize(f3_muls(f3_normalize(light->dir), -1.0f));
                            float cos_surface = fmaxf(0.0f, f3_dot(hit.normal, light_dir));
                            if (cos_surface > 0.0f) {
                                c.shadow++;
                                if (gpu_check_visibility(s, f3_add(hit.pos, f3_muls(hit.normal, OR_EPSILON)), f3_add(hit.pos, f3_muls(light_dir, 1e4f)))) {
                                    v3 brdf = bsdf_evaluate(&hit.mtl, wo, light_dir, hit.normal);
                                    v3 contrib = f3_muls(f3_muls(f3_mul(f3_mul(throughput, brdf), light->illum), cos_surface), (float) num_lights);
                                    if (is_valid_color(contrib)) { final_color = f3_add(final_color, clamp_radiance(contrib, 15.0f)); c.path_samples++; }
                                }
                            }
                        } else {
                            /* random_in_unit_sphere_device + normalize (:153-155) = a uniform direction; drawn from two uniforms */
                            float z = 1.0f - 2.0f * un[1], r = sqrtf(fmaxf(0.0f, 1.0f - z * z)), phi = 2.0f * OR_PI * un[2];
                            v3 d_vec_local = mk3(r * cosf(phi), r * sinf(phi), z);
                            v3 light_pos = f3_add(light->pos, f3_muls(d_vec_local, light->ball_r));
                            v3 wi_light = f3_sub(light_pos, hit.pos);
                            float dist2 = f3_dot(wi_light, wi_light);
                            float dist = sqrtf(dist2);
                            wi_light = f3_divs(wi_light, dist);
                            float cos_surface = fmaxf(0.0f, f3_dot(hit.normal, wi_light));
                            float cos_light = fmaxf(0.0f, f3_dot(d_vec_local, f3_muls(wi_light, -1.0f)));
                            if (cos_surface > 0.0f && cos_light > 0.0f) {
                                int inside_cone = 1;
                                if (light->cutoff > 0.0f && !light->is_parallel) {
                                    v3 main_dir = f3_normalize(light->dir);
                                    if (f3_dot(main_dir, f3_muls(wi_light, -1.0f)) < cosf(light->cutoff)) inside_cone = 0;
                                }
                                if (inside_cone) {
                                    c.shadow++;
                                    if (gpu_check_visibility(s, f3_add(hit.pos, f3_muls(hit.normal, OR_EPSILON)), f3_add(light_pos, f3_muls(d_vec_local, OR_EPSILON)))) {
                                        float area = 4.0f * OR_PI * light->ball_r * light->ball_r;
                                        float pdf_light_area = 1.0f / (num_lights * area);
                                        float pdf_light_dir = pdf_light_area * dist2 / fmaxf(cos_light, 1e-6f);
                                        float pdf_bsdf = bsdf_pdf(&hit.mtl, wo, wi_light, hit.normal);
                                        float p_l = pdf_light_dir * pdf_light_dir, p_b = pdf_bsdf * pdf_bsdf;
                                        float mis_w = p_l / fmaxf(p_l + p_b, 1e-8f);
                                        v3 brdf = bsdf_evaluate(&hit.mtl, wo, wi_light, hit.normal);
                                        v3 contrib = f3_muls(f3_divs(f3_muls(f3_mul(f3_mul(throughput, brdf), light->illum), cos_surface), pdf_light_dir), mis_w);
                                        if (is_valid_color(contrib)) { final_color = f3_add(final_color, clamp_radiance(contrib, 15.0f)); c.path_samples++; }
                                    }
                                }
                            }
                        }
                    }
                    v3 wi = mk3(0, 0, 0), bsdf_val = mk3(0, 0, 0);
                    float pdf_omega = 0.0f, new_eta = 0.0f; int is_delta = 0;
                    float ub[4];
                    philox_unit4(key, (uint32_t) idx, sg, 1u + bounce, 2u, ub);
                    bsdf_sample(&hit.mtl, wo, hit.normal, ub[0], ub[1], ub[2], eyeray_refract, &wi, &bsdf_val, &pdf_omega, &is_delta, &new_eta);
                    if (pdf_omega <= 0.0f) break;                             /* :222; a delta sample with pdf 0 (TIR) ends the path too (see b2_pt.cuh) */
                    if (is_delta) {
                        throughput = f3_mul(throughput, bsdf_val);
                        eyeray_dir = wi; eyeray_refract = new_eta;
                        eyeray_point = f3_add(hit.pos, f3_muls(hit.normal, f3_dot(wi, hit.normal) < 0.0f ? -OR_EPSILON : OR_EPSILON));
                        last_is_delta = 1;
                        if (!is_valid_color(throughput)) break;
                        ++n_delta;
                        if (max_delta > 0 && n_delta >= max_delta) { c.delta_cap++; break; }
                        depth--; continue;
                    }
                    float cos_wi = fabsf(f3_dot(hit.normal, wi));
                    throughput = f3_divs(f3_muls(f3_mul(throughput, bsdf_val), cos_wi), pdf_omega);
                    if (!is_valid_color(throughput)) break;
                    eyeray_dir = wi;
                    eyeray_point = f3_add(hit.pos, f3_muls(hit.normal, OR_EPSILON));
                    last_is_delta = 0;
                }
                if (!is_valid_color(final_color)) final_color = mk3(0, 0, 0);
                color_sum = f3_add(color_sum, final_color);
            }
            st3(image + 3 * (size_t) idx, f3_divs(color_sum, (float) spp));
        }
#pragma omp critical
        { total.closest += c.closest; total.shadow += c.shadow; total.path_samples += c.path_samples; total.eye_samples += c.eye_samples; total.delta_cap += c.delta_cap; }
    }
    if (stats) {
        stats->eye_samples = total.eye_samples; stats->closest_rays = total.closest; stats->shadow_rays = total.shadow;
        stats->path_samples = total.path_samples; stats->delta_cap_hits = total.delta_cap; stats->seconds = now_s() - t0;
    }
}

/* ppm_render_wrapper, ppm_cu.cu:325-398: eye pass, photon pass, resolve.  radius <= 0: PPM_RADIUS 0.05 (ppm_cu.cuh:5). */
typedef struct { v3 pos, normal, wo, thr; Mtl mtl; int pixel, gx, gy, gz; uint64_t key; } OrHitPoint;
static uint64_t cell_key(int gx, int gy, int gz) {
    return ((uint64_t) ((uint32_t) (gx + (1 << 20)) & 0x1fffffu) << 42) | ((uint64_t) ((uint32_t) (gy + (1 << 20)) & 0x1fffffu) << 21) | (uint64_t) ((uint32_t) (gz + (1 << 20)) & 0x1fffffu);
}
static int cmp_hp(const void *a, const void *b) {
    const OrHitPoint *x = (const OrHitPoint *) a, *y = (const OrHitPoint *) b;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return x->pixel - y->pixel;
}
void or_render_ppm(OrScene *s, const float *cam, int eye_depth, int light_depth, int W, int H, int spl, float radius,
                   const OrRenderOpts *opts, float *image, OrStats *stats) {
    if (stats) memset(stats, 0, sizeof(*stats));
    if (!(radius > 0.0f)) radius = 0.05f;
    v3 cam_eye, cUL, cdx, cdy;
    camera_frame(cam, W, H, &cam_eye, &cUL, &cdx, &cdy);
    const int num_lights = s->n_lights, npix = W * H;
    const int nthreads = (opts && opts->nthreads > 0) ? opts->nthreads : omp_get_max_threads();
    const int max_delta = opts ? opts->max_delta : 0;
    const uint32_t sg = (uint32_t) (opts ? opts->sample_begin : 0);
    uint32_t key[2] = { 0u, 0u };
    if (opts) { key[0] = (uint32_t) opts->seed; key[1] = (uint32_t) (opts->seed >> 32); }
    v3 min_bound, max_bound;
    scene_bounds(s, &min_bound, &max_bound);
    const float inv_cell = 1.0f / radius;
    OrHitPoint *hp = (OrHitPoint *) calloc(npix > 0 ? npix : 1, sizeof(OrHitPoint));
    int n_hp = 0;
    uint64_t n_closest = 0, n_cap = 0, n_gather = 0;
    double t0 = now_s();
    memset(image, 0, sizeof(float) * 3 * (size_t) npix);
    /* 1. ppm_eye_trace, :64-152 */
    for (int idx = 0; idx < npix && eye_depth > 0; ++idx) {
        int px = idx % W, py = idx / W;
        float u[4];
        philox_unit4(key, (uint32_t) idx, sg, 0u, 2u, u);
        v3 pixel_pos = f3_add(f3_add(cUL, f3_muls(cdx, (float) px + u[0])), f3_muls(cdy, (float) py + u[1]));
        v3 ro = cam_eye, rd = f3_normalize(f3_sub(pixel_pos, cam_eye));
        float eta = 1.0f;
        v3 thr = mk3(1.0f, 1.0f, 1.0f);
        for (int n_delta = 0;; ++n_delta) {
            Hit hit = gpu_find_closest_hit(s, ro, rd);
            n_closest++;
            if (!hit.hit) break;
            v3 wo = f3_muls(rd, -1.0f);
            if (hit.is_light) {
                v3 contrib = f3_mul(thr, hit.mtl.base_color);
                if (is_valid_color(contrib)) st3(image + 3 * (size_t) idx, clamp_radiance(contrib, 15.0f));
                break;
            }
            v3 wi = mk3(0, 0, 0), bsdf_val = mk3(0, 0, 0);
            float pdf_omega = 0.0f, new_eta = 0.0f; int is_delta = 0;
            float ub[4];
            philox_unit4(key, (uint32_t) idx, sg, 1u + (uint32_t) n_delta, 2u, ub);
            bsdf_sample(&hit.mtl, wo, hit.normal, ub[0], ub[1], ub[2], eta, &wi, &bsdf_val, &pdf_omega, &is_delta, &new_eta);
            if (!is_delta) {
                OrHitPoint *h = &hp[n_hp++];
                h->pos = hit.pos; h->normal = hit.normal; h->wo = wo; h->thr = thr; h->mtl = hit.mtl; h->pixel = idx;
                h->gx = (int) floorf((hit.pos.x - min_bound.x) * inv_cell); h->gy = (int) floorf((hit.pos.y - min_bound.y) * inv_cell); h->gz = (int) floorf((hit.pos.z - min_bound.z) * inv_cell);
                h->key = cell_key(h->gx, h->gy, h->gz);
                break;
            }
            if (pdf_omega <= 0.0f) break;
            thr = f3_mul(thr, bsdf_val);
            rd = wi; eta = new_eta;
            ro = f3_add(hit.pos, f3_muls(hit.normal, f3_dot(wi, hit.normal) < 0.0f ? -OR_EPSILON : OR_EPSILON));
            if (!is_valid_color(thr)) break;
            if (max_delta > 0 && n_delta + 1 >= max_delta) { n_cap++; break; }
        }
    }
    /* 2. cell order (the reference's hash grid, :27-58, visits the same hit points: every one within `radius` lies in the 27 cells) */
    qsort(hp, n_hp, sizeof(OrHitPoint), cmp_hp);
    double *flux = (double *) calloc(3 * (size_t) (n_hp > 0 ? n_hp : 1), sizeof(double));
    /* 3. ppm_photon_trace, :158-287 */
    const int total_photons = num_lights * spl;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 64) reduction(+ : n_closest, n_cap, n_gather)
    for (int idx = 0; idx < total_photons; ++idx) {
        const Light *light = &s->lights[idx % num_lights];
        float u[4];
        philox_unit4(key, (uint32_t) idx, 0u, 0u, 5u, u);
        v3 w = f3_normalize(light->dir), uu, vv;
        if (fabsf(w.x) > 0.9f) uu = mk3(0.0f, 1.0f, 0.0f); else uu = mk3(1.0f, 0.0f, 0.0f);
        vv = f3_normalize(f3_cross(w, uu));
        uu = f3_normalize(f3_cross(vv, w));
        v3 ro, rd;
        if (light->is_parallel) {
            rd = w;
            v3 center = f3_muls(f3_add(min_bound, max_bound), 0.5f);
            float scene_radius = f3_length(f3_sub(max_bound, min_bound)) * 0.5f, plane = scene_radius * 2.0f;
            ro = f3_add(f3_add(f3_sub(center, f3_muls(rd, scene_radius * 2.0f)), f3_muls(uu, (u[0] - 0.5f) * plane)), f3_muls(vv, (u[1] - 0.5f) * plane));
        } else {
            float theta = acosf(1.0f - u[0] * (1.0f - cosf(light->cutoff))), phi = 2.0f * OR_PI * u[1];
            v3 ld = mk3(sinf(theta) * cosf(phi), sinf(theta) * sinf(phi), cosf(theta));
            rd = f3_normalize(f3_add(f3_add(f3_muls(uu, ld.x), f3_muls(vv, ld.y)), f3_muls(w, ld.z)));
            ro = f3_add(light->pos, f3_muls(rd, light->ball_r));
        }
        v3 photon_flux = f3_muls(light->illum, (float) num_lights / fmaxf((float) spl, 1.0f));   /* :209 */
        float eta = 1.0f;
        int n_delta = 0;
        for (int depth = 0; depth < light_depth; depth++) {
            Hit hit = gpu_find_closest_hit(s, ro, rd);
            n_closest++;
            if (!hit.hit || hit.is_light) break;
            v3 wi_light = f3_muls(rd, -1.0f);
            if (gathers_light(&hit.mtl)) {
                int gx = (int) floorf((hit.pos.x - min_bound.x) * inv_cell), gy = (int) floorf((hit.pos.y - min_bound.y) * inv_cell), gz = (int) floorf((hit.pos.z - min_bound.z) * inv_cell);
                for (int z = -1; z <= 1; z++) for (int y = -1; y <= 1; y++) for (int x = -1; x <= 1; x++) {
                    uint64_t k = cell_key(gx + x, gy + y, gz + z);
                    int lo = 0, hi = n_hp;
                    while (lo < hi) { int mid = (lo + hi) / 2; if (hp[mid].key < k) lo = mid + 1; else hi = mid; }
                    for (int j = lo; j < n_hp && hp[j].key == k; ++j) {
                        const OrHitPoint *h = &hp[j];
                        if (!(f3_dot(h->normal, hit.normal) > 0.01f)) continue;
                        v3 dd = f3_sub(h->pos, hit.pos);
                        if (!(f3_dot(dd, dd) < radius * radius)) continue;
                        v3 brdf = bsdf_evaluate(&h->mtl, h->wo, wi_light, h->normal);
                        if (is_valid_color(brdf)) {
                            v3 e = f3_mul(f3_mul(photon_flux, brdf), h->thr);
#pragma omp atomic
                            flux[3 * j] += e.x;
#pragma omp atomic
                            flux[3 * j + 1] += e.y;
#pragma omp atomic
                            flux[3 * j + 2] += e.z;
                            n_gather++;
                        }
                    }
                }
            }
            v3 wi = mk3(0, 0, 0), bsdf_val = mk3(0, 0, 0);
            float pdf_omega = 0.0f, new_eta = 0.0f; int is_delta = 0;
            float ub[4];
            philox_unit4(key, (uint32_t) idx, 0u, 1u + (uint32_t) (depth + n_delta), 5u, ub);
            bsdf_sample(&hit.mtl, wi_light, hit.normal, ub[0], ub[1], ub[2], eta, &wi, &bsdf_val, &pdf_omega, &is_delta, &new_eta);
            if (pdf_omega <= 0.0f) break;
            float cos_wi = fabsf(f3_dot(hit.normal, wi));
            if (is_delta) {
                photon_flux = f3_mul(photon_flux, bsdf_val);
                ++n_delta;
                if (max_delta > 0 && n_delta >= max_delta) { n_cap++; break; }
                depth--;
            } else photon_flux = f3_divs(f3_muls(f3_mul(photon_flux, bsdf_val), cos_wi), pdf_omega);
            if (!is_valid_color(photon_flux)) break;
            rd = wi; eta = new_eta;
            ro = f3_add(hit.pos, f3_muls(hit.normal, f3_dot(wi, hit.normal) < 0.0f ? -OR_EPSILON : OR_EPSILON));
        }
    }
    /* 4. ppm_resolve_image, :293-315 */
    for (int j = 0; j < n_hp; ++j) {
        float area = OR_PI * radius * radius;
        v3 radiance = f3_divs(mk3((float) flux[3 * j], (float) flux[3 * j + 1], (float) flux[3 * j + 2]), fmaxf(area, 1e-6f));
        if (is_valid_color(radiance)) {
            float *px = image + 3 * (size_t) hp[j].pixel;
            st3(px, f3_add(ld3(px), clamp_radiance(radiance, 15.0f)));
        }
    }
    free(flux); free(hp);
    if (stats) {
        stats->eye_samples = (uint64_t) npix; stats->closest_rays = n_closest; stats->path_samples = n_gather; stats->delta_cap_hits = n_cap;
        stats->shadow_rays = (uint64_t) n_hp; stats->seconds = now_s() - t0;
    }
}
