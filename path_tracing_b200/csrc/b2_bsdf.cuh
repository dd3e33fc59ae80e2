// b2_bsdf.cuh -- BSDF evaluate / pdf / sample of the reference material model, re-authored for the device.
// Follows include/geometric.cuh:119-221 (frame, Fresnel, GGX) and :419-562 (bsdf_evaluate / bsdf_pdf /
// bsdf_sample), including its quirks (SURVEY.md 8a rows M,N,O): F == 1 for eta == 0 non-metals through
// the TIR branch of FrDielectric, the 1e-7 regularisers, delta lobes returning f without cos/pdf.
// FMA contraction is allowed here: parity of shading values is to ~1e-6 relative, not bitwise.
#pragma once
#include "b2_common.cuh"

struct Mat { float3 base; float roughness, metallic, eta; };   // CudaMaterial, geometric.cuh:21-27
// The reference's tr_D (geometric.cuh:178-185) squares tan^2 once too often: e = cos^4 (alpha^2 + tan^4) instead of
// cos^4 (alpha^2 + tan^2)^2, so D is not the GGX distribution its own sample_vndf draws from and bsdf_pdf is not the density
// of bsdf_sample.  Parity mode reproduces that bug for bug.  The strategies that are not part of the reference estimator
// (b2_strategies.cuh) need a sampler and a pdf that agree, or the two estimators they validate against each other carry
// different biases: they mark their materials by setting the SIGN BIT of roughness (a material travels through the
// eye-vertex / light-vertex records as six floats, so the flag costs no storage), and D is then the normalised GGX.
B2_HD void mat_mark_fix(Mat &m) { m.roughness = u2f(f2u(m.roughness) | 0x80000000u); }
B2_HD bool mat_fix(const Mat &m) { return (f2u(m.roughness) & 0x80000000u) != 0u; }
B2_HD float mat_rough(const Mat &m) { return fabsf(m.roughness); }

struct Frame { float3 T, B, N; };
B2_HD Frame make_frame(float3 N) {                              // build_local_frame, :119-123
    Frame f; f.N = N;
    if (fabsf(N.z) < 0.999f) f.T = normalize3(cross3(mk3(0, 0, 1), N));
    else f.T = normalize3(cross3(mk3(0, 1, 0), N));
    f.B = cross3(N, f.T);
    return f;
}
B2_HD float3 to_local(const Frame &f, float3 v) { return mk3(dot3(v, f.T), dot3(v, f.B), dot3(v, f.N)); }
B2_HD float3 to_world(const Frame &f, float3 v) {
    return mk3(f.T.x * v.x + f.B.x * v.y + f.N.x * v.z, f.T.y * v.x + f.B.y * v.y + f.N.y * v.z, f.T.z * v.x + f.B.z * v.y + f.N.z * v.z);
}

B2_HD float fr_dielectric(float cosI, float etaI, float etaT) {  // :145-160
    cosI = fmaxf(-1.0f, fminf(1.0f, cosI));
    if (!(cosI > 0.0f)) { float t = etaI; etaI = etaT; etaT = t; cosI = fabsf(cosI); }
    float sinI = fast_sqrt(fmaxf(0.0f, 1.0f - cosI * cosI));
    float sinT = fast_div(etaI, etaT) * sinI;                    // etaT == 0 (opaque non-metal): inf or nan -> see below
    if (sinT >= 1.0f) return 1.0f;
    float cosT = fast_sqrt(fmaxf(0.0f, 1.0f - sinT * sinT));
    float Rparl = fast_div((etaT * cosI) - (etaI * cosT), (etaT * cosI) + (etaI * cosT));
    float Rperp = fast_div((etaI * cosI) - (etaT * cosT), (etaI * cosI) + (etaT * cosT));
    return (Rparl * Rparl + Rperp * Rperp) * 0.5f;
}
B2_HD float3 fr_schlick(float cosI, float3 R0) {                 // :163-167
    float c = fmaxf(0.0f, 1.0f - cosI);
    float c5 = c * c * c * c * c;
    return R0 + (mk3(1.0f, 1.0f, 1.0f) - R0) * c5;
}
B2_HD float roughness_to_alpha(float r) { float x = fmaxf(r, 1e-3f); return x * x; }   // :173
B2_HD float tr_D(float3 wh, float alpha, bool fix) {             // :178-185
    float cos2 = wh.z * wh.z;
    float tan2 = fast_div(fmaxf(0.0f, 1.0f - cos2), cos2 + 1e-7f);
    if (isinf(tan2)) return 0.0f;
    float e = (cos2 * cos2) * (fix ? (alpha * alpha + tan2) * (alpha * alpha + tan2) : alpha * alpha + tan2 * tan2);
    if (e < 1e-12f) return 0.0f;
    return fast_div(alpha * alpha, B2_PI * e);
}
B2_HD float tr_lambda(float3 w, float alpha) {                   // :187-192
    float t = fabsf(fast_div(fast_sqrt(fmaxf(0.0f, 1.0f - w.z * w.z)), w.z + 1e-7f));
    if (isinf(t)) return 0.0f;
    float a2t2 = (alpha * t) * (alpha * t);
    return (-1.0f + fast_sqrt(1.0f + a2t2)) * 0.5f;
}
B2_HD float tr_G(float3 wo, float3 wi, float alpha) { return fast_rcp(1.0f + tr_lambda(wo, alpha) + tr_lambda(wi, alpha)); }
B2_HD float3 sample_vndf(float3 wo, float alpha, float u1, float u2) {  // :200-221
    float3 V = normalize3(mk3(alpha * wo.x, alpha * wo.y, wo.z));
    float3 T1 = (V.z < 0.9999f) ? normalize3(cross3(mk3(0, 0, 1), V)) : mk3(1, 0, 0);
    float3 T2 = cross3(V, T1);
    float r = fast_sqrt(u1);
    float sphi, cphi;
    sincos_2pi(u2, sphi, cphi);
    float t1 = r * cphi;
    float t2 = r * sphi;
    float s = 0.5f * (1.0f + V.z);
    t2 = (1.0f - s) * fast_sqrt(fmaxf(0.0f, 1.0f - t1 * t1)) + s * t2;
    float3 Nh = T1 * t1 + T2 * t2 + V * fast_sqrt(fmaxf(0.0f, 1.0f - t1 * t1 - t2 * t2));
    return normalize3(mk3(alpha * Nh.x, alpha * Nh.y, fmaxf(0.0f, Nh.z)));
}

// local-space cores (callers that evaluate several lobes for one normal reuse the frame)
B2_HD float3 bsdf_eval_local(const Mat &m, float3 wo, float3 wi) {      // :426-455
    if (wo.z == 0.0f || wi.z == 0.0f) return mk3(0, 0, 0);
    if (m.eta > 0.0f && mat_rough(m) < 0.001f) return mk3(0, 0, 0);
    float alpha = roughness_to_alpha(mat_rough(m));
    float3 whv = wo + wi;
    if (length3(whv) < 1e-6f) return mk3(0, 0, 0);
    float3 wh = normalize3(whv);
    if (wh.z < 0.0f) wh = wh * -1.0f;
    float3 diffuse = m.base * ((1.0f / B2_PI) * (1.0f - m.metallic));
    if (wo.z * wi.z < 0.0f) diffuse = mk3(0, 0, 0);
    float D = tr_D(wh, alpha, mat_fix(m));
    float G = tr_G(wo, wi, alpha);
    float3 F;
    if (m.metallic > 0.0f) F = fr_schlick(fabsf(wo.z), m.base);
    else { float fr = fr_dielectric(dot3(wo, wh), 1.0f, m.eta); F = mk3(fr, fr, fr); }
    float3 spec = (F * D * G) / fmaxf(4.0f * fabsf(wo.z) * fabsf(wi.z), 1e-4f);   // operator/ = one reciprocal
    if (wo.z * wi.z > 0.0f) return diffuse + spec;
    return diffuse;
}
B2_HD float bsdf_pdf_local(const Mat &m, float3 wo, float3 wi) {        // :464-483
    if (wo.z * wi.z <= 0.0f) return 0.0f;
    if (m.eta > 0.0f && mat_rough(m) < 0.001f) return 0.0f;
    float alpha = roughness_to_alpha(mat_rough(m));
    float3 whv = wo + wi;
    if (length3(whv) < 1e-6f) return 0.0f;
    float3 wh = normalize3(whv);
    if (wh.z < 0.0f) wh = wh * -1.0f;
    float pdf_diffuse = fabsf(wi.z) * (1.0f / B2_PI);
    float G1 = fast_rcp(1.0f + tr_lambda(wo, alpha));
    float owh = dot3(wo, wh);
    float pdf_wh = fast_div(tr_D(wh, alpha, mat_fix(m)) * G1 * fmaxf(0.0f, owh), fabsf(wo.z));
    float pdf_spec = fast_div(pdf_wh, 4.0f * owh + 1e-7f);
    float sw = m.metallic > 0.0f ? 1.0f : 0.5f;
    return (1.0f - sw) * pdf_diffuse + sw * pdf_spec;
}
// The world-space entry points are deliberately NOT inlined on the device: the connect stage calls each of them twice
// per pair, and inlining every copy made the kernel miss in the instruction cache (ncu: 41 % stall_no_inst).
// B2_MAT_LOCAL_COPY: the callers in the connection stage (b2_bdpt.cuh) hand these a 24-byte local copy of the material, not a
// reference into the 80-byte vertex record, which would force the whole record into local memory.  With 28 KB of L1 per SM
// the big records were the faster form (a hand-placed spill of cold data: C3 1198 vs 1221 ms); since 0.2.3 leaves L1 60 KB
// the 51 GB of local traffic per launch they cause costs more than it saves: C3 1042 -> 1004 ms, C1 250 -> 228 ms
// (profiles/r02_connect_summary.md).  B2_MAT_BY_VALUE: Mat in registers through the call (measured slower both times).
#ifndef B2_MAT_LOCAL_COPY
#define B2_MAT_LOCAL_COPY 1
#endif
#ifndef B2_MAT_BY_VALUE
#define B2_MAT_BY_VALUE 0
#endif
#if B2_MAT_BY_VALUE
typedef Mat MatArg;
#else
typedef const Mat &MatArg;
#endif
B2_NOINLINE float3 bsdf_evaluate(MatArg m, float3 wo_w, float3 wi_w, float3 N) {
    Frame f = make_frame(N);
    return bsdf_eval_local(m, to_local(f, wo_w), to_local(f, wi_w));
}
B2_NOINLINE float bsdf_pdf(MatArg m, float3 wo_w, float3 wi_w, float3 N) {
    Frame f = make_frame(N);
    return bsdf_pdf_local(m, to_local(f, wo_w), to_local(f, wi_w));
}

struct BsdfSample { float3 wi, f; float pdf, new_eta; bool is_delta; };
// bsdf_sample, :486-562.  On the early returns (pdf = 0) wi/f keep the caller-provided values,
// like the by-reference outputs of the reference.
#ifndef B2_SAMPLE_NOINLINE
#define B2_SAMPLE_NOINLINE 0
#endif
#if B2_SAMPLE_NOINLINE
B2_NOINLINE
#else
B2_HD
#endif
void bsdf_sample(const Mat &m, float3 wo_w, float3 N, float u_rr, float u1, float u2, float cur_eta, BsdfSample &s) {
    s.is_delta = false;
    s.new_eta = cur_eta;
    Frame fr = make_frame(N);
    float3 wo = to_local(fr, wo_w);
    float3 wi;
    if (m.eta > 0.0f && mat_rough(m) < 0.001f && m.metallic < 0.01f) {
        s.is_delta = true;
        float F = fr_dielectric(wo.z, cur_eta, m.eta);
        if (u_rr < F) {
            wi = mk3(-wo.x, -wo.y, wo.z);
            s.pdf = F;
            s.f = mk3(F, F, F) / fabsf(wi.z);
        } else {
            float eta = wo.z > 0.0f ? (cur_eta / m.eta) : (m.eta / cur_eta);
            float sin2I = fmaxf(0.0f, 1.0f - wo.z * wo.z);
            float sin2T = eta * eta * sin2I;
            if (sin2T >= 1.0f) { s.pdf = 0.0f; return; }
            float cosT = fast_sqrt(1.0f - sin2T);
            if (wo.z > 0.0f) cosT = -cosT;
            wi = mk3(-eta * wo.x, -eta * wo.y, cosT);
            s.new_eta = (wo.z > 0.0f) ? m.eta : 1.0f;
            s.pdf = 1.0f - F;
            s.f = m.base * (1.0f - F) / fabsf(wi.z);
        }
        s.wi = to_world(fr, wi);
        return;
    }
    if (m.metallic > 0.99f && mat_rough(m) < 0.001f) {
        s.is_delta = true;
        wi = mk3(-wo.x, -wo.y, wo.z);
        s.pdf = 1.0f;
        s.f = fr_schlick(fabsf(wo.z), m.base) / fabsf(wi.z);
        s.wi = to_world(fr, wi);
        return;
    }
    float alpha = roughness_to_alpha(mat_rough(m));
    float sw = m.metallic > 0.0f ? 1.0f : 0.5f;
    if (u_rr < sw) {
        float3 wh = sample_vndf(wo.z > 0 ? wo : wo * -1.0f, alpha, u1, u2);
        if (wo.z < 0.0f) wh = wh * -1.0f;
        float3 I = wo * -1.0f;
        wi = I - wh * 2.0f * dot3(wh, I);              // reflect(), :99
        if (wo.z * wi.z <= 0.0f) { s.pdf = 0.0f; return; }
    } else {
        float r = fast_sqrt(u1);
        float sphi, cphi;
        sincos_2pi(u2, sphi, cphi);
        wi = mk3(r * cphi, r * sphi, fast_sqrt(fmaxf(0.0f, 1.0f - u1)));
        if (wo.z < 0.0f) wi.z *= -1.0f;
    }
    s.wi = to_world(fr, wi);
    // the reference re-derives both from world space (:560-561): keep that (wi_w -> local again)
    float3 wi_l = to_local(fr, s.wi);
    s.pdf = bsdf_pdf_local(m, wo, wi_l);
    s.f = bsdf_eval_local(m, wo, wi_l);
}
