// The three active spectral BSDFs and the material dispatch.
//
// Reference behaviour: src/scene/basic_diffuse.glsl.inc, basic_metal.glsl.inc,
// basic_translucent.glsl.inc and the dispatch macros of
// src/scene/scene.glsl.inc:687-762.  Direction naming follows the material
// files: `wi` is the given direction, `wo` the sampled one (the integrator's
// naming is the inverse).  Type 3 (OpenPBR) and unknown types
// fall through the dispatch exactly like the reference (include disabled at
// scene.glsl.inc:685): such a surface ends the path.
#pragma once
#include "ptb_scene.cuh"

namespace ptb {

struct BsdfParams { u32 material; f2 uv; f4 lambda; f4 exterior_ior; };

struct Medium { u32 priority; f4 ior, absorption, scattering; float anisotropy; };

// attribute word offsets (tables at the top of each material's .glsl.inc)
#define PTB_DIFFUSE_BASE 1u
#define PTB_METAL_BASE 1u
#define PTB_METAL_SPECULAR 5u
#define PTB_METAL_ROUGHNESS 9u
#define PTB_METAL_ANISOTROPY 11u
#define PTB_TRANS_IOR 1u
#define PTB_TRANS_ABBE 2u
#define PTB_TRANS_ROUGHNESS 3u
#define PTB_TRANS_ANISOTROPY 5u
#define PTB_TRANS_TRANSMISSION 7u
#define PTB_TRANS_DEPTH 10u
#define PTB_TRANS_SCATTERING 11u
#define PTB_TRANS_SCATTER_ANISOTROPY 14u

// ---- BasicDiffuse (basic_diffuse.glsl.inc:19-50) ----
// The pdf is wi.z/pi of the GIVEN direction (reference quirk Q7), kept as is.
PTB_HDN bool diffuse_eval(const DeviceScene& s, const BsdfParams& p, f3 wi, f4& throughput, f4& pdf) {
    f4 reflectance = material_reflectance(s, p.material, PTB_DIFFUSE_BASE, p.lambda, p.uv);
    pdf = mk4(wi.z / PTB_PI);
    throughput = pdf * reflectance;
    return true;
}
PTB_HDN bool diffuse_sample(const DeviceScene& s, const BsdfParams& p, f3 wi, Rng& g, f3& wo, f4& throughput, f4& pdf) {
    f3 d = random_direction(g);
    wo = safe_normalize(d + mk3(0, 0, 1));
    return diffuse_eval(s, p, wi, throughput, pdf);
}

// ---- BasicMetal (basic_metal.glsl.inc) ----
struct MetalParams { f4 base, specular; f2 alpha; bool rough; };
PTB_HD MetalParams metal_params(const DeviceScene& s, const BsdfParams& p) {      // :6-28
    MetalParams m;
    m.base = material_reflectance(s, p.material, PTB_METAL_BASE, p.lambda, p.uv);
    m.specular = material_reflectance(s, p.material, PTB_METAL_SPECULAR, p.lambda, p.uv);
    float r = material_value(s, p.material, PTB_METAL_ROUGHNESS, p.uv);
    float a = material_value(s, p.material, PTB_METAL_ANISOTROPY, p.uv);
    m.alpha = ggx_alpha(r, a);
    m.rough = m.alpha.x * m.alpha.y > PTB_EPSILON;
    return m;
}
PTB_HDN bool metal_eval(const DeviceScene& s, const BsdfParams& p, f3 wi, f3 wo, f4& throughput, f4& pdf) {   // :46-84
    MetalParams m = metal_params(s, p);
    if (wi.z <= 0.0f || wo.z <= 0.0f || !m.rough) return false;
    f3 half = safe_normalize(wi + wo);
    float gm = ggx_g1(wi, m.alpha);
    float d = ggx_d(half, m.alpha);
    pdf = mk4((gm * d) / (4.0f * wi.z));
    float gs = ggx_g1(wo, m.alpha);
    f4 f = schlick_metal(m.base, m.specular, dot(wi, half));
    throughput = (pdf * gs) * f;
    return true;
}
PTB_HDN bool metal_sample(const DeviceScene& s, const BsdfParams& p, f3 wi, Rng& g, f3& wo, f4& throughput, f4& pdf) {   // :87-141
    MetalParams m = metal_params(s, p);
    if (wi.z <= 0.0f) return false;
    float u1 = g.uniform();
    float u2 = g.uniform();
    f3 n = ggx_visible_normal(wi, m.alpha, u1, u2);
    float c = fminf(dot(n, wi), 1.0f);
    wo = ((2.0f * c) * n) - wi;
    if (wo.z <= 0.0f) return false;
    pdf = mk4(1.0f);
    if (m.rough) {
        float gm = ggx_g1(wi, m.alpha);
        float d = ggx_d(n, m.alpha);
        pdf = pdf * mk4((gm * d) / (4.0f * wi.z));
    }
    float gs = ggx_g1(wo, m.alpha);
    f4 f = schlick_metal(m.base, m.specular, c);
    throughput = (pdf * gs) * f;
    return true;
}

// ---- BasicTranslucent (basic_translucent.glsl.inc) ----
struct TransParams { f4 eta; f2 alpha; bool rough; };
PTB_HD TransParams trans_params(const DeviceScene& s, const BsdfParams& p, f3 wi) {   // :10-48
    TransParams t;
    f4 interior = cauchy_ior(material_f32(s, p.material, PTB_TRANS_IOR), material_f32(s, p.material, PTB_TRANS_ABBE), p.lambda);
    t.eta = wi.z < 0.0f ? interior / p.exterior_ior : p.exterior_ior / interior;
    float r = material_value(s, p.material, PTB_TRANS_ROUGHNESS, p.uv);
    float a = material_value(s, p.material, PTB_TRANS_ANISOTROPY, p.uv);
    t.alpha = ggx_alpha(r, a);
    t.rough = t.alpha.x * t.alpha.y > PTB_EPSILON;
    return t;
}
PTB_HDN void trans_load_medium(const DeviceScene& s, u32 material, f4 lambda, Medium& m) {   // :55-82
    m.ior = cauchy_ior(material_f32(s, material, PTB_TRANS_IOR), material_f32(s, material, PTB_TRANS_ABBE), lambda);
    float depth = material_f32(s, material, PTB_TRANS_DEPTH);
    if (depth > 0.0f) {
        f4 extinction = (-log4(parametric_spectrum(material_f3(s, material, PTB_TRANS_TRANSMISSION), lambda))) / depth;
        f4 scattering = parametric_spectrum(material_f3(s, material, PTB_TRANS_SCATTERING), lambda) / depth;
        m.absorption = max4s(extinction - scattering, 0.0f);
        m.scattering = scattering;
        m.anisotropy = material_f32(s, material, PTB_TRANS_SCATTER_ANISOTROPY);
    } else {
        m.absorption = mk4(0.0f); m.scattering = mk4(0.0f); m.anisotropy = 0.0f;
    }
}
// Per-wavelength refraction pdf terms shared by eval and sample (:128-160, :283-323).
PTB_HD f4 trans_refraction_pdf(f3 wi, f3 wo, const TransParams& t, f3 h1, f4 ci, f4 co, bool first_always) {
    f3 h2 = safe_normalize(wo + wi * t.eta.y);
    f3 h3 = safe_normalize(wo + wi * t.eta.z);
    f3 h4 = safe_normalize(wo + wi * t.eta.w);
    ci.y = dot(wi, h2); ci.z = dot(wi, h3); ci.w = dot(wi, h4);
    co.y = dot(wo, h2); co.z = dot(wo, h3); co.w = dot(wo, h4);
    f4 f = fresnel_dielectric(t.eta, ci, co);
    f4 d = mk4(0.0f);
    if (first_always || ci.x * co.x < 0.0f) d.x = ggx_d(h1, t.alpha);
    if (ci.y * co.y < 0.0f) d.y = ggx_d(h2, t.alpha);
    if (ci.z * co.z < 0.0f) d.z = ggx_d(h3, t.alpha);
    if (ci.w * co.w < 0.0f) d.w = ggx_d(h4, t.alpha);
    float gm = ggx_g1(wi, t.alpha);
    f4 q = ci * t.eta + co;
    f4 jac = abs4(co) / (q * q);
    return (((d * (1.0f - f)) * gm) * jac) * abs4(ci / wi.z);
}
PTB_HDN bool trans_eval(const DeviceScene& s, const BsdfParams& p, f3 wi, f3 wo, f4& throughput, f4& pdf) {   // :90-169
    TransParams t = trans_params(s, p, wi);
    if (!t.rough) { pdf = mk4(0.0f); throughput = mk4(0.0f); return true; }
    if (wi.z * wo.z > 0.0f) {
        float gm = ggx_g1(wi, t.alpha);
        f3 half = safe_normalize(wo + wi);
        float c = dot(half, wi);
        f4 f = fresnel_dielectric(t.eta, mk4(c));
        float d = ggx_d(half, t.alpha);
        pdf = ((f * gm) * d) / (4.0f * wi.z);
    } else {
        f3 h1 = safe_normalize(wo + wi * t.eta.x);
        f4 ci = mk4(dot(wi, h1), 0, 0, 0), co = mk4(dot(wo, h1), 0, 0, 0);
        pdf = trans_refraction_pdf(wi, wo, t, h1, ci, co, false);
    }
    float gs = ggx_g1(wo, t.alpha);
    throughput = pdf * gs;
    return true;
}
PTB_HDN bool trans_sample(const DeviceScene& s, const BsdfParams& p, f3 wi, Rng& g, f3& wo, f4& throughput, f4& pdf) {   // :172-339
    TransParams t = trans_params(s, p, wi);
    float u1 = g.uniform();
    float u2 = g.uniform();
    f3 n = ggx_visible_normal(wi * gsign(wi.z), t.alpha, u1, u2);
    float c = gclamp(dot(n, wi), -1.0f, 1.0f);
    float cr = cos_refracted(t.eta.x, c);
    float reflectance = fresnel_dielectric(t.eta.x, c, cr);
    if (g.uniform() < reflectance) {
        wo = ((2.0f * c) * n) - wi;
        if (wo.z * wi.z <= 0.0f) return false;
        pdf = fresnel_dielectric(t.eta, mk4(c));
        if (t.rough) {
            float gm = ggx_g1(wi, t.alpha);
            float d = ggx_d(n, t.alpha);
            pdf = pdf * ((gm * d) / (4.0f * fabsf(wi.z)));
        }
        float gs = ggx_g1(wo, t.alpha);
        throughput = pdf * gs;
        return true;
    }
    wo = ((cr + t.eta.x * c) * n) - (t.eta.x * wi);
    if (wo.z * wi.z >= 0.0f) return false;
    if (t.rough) {
        pdf = trans_refraction_pdf(wi, wo, t, n, mk4(c, 0, 0, 0), mk4(cr, 0, 0, 0), true);
    } else {
        pdf = mk4(1.0f - reflectance, 0.0f, 0.0f, 0.0f);      // smooth refraction drops the 3 secondary wavelengths (Q22)
    }
    float gs = ggx_g1(wo, t.alpha);
    throughput = pdf * gs;
    return true;
}

// ---- dispatch (scene.glsl.inc:687-762) ----
PTB_HD u32 material_type(const DeviceScene& s, u32 slot) { return material_u32(s, slot, 0); }

PTB_HD bool material_load_medium(const DeviceScene& s, u32 material, f4 lambda, Medium& m) {
    if ((s.features & PTB_FEAT_TRANSLUCENT) && material_type(s, material) == 2u) { trans_load_medium(s, material, lambda, m); return true; }
    return false;
}
PTB_HD bool material_has_dirac(const DeviceScene& s, const BsdfParams& p) {
    u32 type = material_type(s, p.material);
    if ((s.features & PTB_FEAT_METAL) && type == 1u) return material_value(s, p.material, PTB_METAL_ROUGHNESS, p.uv) < 1e-3f;
    if ((s.features & PTB_FEAT_TRANSLUCENT) && type == 2u) return material_value(s, p.material, PTB_TRANS_ROUGHNESS, p.uv) < 1e-3f;
    return false;
}
PTB_HD bool material_eval(const DeviceScene& s, const BsdfParams& p, f3 wi, f3 wo, f4& throughput, f4& pdf) {
    u32 type = material_type(s, p.material);
    if ((s.features & PTB_FEAT_DIFFUSE) && type == 0u) return diffuse_eval(s, p, wi, throughput, pdf);
    if ((s.features & PTB_FEAT_METAL) && type == 1u) return metal_eval(s, p, wi, wo, throughput, pdf);
    if ((s.features & PTB_FEAT_TRANSLUCENT) && type == 2u) return trans_eval(s, p, wi, wo, throughput, pdf);
    return false;
}
PTB_HD bool material_sample(const DeviceScene& s, const BsdfParams& p, f3 wi, Rng& g, f3& wo, f4& throughput, f4& pdf) {
    u32 type = material_type(s, p.material);
    if ((s.features & PTB_FEAT_DIFFUSE) && type == 0u) return diffuse_sample(s, p, wi, g, wo, throughput, pdf);
    if ((s.features & PTB_FEAT_METAL) && type == 1u) return metal_sample(s, p, wi, g, wo, throughput, pdf);
    if ((s.features & PTB_FEAT_TRANSLUCENT) && type == 2u) return trans_sample(s, p, wi, g, wo, throughput, pdf);
    return false;
}

}  // namespace ptb
