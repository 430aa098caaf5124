// RNG, direction sampling, microfacet and Fresnel helpers of the integrator.
// Behaviour follows the reference's src/core/common.glsl.inc (line ranges cited
// per function); random draws are sequenced in statements in the order the
// GLSL evaluates them.
#pragma once
#include "ptb_vec.cuh"

namespace ptb {

// common.glsl.inc:4-12 — PI/TAU are the reference's truncated literals.
#define PTB_INFINITY 1e30f
#define PTB_EPSILON 1e-9f
#define PTB_PI 3.141592653f
#define PTB_TAU 6.283185306f
#define PTB_HIT_TIME_LIMIT 1048576.0f
#define PTB_LAMBDA_MIN 360.0f
#define PTB_LAMBDA_MAX 830.0f

// common.glsl.inc:189-203: 32-bit LCG state with the PCG RXS-M-XS output
// permutation; Random0To1() divides by 2^32 in float (can return 1.0).
struct Rng {
    u32 state;
    PTB_HD u32 next() {
        state = state * 747796405u + 2891336453u;
        u32 s = state;
        u32 w = ((s >> ((s >> 28u) + 4u)) ^ s) * 277803737u;
        return (w >> 22u) ^ w;
    }
    PTB_HD float uniform() { return (float)next() / 4294967296.0f; }
};

// common.glsl.inc:113-125
PTB_HD f3 tangent_vector(f3 n) {
    f3 v = fabsf(n.x) < 0.9f ? mk3(1, 0, 0) : mk3(0, 1, 0);
    return normalize(cross(v, n));
}
PTB_HD void coordinate_frame(f3 z, f3& x, f3& y) {
    f3 v = fabsf(z.x) < 0.9f ? mk3(1, 0, 0) : mk3(0, 1, 0);
    x = normalize(cross(v, z));
    y = cross(x, z);
}

// sin and cos of one angle with one argument reduction (the same values sinf / cosf return, CUDA math API).
PTB_HD void sin_cos(float a, float& s, float& c) { sincosf(a, &s, &c); }

// common.glsl.inc:205-218
PTB_HD f2 random_point_on_disk(Rng& g) {
    float r = sqrtf(g.uniform());
    float theta = g.uniform() * PTB_TAU;
    float sn, cs;
    sin_cos(theta, sn, cs);
    return mk2(r * cs, r * sn);
}
PTB_HD f3 random_direction(Rng& g) {
    float z = 2.0f * g.uniform() - 1.0f;
    float r = sqrtf(1.0f - z * z);
    float phi = PTB_TAU * g.uniform();
    float sn, cs;
    sin_cos(phi, sn, cs);
    return mk3(r * cs, r * sn, z);
}

// common.glsl.inc:228-254 — von Mises-Fisher lobe used for sky sampling.
PTB_HD f3 random_vmf(Rng& g, float kappa, f3 mu) {
    float xi = g.uniform();
    float z = 1.0f + (1.0f / kappa) * logf(xi + (1.0f - xi) * expf(-2.0f * kappa));
    float r = sqrtf(1.0f - z * z);
    float phi = g.uniform() * PTB_TAU;
    float sn, cs;
    sin_cos(phi, sn, cs);
    f3 v = mk3(r * cs, r * sn, z);
    f3 mx, my;
    coordinate_frame(mu, mx, my);
    return safe_normalize((v.x * mx + v.y * my) + v.z * mu);
}
PTB_HD float vmf_pdf(float kappa, f3 mu, f3 d) {
    if (kappa < PTB_EPSILON) return 1.0f / (4.0f * PTB_PI);
    float c = kappa / ((2.0f * PTB_PI) * (1.0f - expf(-2.0f * kappa)));
    return c * expf(kappa * (dot(mu, d) - 1.0f));
}

// common.glsl.inc:259-276 — Henyey-Greenstein direction.
PTB_HD f3 sample_hg(float aniso, float u1, float u2) {
    float z;
    if (fabsf(aniso) < 1e-3f) {
        z = 1.0f - 2.0f * u1;
    } else {
        float g = aniso;
        float s = (1.0f - g * g) / ((1.0f + g) - (2.0f * g) * u1);
        z = -((1.0f + g * g) - s * s) / (2.0f * g);
    }
    float r = sqrtf(1.0f - z * z);
    float phi = u2 * PTB_TAU;
    float sn, cs;
    sin_cos(phi, sn, cs);
    return mk3(r * cs, r * sn, z);
}

// common.glsl.inc:281-354 — GGX.
PTB_HD f2 ggx_alpha(float roughness, float anisotropy) {
    float s = 1.0f - anisotropy;
    float ax = (roughness * roughness) * sqrtf(2.0f / (1.0f + s * s));
    return mk2(ax, s * ax);
}
PTB_HD float ggx_g1(f3 d, f2 alpha) {
    f3 d2 = d * d;
    if (d2.z < PTB_EPSILON) return 0.0f;
    f2 a2 = alpha * alpha;
    float k = dot(a2, mk2(d2.x, d2.y)) / d2.z;
    return 2.0f / (1.0f + sqrtf(1.0f + k));
}
PTB_HD f3 ggx_visible_normal(f3 d, f2 alpha, float u1, float u2) {
    f3 vz = safe_normalize(mk3(alpha.x * d.x, alpha.y * d.y, d.z));
    float l2 = vz.x * vz.x + vz.y * vz.y;
    f3 vx = l2 > 0.0f ? mk3(-vz.y, vz.x, 0.0f) / sqrtf(l2) : mk3(1, 0, 0);
    f3 vy = cross(vz, vx);
    float r = sqrtf(u1);
    float phi = PTB_TAU * u2;
    float s = 0.5f * (1.0f + vz.z);
    float sn, cs;
    sin_cos(phi, sn, cs);
    float tx = r * cs;
    float ty = (1.0f - s) * sqrtf(1.0f - tx * tx) + (s * r) * sn;
    float tz = sqrtf(fmaxf(0.0f, (1.0f - tx * tx) - ty * ty));
    f3 n = (tx * vx + ty * vy) + tz * vz;
    return safe_normalize(mk3(alpha.x * n.x, alpha.y * n.y, fmaxf(0.0f, n.z)));
}
PTB_HD float ggx_d(f3 n, f2 alpha) {
    f2 a = mk2(1.0f / alpha.x, 1.0f / alpha.y);
    float b = dot(n * n, mk3(a.x * a.x, a.y * a.y, 1.0f));
    return 1.0f / ((((PTB_PI * alpha.x) * alpha.y) * b) * b);
}

// common.glsl.inc:360-371 — Cauchy dispersion from (IOR at d-line, Abbe number).
PTB_HD f4 cauchy_ior(float base_ior, float abbe, f4 lambda) {
    const float lc = 656.3f, ld = 587.6f, lf = 486.1f;
    float b = (base_ior - 1.0f) / (abbe * (1.0f / (lf * lf) - 1.0f / (lc * lc)));
    float a = base_ior - b / (ld * ld);
    return a + mk4(b) / (lambda * lambda);
}

// common.glsl.inc:379-420 — refraction cosine and dielectric Fresnel.
PTB_HD float cos_refracted(float eta, float c) {
    float c2 = 1.0f - (eta * eta) * (1.0f - c * c);
    return -gsign(c) * sqrtf(fmaxf(c2, 0.0f));
}
PTB_HD f4 cos_refracted(f4 eta, f4 c) {
    f4 c2 = 1.0f - (eta * eta) * (1.0f - c * c);
    return (-gsign(c)) * sqrt4(max4s(c2, 0.0f));
}
PTB_HD float fresnel_dielectric(float eta, float c1, float c2) {
    float ks = eta * c1;
    float rs = (ks + c2) / (ks - c2);
    float kp = eta * c2;
    float rp = (kp + c1) / (kp - c1);
    return 0.5f * (rs * rs + rp * rp);
}
PTB_HD f4 fresnel_dielectric(f4 eta, f4 c1, f4 c2) {
    f4 ks = eta * c1;
    f4 rs = (ks + c2) / (ks - c2);
    f4 kp = eta * c2;
    f4 rp = (kp + c1) / (kp - c1);
    return 0.5f * (rs * rs + rp * rp);
}
PTB_HD f4 fresnel_dielectric(f4 eta, f4 c1) { return fresnel_dielectric(eta, c1, cos_refracted(eta, c1)); }

// common.glsl.inc:425-436 — F82-tint Schlick for conductors.
PTB_HD f4 schlick_metal(f4 base, f4 specular, float c) {
    const float cmax = 1.0f / 7.0f;
    f4 fs = base + (1.0f - base) * powf(1.0f - c, 5.0f);
    f4 fsmax = base + (1.0f - base) * powf(1.0f - cmax, 5.0f);
    f4 fmax = specular * fsmax;
    float den = cmax * powf(1.0f - cmax, 6.0f);
    float num = c * powf(1.0f - c, 6.0f);
    return fs - (num / den) * (fsmax - fmax);
}

}  // namespace ptb
