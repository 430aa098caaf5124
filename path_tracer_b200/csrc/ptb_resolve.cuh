// Film resolve: accumulator -> display colour.
// Reference behaviour: src/integrator/resolve.glsl:60-130 (normalise by alpha,
// XYZ -> linear sRGB, Brightness, one of four tone-mapping operators) followed by
// the sRGB OETF + UNORM8 conversion the reference gets from its B8G8R8A8_SRGB
// swapchain (src/core/vulkan.cpp:1407).  The reference stretches the film over
// the window with a bilinear sampler; the film here is resolved 1:1.
#pragma once
#include "ptb_scene.cuh"

namespace ptb {

PTB_HD float luminance(f3 c) { return dot(c, mk3(0.2126f, 0.7152f, 0.0722f)); }

// resolve.glsl:65-72 — no zero-luminance guard in the reference (0/0 = NaN on black), kept.
PTB_HD f3 tonemap_reinhard(f3 c, float white) {
    float old_l = luminance(c);
    float n = old_l * (1.0f + (old_l / (white * white)));
    float new_l = n / (1.0f + old_l);
    return (c * new_l) / old_l;
}
// resolve.glsl:74-88
PTB_HD f3 hable_partial(f3 x) {
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    f3 num = x * (A * x + mk3(C * B)) + mk3(D * E);
    f3 den = x * (A * x + mk3(B)) + mk3(D * F);
    return (num / den) - mk3(E / F);
}
PTB_HD f3 tonemap_hable(f3 c) {
    f3 cur = hable_partial(c * 2.0f);
    f3 white_scale = mk3(1.0f) / hable_partial(mk3(11.2f));
    return cur * white_scale;
}
// resolve.glsl:90-110 — Hill's ACES fit, column-major literals.
PTB_HD f3 tonemap_aces(f3 c) {
    const f3 i0 = mk3(0.59719f, 0.07600f, 0.02840f), i1 = mk3(0.35458f, 0.90834f, 0.13383f), i2 = mk3(0.04823f, 0.01566f, 0.83777f);
    const f3 o0 = mk3(1.60475f, -0.10208f, -0.00327f), o1 = mk3(-0.53108f, 1.10813f, -0.07276f), o2 = mk3(-0.07367f, -0.00605f, 1.07602f);
    f3 v = (i0 * c.x + i1 * c.y) + i2 * c.z;
    f3 a = v * (v + mk3(0.0245786f)) - mk3(0.000090537f);
    f3 b = v * (0.983729f * v + mk3(0.4329510f)) + mk3(0.238081f);
    f3 r = a / b;
    return (o0 * r.x + o1 * r.y) + o2 * r.z;
}
// resolve.glsl:112-130
PTB_HD f3 resolve_pixel(float4 value, float brightness, u32 mode, float white) {
    f3 c = mk3(0, 0, 0);
    if (value.w > 0.0f) c = xyz_to_srgb((brightness * mk3(value.x, value.y, value.z)) / value.w);
    if (mode == PTB_TONE_MAPPING_MODE_CLAMP) c = mk3(gclamp(c.x, 0.0f, 1.0f), gclamp(c.y, 0.0f, 1.0f), gclamp(c.z, 0.0f, 1.0f));
    if (mode == PTB_TONE_MAPPING_MODE_REINHARD) c = tonemap_reinhard(c, white);
    if (mode == PTB_TONE_MAPPING_MODE_HABLE) c = tonemap_hable(c);
    if (mode == PTB_TONE_MAPPING_MODE_ACES) c = tonemap_aces(c);
    return c;
}
// Vulkan "sRGB EOTF^-1" then UNORM8 (round to nearest); NaN and negatives map to 0.
PTB_HD u32 linear_to_srgb8(float l) {
    if (!(l > 0.0f)) l = 0.0f;
    if (l > 1.0f) l = 1.0f;
    float e = l <= 0.0031308f ? l * 12.92f : 1.055f * powf(l, 1.0f / 2.4f) - 0.055f;
    return (u32)(int)(e * 255.0f + 0.5f);
}

}  // namespace ptb
