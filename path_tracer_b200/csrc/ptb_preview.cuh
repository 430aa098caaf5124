// Preview visualiser: the reference's debug view as a CUDA kernel body.
// Follows src/application/preview_render.glsl:96-178 — one camera ray per pixel, closest hit,
// colour by render mode — plus the helpers only the preview uses: the 16-sample D65 observation
// (src/core/spectrum.glsl.inc:159-164,197-215) and MaterialBaseColor (src/scene/scene.glsl.inc:696-701).
// No graphics API: the image goes to a device buffer.
#pragma once
#include "ptb_bsdf.cuh"
#include "ptb_wide.cuh"

namespace ptb {

#define PTB_PREVIEW_BASE_COLOR 0u
#define PTB_PREVIEW_BASE_COLOR_SHADED 1u
#define PTB_PREVIEW_NORMAL 2u
#define PTB_PREVIEW_MATERIAL_INDEX 3u
#define PTB_PREVIEW_PRIMITIVE_INDEX 4u
#define PTB_PREVIEW_MESH_COMPLEXITY 5u
#define PTB_PREVIEW_SCENE_COMPLEXITY 6u

#if defined(__CUDACC__)
__device__ const float k_d65[471] = {
#include "../host/d65_table.inc"
};
__device__ const float k_preview_colors[20][3] = {        // preview_render.glsl:16-38
    {0.902f, 0.098f, 0.294f}, {0.235f, 0.706f, 0.294f}, {1.000f, 0.882f, 0.098f}, {0.263f, 0.388f, 0.847f}, {0.961f, 0.510f, 0.192f},
    {0.569f, 0.118f, 0.706f}, {0.275f, 0.941f, 0.941f}, {0.941f, 0.196f, 0.902f}, {0.737f, 0.965f, 0.047f}, {0.980f, 0.745f, 0.745f},
    {0.000f, 0.502f, 0.502f}, {0.902f, 0.745f, 1.000f}, {0.604f, 0.388f, 0.141f}, {1.000f, 0.980f, 0.784f}, {0.502f, 0.000f, 0.000f},
    {0.667f, 1.000f, 0.765f}, {0.502f, 0.502f, 0.000f}, {1.000f, 0.847f, 0.694f}, {0.000f, 0.000f, 0.459f}, {0.502f, 0.502f, 0.502f},
};

// spectrum.glsl.inc:159-164
__device__ __forceinline__ float illuminant_d65(float nl) {
    const float offset = nl * 470.0f;
    int index = (int)offset;
    index = index < 0 ? 0 : (index > 469 ? 469 : index);
    return gmix(k_d65[index], k_d65[index + 1], offset - (float)index);
}
// spectrum.glsl.inc:197-215: XYZ of a parametric spectrum lit by D65, 16 samples
__device__ __forceinline__ f3 observe_under_d65(f4 beta_and_intensity) {
    const float delta = (PTB_LAMBDA_MAX - PTB_LAMBDA_MIN) / 16.0f;
    f3 color = mk3(0.0f);
    for (int i = 0; i < 16; i++) {
        const float nl = (float)i / 15.0f;
        const float d65 = illuminant_d65(nl) / 10566.864005f;
        const float lambda = gmix(PTB_LAMBDA_MIN, PTB_LAMBDA_MAX, nl);
        const float sp = beta_and_intensity.w * parametric_spectrum(xyz(beta_and_intensity), lambda);
        color = color + (standard_observer(lambda) * (sp * d65)) * delta;
    }
    return color;
}
// scene.glsl.inc:254-272 (MaterialColor / MaterialTexturableColor) and :696-701 (MaterialBaseColor)
__device__ __forceinline__ f3 material_color(const DeviceScene& s, u32 slot, u32 attr) {
    return observe_under_d65(mk4(material_f3(s, slot, attr), 1.0f));
}
__device__ __forceinline__ f3 material_texturable_color(const DeviceScene& s, u32 slot, u32 attr, f2 uv) {
    f3 color = material_color(s, slot, attr);
    const u32 tex = material_u32(s, slot, attr + 3);
    if (tex != 0xFFFFFFFFu) color = color * observe_under_d65(mk4(xyz(sample_texture(s, tex, uv)), 1.0f));
    return color;
}
__device__ __forceinline__ f3 material_base_color(const DeviceScene& s, u32 material, f2 uv) {
    const u32 type = material_type(s, material);
    if (type == 0u) return material_texturable_color(s, material, PTB_DIFFUSE_BASE, uv);
    if (type == 1u) return material_texturable_color(s, material, PTB_METAL_BASE, uv);
    if (type == 2u) return material_color(s, material, PTB_TRANS_TRANSMISSION);
    return mk3(0.0f);
}

struct PreviewConsts {             // preview_render.glsl:48-60
    float to[16];                  // CameraTransform.To (column-major)
    u32 mode;
    float brightness;
    u32 selected_shape;
    u32 width, height, mouse_x, mouse_y;
};

// One pixel of the preview image; *shape receives Hit.ShapeIndex.  MODE: 0/1 BVH2 faithful/canonical, 2 wide.
template <int MODE>
__device__ __forceinline__ f3 preview_pixel(const DeviceScene& s, const PreviewConsts& p, u32 px, u32 py, u32* shape) {
    const f2 screen = mk2(((float)px + 0.5f) / (float)p.width, ((float)py + 0.5f) / (float)p.height);
    const float aspect = (float)p.width / (float)p.height;
    Ray ray;
    ray.origin = mat_mul_point(p.to, mk3(0.0f, 0.0f, 0.0f));
    ray.velocity = mat_mul_vector(p.to, normalize(mk3((screen.x - 0.5f) * aspect, 0.5f - screen.y, -1.0f)));
    ray.duration = PTB_HIT_TIME_LIMIT;

    Complexity cx;
    cx.scene = cx.mesh = 0u;
    HitId h;
    if (MODE == 0) h = closest_hit_bvh2<false>(s, ray, &cx);
    else if (MODE == 1) h = closest_hit_bvh2<true>(s, ray, &cx);
    else h = closest_hit_wide(s, ray, &cx);
    const bool hit = h.shape != PTB_SHAPE_NONE;
    Surface sf;
    if (hit) sf = finish_hit(s, h);

    f3 color = mk3(0.0f);
    switch (p.mode) {
        case PTB_PREVIEW_BASE_COLOR:
        case PTB_PREVIEW_BASE_COLOR_SHADED:
            if (!hit) {
                f4 spectrum = mk4(0, 0, 100, 1);
                if (s.globals.SkyboxTextureIndex != 0xFFFFFFFFu) {
                    const float phi = atan2f(ray.velocity.y, ray.velocity.x);
                    const float theta = asinf(ray.velocity.z);
                    spectrum = sample_texture(s, s.globals.SkyboxTextureIndex, mk2(0.5f + phi / PTB_TAU, 0.5f + theta / PTB_PI));
                }
                color = xyz_to_srgb(observe_under_d65(spectrum));
            } else {
                color = xyz_to_srgb(material_base_color(s, sf.material, sf.uv));
                if (p.mode == PTB_PREVIEW_BASE_COLOR_SHADED) color = color * dot(sf.normal, -ray.velocity);
            }
            break;
        case PTB_PREVIEW_NORMAL:
            color = hit ? 0.5f * (sf.normal + mk3(1.0f)) : 0.5f * (mk3(1.0f) - ray.velocity);
            break;
        case PTB_PREVIEW_MATERIAL_INDEX:
            if (hit) { const float* c = k_preview_colors[sf.material % 20u]; color = mk3(c[0], c[1], c[2]); }
            break;
        case PTB_PREVIEW_PRIMITIVE_INDEX:
            if (hit) { const float* c = k_preview_colors[(h.shape_type == PTB_SHAPE_MESH ? h.prim : 0u) % 20u]; color = mk3(c[0], c[1], c[2]); }
            break;
        case PTB_PREVIEW_MESH_COMPLEXITY: color = (mk3(0, 1, 0) * (float)cx.mesh) / 256.0f; break;
        case PTB_PREVIEW_SCENE_COMPLEXITY: color = (mk3(0, 1, 0) * (float)cx.scene) / 256.0f; break;
        default: break;
    }
    if (h.shape == p.selected_shape) color = color * mk3(1.0f, 0.5f, 0.5f);
    color = color * p.brightness;
    *shape = h.shape;
    return color;
}
#endif

}  // namespace ptb
