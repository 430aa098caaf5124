// Device view of an uploaded scene + the scene-side device functions of the
// hot path: spectral helpers, atlas texture fetch, skybox, material attribute
// fetch and camera-ray generation.
//
// Reference behaviour: src/core/spectrum.glsl.inc:10-55,169-192 and
// src/scene/scene.glsl.inc:181-302,613-655.  Layout differences from the
// reference are deliberate: packed records are read through
// 16-byte vector loads, textures are plain RGBA32F arrays addressed with manual
// fp32 filtering (no sampler hardware: weights stay in fp32).
#pragma once
#include "ptb_sampling.cuh"
#include "ptb.h"

namespace ptb {

// Everything a kernel needs to see of a scene: raw device pointers to the packed
// arrays exactly as PackSceneData() lays them out (src/scene/scene.hpp:82-173),
// plus the re-flattened wide BVH built at ptb_scene_update().
struct DeviceScene {
    const ptb_packed_texture*      textures;
    const u32*                     material_data;     // 32 words per slot
    const ptb_packed_shape*        shapes;
    const ptb_packed_shape_node*   shape_nodes;
    const ptb_packed_mesh_face*    mesh_faces;
    const ptb_packed_mesh_vertex*  mesh_vertices;
    const ptb_packed_mesh_node*    mesh_nodes;
    const uint4*                   face_attrs;        // 2 x uint4 per face: the three vertices' {packed normal, packed uv} gathered once
                                                      // (k_face_attrs), so that deriving a hit's attributes is one 32-byte fetch by face
                                                      // index instead of face -> 3 indices -> 3 scattered vertex records; null = not built
    const ptb_packed_camera*       cameras;
    const float* const*            atlas_layers;      // device array of device pointers, RGBA32F
    u32 atlas_width, atlas_height;
    u32 shape_count, shape_node_count, mesh_face_count, mesh_node_count, material_word_count, texture_count, camera_count;
    ptb_packed_scene_globals       globals;           // by value: lives in the kernel parameter bank
    // wide BVH (ptb_wide.cuh); null until built
    const uint4*  wide_nodes;        // 80-byte nodes as 5 x uint4
    const float4* wide_tris;         // 3 x float4 per leaf primitive: object-space vertices; .w = face index, shape index, tag (rank<<3 | identity<<2 | kind)
    const u32*    wide_planes;       // {shape index, tag} of unbounded shapes (planes), tested linearly
    u32 wide_plane_count;
    u32 wide_tlas_root;              // 0 when the tree exists, 0xFFFFFFFF when the scene has no bounded primitive
    u32 features;                    // PTB_FEAT_*: what the scene can make scatter do (set at ptb_scene_update)
};

// Scene features.  k_scatter is instantiated for a few masks: the kernel overwrites `features` of its
// by-value scene with the compile-time mask, so every test below folds and the code of absent
// features is never emitted.  The host picks the smallest instantiation that covers the scene.
#define PTB_FEAT_DIFFUSE       1u     // BasicDiffuse materials
#define PTB_FEAT_METAL         2u     // BasicMetal
#define PTB_FEAT_TRANSLUCENT   4u     // BasicTranslucent (dielectric + interior medium)
#define PTB_FEAT_TEXTURES      8u     // any texture in the scene
#define PTB_FEAT_SKY_TEXTURE  16u     // skybox texture set
#define PTB_FEAT_SCENE_SCATTER 32u    // SceneScatterRate > 0
#define PTB_FEAT_ALL          63u

struct Ray { f3 origin; f3 velocity; float duration; };

// ---- spectrum.glsl.inc ----
// :10-34  Wyman-Sloan-Shirley multi-lobe fit of the CIE 1931 observer.
PTB_HDN f3 standard_observer(float l) {
    f3 r;
    {
        float t1 = (l - 442.0f) * (l < 442.0f ? 0.0624f : 0.0374f);
        float t2 = (l - 599.8f) * (l < 599.8f ? 0.0264f : 0.0323f);
        float t3 = (l - 501.1f) * (l < 501.1f ? 0.0490f : 0.0382f);
        r.x = (0.362f * expf((-0.5f * t1) * t1) + 1.056f * expf((-0.5f * t2) * t2)) - 0.065f * expf((-0.5f * t3) * t3);
    }
    {
        float t1 = (l - 568.8f) * (l < 568.8f ? 0.0213f : 0.0247f);
        float t2 = (l - 530.9f) * (l < 530.9f ? 0.0613f : 0.0322f);
        r.y = 0.821f * expf((-0.5f * t1) * t1) + 0.286f * expf((-0.5f * t2) * t2);
    }
    {
        float t1 = (l - 437.0f) * (l < 437.0f ? 0.0845f : 0.0278f);
        float t2 = (l - 459.0f) * (l < 459.0f ? 0.0385f : 0.0725f);
        r.z = 1.217f * expf((-0.5f * t1) * t1) + 0.681f * expf((-0.5f * t2) * t2);
    }
    return r;
}
// :38-47  mat4x3(observer columns) * weights
PTB_HD f3 observer_times(f4 lambda, f4 w) {
    f3 c0 = standard_observer(lambda.x), c1 = standard_observer(lambda.y);
    f3 c2 = standard_observer(lambda.z), c3 = standard_observer(lambda.w);
    return ((c0 * w.x + c1 * w.y) + c2 * w.z) + c3 * w.w;
}
// :50-55
PTB_HD f3 xyz_to_srgb(f3 c) {
    const f3 m0 = mk3(+3.2406f, -0.9689f, +0.0557f);
    const f3 m1 = mk3(-1.5372f, +1.8758f, -0.2040f);
    const f3 m2 = mk3(-0.4986f, +0.0415f, +1.0570f);
    return (m0 * c.x + m1 * c.y) + m2 * c.z;
}
// :169-192  Jakob-Hanika sigmoid of a quadratic.
PTB_HD float parametric_spectrum(f3 beta, float l) {
    float x = (beta.x * l + beta.y) * l + beta.z;
    return 0.5f + x / (2.0f * sqrtf(1.0f + x * x));
}
PTB_HDN f4 parametric_spectrum(f3 beta, f4 l) {
    return mk4(parametric_spectrum(beta, l.x), parametric_spectrum(beta, l.y), parametric_spectrum(beta, l.z), parametric_spectrum(beta, l.w));
}
PTB_HD f4 parametric_spectrum(f4 beta_and_intensity, f4 l) { return beta_and_intensity.w * parametric_spectrum(xyz(beta_and_intensity), l); }

// ---- textures (scene.glsl.inc:181-205) ----
// Vulkan REPEAT addressing at LOD 0 on the atlas layer: nearest = floor(u),
// linear = texels floor(u-0.5), +1 with fp32 weights frac(u-0.5).
PTB_HDN int wrap_repeat(int i, int n) { int m = i % n; return m < 0 ? m + n : m; }

// x, y already wrapped into the layer
PTB_HD f4 fetch_texel(const DeviceScene& s, u32 layer, int x, int y) {
    const float4* base = (const float4*)s.atlas_layers[layer];
    size_t o = (size_t)y * s.atlas_width + (size_t)x;
    float4 t = PTB_LDG(&base[o]);
    return mk4(t.x, t.y, t.z, t.w);
}

PTB_HDN f4 sample_texture(const DeviceScene& s, u32 index, f2 uv) {
    const uint4* rec = (const uint4*)&s.textures[index];
    uint4 r0 = PTB_LDG(&rec[0]);     // placement min.xy, max.xy
    uint4 r1 = PTB_LDG(&rec[1]);     // layer, type, flags, pad
    float u = gmix(u2f(r0.x), u2f(r0.z), gfract(uv.x)) * (float)s.atlas_width;
    float v = gmix(u2f(r0.y), u2f(r0.w), gfract(uv.y)) * (float)s.atlas_height;
    const int W = (int)s.atlas_width, H = (int)s.atlas_height;
    if (r1.z & 1u) return fetch_texel(s, r1.x, wrap_repeat((int)floorf(u), W), wrap_repeat((int)floorf(v), H));
    float ub = u - 0.5f, vb = v - 0.5f;
    float fu = floorf(ub), fv = floorf(vb);
    float a = ub - fu, b = vb - fv;
    const int i0 = wrap_repeat((int)fu, W), j0 = wrap_repeat((int)fv, H);
    const int i1 = i0 + 1 == W ? 0 : i0 + 1, j1 = j0 + 1 == H ? 0 : j0 + 1;
    f4 t00 = fetch_texel(s, r1.x, i0, j0), t10 = fetch_texel(s, r1.x, i1, j0);
    f4 t01 = fetch_texel(s, r1.x, i0, j1), t11 = fetch_texel(s, r1.x, i1, j1);
    f4 row0 = t00 * (1.0f - a) + t10 * a;
    f4 row1 = t01 * (1.0f - a) + t11 * a;
    return row0 * (1.0f - b) + row1 * b;
}

// scene.glsl.inc:209-229 — equirectangular sky; z-up world.
PTB_HDN f4 skybox_radiance(const DeviceScene& s, f3 dir, f4 lambda) {
    f4 spectrum;
    if (!(s.features & PTB_FEAT_SKY_TEXTURE) || s.globals.SkyboxTextureIndex == 0xFFFFFFFFu) {
        spectrum = mk4(0, 0, 100, 1);
    } else {
        float phi = atan2f(dir.y, dir.x);
        float theta = asinf(dir.z);
        spectrum = sample_texture(s, s.globals.SkyboxTextureIndex, mk2(0.5f + phi / PTB_TAU, 0.5f + theta / PTB_PI));
    }
    return parametric_spectrum(spectrum, lambda) * s.globals.SkyboxBrightness;
}

// ---- material attribute fetch (scene.glsl.inc:231-302) ----
PTB_HD u32 material_u32(const DeviceScene& s, u32 slot, u32 attr) { return PTB_LDG(&s.material_data[32u * slot + attr]); }
PTB_HD float material_f32(const DeviceScene& s, u32 slot, u32 attr) { return u2f(material_u32(s, slot, attr)); }
PTB_HD f3 material_f3(const DeviceScene& s, u32 slot, u32 attr) {
    return mk3(material_f32(s, slot, attr), material_f32(s, slot, attr + 1), material_f32(s, slot, attr + 2));
}
// reflectance = sigmoid-poly(beta) [* sigmoid-poly(texel.xyz)]
PTB_HDN f4 material_reflectance(const DeviceScene& s, u32 slot, u32 attr, f4 lambda, f2 uv) {
    f4 value = parametric_spectrum(material_f3(s, slot, attr), lambda);
    u32 tex = material_u32(s, slot, attr + 3);
    if ((s.features & PTB_FEAT_TEXTURES) && tex != 0xFFFFFFFFu) value = value * parametric_spectrum(xyz(sample_texture(s, tex, uv)), lambda);
    return value;
}
// scalar [* texel.r]
PTB_HDN float material_value(const DeviceScene& s, u32 slot, u32 attr, f2 uv) {
    float value = material_f32(s, slot, attr);
    u32 tex = material_u32(s, slot, attr + 1);
    if ((s.features & PTB_FEAT_TEXTURES) && tex != 0xFFFFFFFFu) value *= sample_texture(s, tex, uv).x;
    return value;
}

// ---- cameras (scene.glsl.inc:613-655) ----
// Pinhole and thin lens both draw two randoms for the aperture disk even at
// radius 0; the 360 camera draws none.  Output is in world space.
PTB_HDN Ray generate_camera_ray(const DeviceScene& s, u32 camera_index, f2 nsp, Rng& g) {
    const ptb_packed_camera* cam = &s.cameras[camera_index];
    const uint4 head = PTB_LDG((const uint4*)cam);            // model, focal, aperture, sensor distance
    const float2 sensor = PTB_LDG((const float2*)&cam->SensorSize[0]);
    const u32 model = head.x;
    const float focal = u2f(head.y), aperture = u2f(head.z), sensor_distance = u2f(head.w);
    Ray r;
    r.origin = mk3(0, 0, 0);
    r.velocity = mk3(0, 0, 0);
    r.duration = PTB_HIT_TIME_LIMIT;
    if (model == 0u) {
        f3 sp = mk3(-sensor.x * (nsp.x - 0.5f), -sensor.y * (0.5f - nsp.y), sensor_distance);
        f2 d = random_point_on_disk(g);
        r.origin = mk3(aperture * d.x, aperture * d.y, 0.0f);
        r.velocity = normalize(r.origin - sp);
    } else if (model == 1u) {
        f3 sp = mk3(-sensor.x * (nsp.x - 0.5f), -sensor.y * (0.5f - nsp.y), sensor_distance);
        f3 op = ((-sp) * focal) / (sp.z - focal);
        f2 d = random_point_on_disk(g);
        r.origin = mk3(aperture * d.x, aperture * d.y, 0.0f);
        r.velocity = normalize(op - r.origin);
    } else if (model == 2u) {
        float phi = (nsp.x - 0.5f) * PTB_TAU;
        float theta = (0.5f - nsp.y) * PTB_PI;
        float st, ct, sp2, cp2;
        sin_cos(theta, st, ct);
        sin_cos(phi, sp2, cp2);
        r.velocity = mk3(ct * sp2, st, -ct * cp2);
    }
    float to[16];
    const float4* m = (const float4*)&cam->Transform.To[0];
#pragma unroll
    for (int c = 0; c < 4; c++) { float4 col = PTB_LDG(&m[c]); to[4 * c] = col.x; to[4 * c + 1] = col.y; to[4 * c + 2] = col.z; to[4 * c + 3] = col.w; }
    Ray w;
    w.origin = mat_mul_point(to, r.origin);
    w.velocity = mat_mul_vector(to, r.velocity);
    w.duration = r.duration;
    return w;
}

}  // namespace ptb
