// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// The FRAGMENT halves of the reference's two full-screen shaders compiled as C++ (second translation unit of
// oracle/_ref/libref.so): src/integrator/resolve.glsl (film -> tone-mapped linear RGB) and
// src/application/preview_render.glsl (the preview visualiser and its pick query).  One call per fragment, with the
// interpolated stage input at the fragment centre, (x + 0.5) / W, as the rasteriser delivers it for the reference's
// full-screen quad; the colour attachment's linear -> sRGB8 conversion is fixed-function hardware, not shader code,
// and stays with the known-answer tests.  Each shader gets its own namespace: they include common.glsl.inc /
// scene.glsl.inc again, whose interface blocks are globals here.
#include <cstring>

#include "ptb.h"
#include "ref_glsl.h"

#define VERTEX 0
#define FRAGMENT 1

namespace glsl {
// sampler2D over the film (RGBA32F).  Framebuffer and film have the same size, so every fragment samples its own
// texel centre and the filter mode does not matter.
struct sampler2D { const float* rgba = nullptr; int width = 0, height = 0; };
inline vec4 texture(const sampler2D& S, vec2 uv) {
    int x = (int)floorf(uv.x * (float)S.width), y = (int)floorf(uv.y * (float)S.height);
    x = x < 0 ? 0 : (x >= S.width ? S.width - 1 : x);
    y = y < 0 ? 0 : (y >= S.height ? S.height - 1 : y);
    const float* t = S.rgba + ((size_t)y * S.width + x) * 4;
    return vec4(t[0], t[1], t[2], t[3]);
}
}  // namespace glsl

namespace resolve_shader {
using namespace glsl;
vec4 gl_Position; int gl_VertexIndex;
#include "integrator/resolve.glsl"
}  // namespace resolve_shader

#undef COMMON_GLSL_INC
#undef SPECTRUM_GLSL_INC

namespace preview_shader {
using namespace glsl;
vec4 gl_Position; int gl_VertexIndex;
#include "application/preview_render.glsl"
void bind_scene(const ptb_scene_view* v) {
#include "ref_bind_scene.inc"
}
}  // namespace preview_shader

extern "C" {

// resolve.glsl:112-131 per pixel; out: w*h*3 floats (OutColor.rgb, before the attachment's sRGB encoding).
void ref_resolve(const float* accum, uint32_t w, uint32_t h, float brightness, uint32_t mode, float white, float* out_linear_rgb) {
    using namespace resolve_shader;
    SampleAccumulatorImageSampler.rgba = accum; SampleAccumulatorImageSampler.width = (int)w; SampleAccumulatorImageSampler.height = (int)h;
    Brightness = brightness; ToneMappingMode = mode; ToneMappingWhiteLevel = white;
    for (uint32_t y = 0; y < h; y++) for (uint32_t x = 0; x < w; x++) {
        FragmentUV = glsl::vec2(((float)x + 0.5f) / (float)w, ((float)y + 0.5f) / (float)h);
        resolve_main();
        float* o = out_linear_rgb + ((size_t)y * w + x) * 3;
        o[0] = OutColor.x; o[1] = OutColor.y; o[2] = OutColor.z;
    }
}

struct ref_preview_params {
    float To[16], From[16];
    uint32_t RenderMode; float Brightness; uint32_t SelectedShapeIndex;
    uint32_t RenderSizeX, RenderSizeY, MouseX, MouseY;
};

// preview_render.glsl:96-178 per pixel; rgba: w*h*4 floats; *hit_shape: the QuerySSBO word.
void ref_preview(const ptb_scene_view* v, const ref_preview_params* p, float* rgba, uint32_t* hit_shape) {
    using namespace preview_shader;
    bind_scene(v);
    std::memcpy(&CameraTransform, p->To, 128);
    RenderMode = p->RenderMode; Brightness = p->Brightness; SelectedShapeIndex = p->SelectedShapeIndex;
    RenderSizeX = p->RenderSizeX; RenderSizeY = p->RenderSizeY; MouseX = p->MouseX; MouseY = p->MouseY;
    HitShapeIndex = *hit_shape;
    for (uint32_t y = 0; y < p->RenderSizeY; y++) for (uint32_t x = 0; x < p->RenderSizeX; x++) {
        ScreenXY = glsl::vec2(((float)x + 0.5f) / (float)p->RenderSizeX, ((float)y + 0.5f) / (float)p->RenderSizeY);
        preview_render_main();
        float* o = rgba + ((size_t)y * p->RenderSizeX + x) * 4;
        o[0] = OutColor.x; o[1] = OutColor.y; o[2] = OutColor.z; o[3] = OutColor.w;
    }
    *hit_shape = HitShapeIndex;
}

}  // extern "C"
