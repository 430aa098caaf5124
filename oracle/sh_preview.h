// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// Restatement of the reference's preview visualiser (src/application/preview_render.glsl:96-178):
// one camera ray per pixel, Trace(), a colour chosen by the render mode, and the pick query
// (shape under the mouse pixel).  The fragment stage's interpolated ScreenXY is the pixel centre.
#pragma once
#include "sh_materials.h"

namespace orc {

struct preview_params {                      // preview_render.glsl:48-60 (push constants) + traversal tie rule
    packed_transform CameraTransform;
    uint RenderMode;
    float Brightness;
    uint SelectedShapeIndex;
    uint RenderSizeX, RenderSizeY;
    uint MouseX, MouseY;
    uint Tie;
};

static const float PREVIEW_COLORS[20][3] = {   // preview_render.glsl:16-38
    {0.902f, 0.098f, 0.294f}, {0.235f, 0.706f, 0.294f}, {1.000f, 0.882f, 0.098f}, {0.263f, 0.388f, 0.847f}, {0.961f, 0.510f, 0.192f},
    {0.569f, 0.118f, 0.706f}, {0.275f, 0.941f, 0.941f}, {0.941f, 0.196f, 0.902f}, {0.737f, 0.965f, 0.047f}, {0.980f, 0.745f, 0.745f},
    {0.000f, 0.502f, 0.502f}, {0.902f, 0.745f, 1.000f}, {0.604f, 0.388f, 0.141f}, {1.000f, 0.980f, 0.784f}, {0.502f, 0.000f, 0.000f},
    {0.667f, 1.000f, 0.765f}, {0.502f, 0.502f, 0.000f}, {1.000f, 0.847f, 0.694f}, {0.000f, 0.000f, 0.459f}, {0.502f, 0.502f, 0.502f},
};

// preview_render.glsl:96-178.  Returns the pixel colour; *Shape receives Hit.ShapeIndex.
inline vec3 PreviewPixel(const scene_view& S, const preview_params& P, uint PixelX, uint PixelY, uint* Shape) {
    vec2 ScreenXY = V2(((float)PixelX + 0.5f) / (float)P.RenderSizeX, ((float)PixelY + 0.5f) / (float)P.RenderSizeY);
    float AspectRatio = (float)P.RenderSizeX / (float)P.RenderSizeY;
    float NearX = (ScreenXY.x - 0.5f) * AspectRatio;
    float NearY = 0.5f - ScreenXY.y;

    ray Ray;
    Ray.Origin = V3(0, 0, 0);
    Ray.Velocity = normalize(V3(NearX, NearY, -1.0f));
    Ray.Duration = HIT_TIME_LIMIT;
    Ray = TransformRay(Ray, P.CameraTransform);

    hit Hit = Trace(S, Ray, (tie_mode)P.Tie);
    vec3 Color = V3(0, 0, 0);
    switch (P.RenderMode) {
        case 0: case 1:
            if (Hit.ShapeIndex == SHAPE_INDEX_NONE) {
                Color = XYZToSRGB(ObserveParametricSpectrumUnderD65(SampleSkyboxSpectrum(S, Ray.Velocity)));
            } else {
                Color = XYZToSRGB(MaterialBaseColor(S, Hit.MaterialIndex, Hit.UV));
                if (P.RenderMode == 1) Color = Color * dot(Hit.Normal, -Ray.Velocity);
            }
            break;
        case 2:
            if (Hit.ShapeIndex == SHAPE_INDEX_NONE) Color = 0.5f * (V3(1, 1, 1) - Ray.Velocity);
            else Color = 0.5f * (Hit.Normal + V3(1, 1, 1));
            break;
        case 3:
            if (Hit.ShapeIndex != SHAPE_INDEX_NONE) { const float* c = PREVIEW_COLORS[Hit.MaterialIndex % 20]; Color = V3(c[0], c[1], c[2]); }
            break;
        case 4:
            if (Hit.ShapeIndex != SHAPE_INDEX_NONE) { const float* c = PREVIEW_COLORS[Hit.PrimitiveIndex % 20]; Color = V3(c[0], c[1], c[2]); }
            break;
        case 5: Color = (V3(0, 1, 0) * (float)Hit.MeshComplexity) / 256.0f; break;
        case 6: Color = (V3(0, 1, 0) * (float)Hit.SceneComplexity) / 256.0f; break;
        default: break;
    }
    if (Hit.ShapeIndex == P.SelectedShapeIndex) Color = Color * V3(1.0f, 0.5f, 0.5f);
    Color = Color * P.Brightness;
    *Shape = Hit.ShapeIndex;
    return Color;
}

}  // namespace orc
