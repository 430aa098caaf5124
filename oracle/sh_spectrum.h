// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Restatement of /root/reference/src/core/spectrum.glsl.inc.
#pragma once
#include "sh_common.h"

namespace orc {

// spectrum.glsl.inc:10-34  Wyman et al. multi-lobe fit of the CIE 1931 observer.
inline vec3 SampleStandardObserver(float Lambda) {
    vec3 R;
    {
        float T1 = (Lambda - 442.0f) * (Lambda < 442.0f ? 0.0624f : 0.0374f);
        float T2 = (Lambda - 599.8f) * (Lambda < 599.8f ? 0.0264f : 0.0323f);
        float T3 = (Lambda - 501.1f) * (Lambda < 501.1f ? 0.0490f : 0.0382f);
        R.x = (0.362f * expf((-0.5f * T1) * T1) + 1.056f * expf((-0.5f * T2) * T2)) - 0.065f * expf((-0.5f * T3) * T3);
    }
    {
        float T1 = (Lambda - 568.8f) * (Lambda < 568.8f ? 0.0213f : 0.0247f);
        float T2 = (Lambda - 530.9f) * (Lambda < 530.9f ? 0.0613f : 0.0322f);
        R.y = 0.821f * expf((-0.5f * T1) * T1) + 0.286f * expf((-0.5f * T2) * T2);
    }
    {
        float T1 = (Lambda - 437.0f) * (Lambda < 437.0f ? 0.0845f : 0.0278f);
        float T2 = (Lambda - 459.0f) * (Lambda < 459.0f ? 0.0385f : 0.0725f);
        R.z = 1.217f * expf((-0.5f * T1) * T1) + 0.681f * expf((-0.5f * T2) * T2);
    }
    return R;
}

// spectrum.glsl.inc:38-47: mat4x3 of four observer columns; the only use is
// `M * vec4` (basic_scatter.glsl:170) = sum of columns weighted by the vec4.
inline vec3 ObserverTimes(vec4 Lambda, vec4 W) {
    vec3 C0 = SampleStandardObserver(Lambda.x);
    vec3 C1 = SampleStandardObserver(Lambda.y);
    vec3 C2 = SampleStandardObserver(Lambda.z);
    vec3 C3 = SampleStandardObserver(Lambda.w);
    return ((C0 * W.x + C1 * W.y) + C2 * W.z) + C3 * W.w;
}

// spectrum.glsl.inc:50-55 (column-major literal): returns M * XYZ.
inline vec3 XYZToSRGB(vec3 C) {
    const vec3 M0 = {+3.2406f, -0.9689f, +0.0557f};
    const vec3 M1 = {-1.5372f, +1.8758f, -0.2040f};
    const vec3 M2 = {-0.4986f, +0.0415f, +1.0570f};
    return (M0 * C.x + M1 * C.y) + M2 * C.z;
}

static const float CIE_ILLUMINANT_D65[471] = {
#include "d65_table.inc"
};

// spectrum.glsl.inc:159-164
inline float SampleIlluminantD65(float NormalizedLambda) {
    float Offset = NormalizedLambda * 470.0f;
    int Index = gl_clampi((int)Offset, 0, 469);
    return gl_mix(CIE_ILLUMINANT_D65[Index], CIE_ILLUMINANT_D65[Index + 1], Offset - (float)Index);
}

// spectrum.glsl.inc:169-192  Jakob–Hanika sigmoid-quadratic.
inline float SampleParametricSpectrum(vec3 Beta, float Lambda) {
    float X = (Beta.x * Lambda + Beta.y) * Lambda + Beta.z;
    return 0.5f + X / (2.0f * sqrtf(1.0f + X * X));
}
inline vec4 SampleParametricSpectrum(vec3 Beta, vec4 L) {
    return V4(SampleParametricSpectrum(Beta, L.x), SampleParametricSpectrum(Beta, L.y),
              SampleParametricSpectrum(Beta, L.z), SampleParametricSpectrum(Beta, L.w));
}
inline float SampleParametricSpectrum(vec4 BetaAndIntensity, float Lambda) {
    return BetaAndIntensity.w * SampleParametricSpectrum(xyz(BetaAndIntensity), Lambda);
}
inline vec4 SampleParametricSpectrum(vec4 BetaAndIntensity, vec4 L) {
    return BetaAndIntensity.w * SampleParametricSpectrum(xyz(BetaAndIntensity), L);
}

// spectrum.glsl.inc:197-215  16-sample D65 observation (preview colours only).
inline vec3 ObserveParametricSpectrumUnderD65(vec4 BetaAndIntensity) {
    const int SampleCount = 16;
    const float DeltaLambda = (CIE_LAMBDA_MAX - CIE_LAMBDA_MIN) / (float)SampleCount;
    vec3 Color = V3(0, 0, 0);
    for (int I = 0; I < SampleCount; I++) {
        float NormalizedLambda = (float)I / (float)(SampleCount - 1);
        float D65 = SampleIlluminantD65(NormalizedLambda) / 10566.864005f;
        float Lambda = gl_mix(CIE_LAMBDA_MIN, CIE_LAMBDA_MAX, NormalizedLambda);
        Color += ((SampleParametricSpectrum(BetaAndIntensity, Lambda) * D65) * SampleStandardObserver(Lambda)) * DeltaLambda;
    }
    return Color;
}
inline vec3 ObserveParametricSpectrumUnderD65(vec3 Beta) { return ObserveParametricSpectrumUnderD65(V4(Beta, 1.0f)); }

}  // namespace orc
