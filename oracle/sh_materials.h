// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Restatement of the three active BSDFs and the material dispatch:
//   /root/reference/src/scene/basic_diffuse.glsl.inc
//   /root/reference/src/scene/basic_metal.glsl.inc
//   /root/reference/src/scene/basic_translucent.glsl.inc
//   /root/reference/src/scene/scene.glsl.inc:659-762
// Naming follows the material files (In = given direction, Out = sampled), which
// is the inverse of the integrator's naming (quirk Q8).
#pragma once
#include "sh_scene.h"

namespace orc {

// scene.glsl.inc:659-665
struct bsdf_parameters {
    uint MaterialIndex;
    vec2 TextureUV;
    vec4 Lambda;
    vec4 ExteriorIOR;
};

// ---------------- BasicDiffuse (basic_diffuse.glsl.inc) ----------------
static const uint BASIC_DIFFUSE_BASE_SPECTRUM = 1;  // :1

// :19-35  (pdf uses In.z of the *given* direction: quirk Q7)
inline bool BasicDiffuse_EvaluateBSDF(const scene_view& S, const bsdf_parameters& P, vec3 In, vec3 Out, vec4& Throughput, vec4& Probability) {
    (void)Out;
    vec4 Reflectance = MaterialTexturableReflectance(S, P.MaterialIndex, BASIC_DIFFUSE_BASE_SPECTRUM, P.Lambda, P.TextureUV);
    Probability = V4(In.z / K_PI);
    Throughput = Probability * Reflectance;
    return true;
}
// :38-50
inline bool BasicDiffuse_SampleBSDF(const scene_view& S, const bsdf_parameters& P, vec3 In, rng& G, vec3& Out, vec4& Throughput, vec4& Probability) {
    vec3 D = RandomDirection(G);
    Out = SafeNormalize(D + V3(0, 0, 1));
    return BasicDiffuse_EvaluateBSDF(S, P, In, Out, Throughput, Probability);
}

// ---------------- BasicMetal (basic_metal.glsl.inc) ----------------
static const uint BASIC_METAL_BASE_SPECTRUM = 1, BASIC_METAL_SPECULAR_SPECTRUM = 5;   // :1-4
static const uint BASIC_METAL_ROUGHNESS = 9, BASIC_METAL_ROUGHNESS_ANISOTROPY = 11;

// :6-28
inline void BasicMetal_GetParameters(const scene_view& S, const bsdf_parameters& P, vec4& Base, vec4& Specular, vec2& Alpha, bool& Rough) {
    Base = MaterialTexturableReflectance(S, P.MaterialIndex, BASIC_METAL_BASE_SPECTRUM, P.Lambda, P.TextureUV);
    Specular = MaterialTexturableReflectance(S, P.MaterialIndex, BASIC_METAL_SPECULAR_SPECTRUM, P.Lambda, P.TextureUV);
    float R = MaterialTexturableValue(S, P.MaterialIndex, BASIC_METAL_ROUGHNESS, P.TextureUV);
    float A = MaterialTexturableValue(S, P.MaterialIndex, BASIC_METAL_ROUGHNESS_ANISOTROPY, P.TextureUV);
    Alpha = GGXRoughnessAlpha(R, A);
    Rough = Alpha.x * Alpha.y > K_EPSILON;
}
// :40-43
inline bool BasicMetal_HasDiracBSDF(const scene_view& S, const bsdf_parameters& P) {
    return MaterialTexturableValue(S, P.MaterialIndex, BASIC_METAL_ROUGHNESS, P.TextureUV) < 1e-3f;
}
// :46-84
inline bool BasicMetal_EvaluateBSDF(const scene_view& S, const bsdf_parameters& P, vec3 In, vec3 Out, vec4& Throughput, vec4& Probability) {
    vec4 Base, Specular; vec2 Alpha; bool Rough;
    BasicMetal_GetParameters(S, P, Base, Specular, Alpha, Rough);
    if (In.z <= 0.0f || Out.z <= 0.0f || !Rough) return false;
    vec3 Half = SafeNormalize(In + Out);
    float Gm = GGXSmithG1(In, Alpha);
    float D = GGXDistribution(Half, Alpha);
    Probability = V4((Gm * D) / (4.0f * In.z));
    float Gs = GGXSmithG1(Out, Alpha);
    vec4 F = SchlickFresnelMetal(Base, Specular, dot(In, Half));
    Throughput = (Probability * Gs) * F;
    return true;
}
// :87-141
inline bool BasicMetal_SampleBSDF(const scene_view& S, const bsdf_parameters& P, vec3 In, rng& G, vec3& Out, vec4& Throughput, vec4& Probability) {
    vec4 Base, Specular; vec2 Alpha; bool Rough;
    BasicMetal_GetParameters(S, P, Base, Specular, Alpha, Rough);
    if (In.z <= 0.0f) return false;
    float U1 = G.Uniform();
    float U2 = G.Uniform();
    vec3 Normal = GGXVisibleNormal(In, Alpha, U1, U2);
    float CosThetaIn = fminf(dot(Normal, In), 1.0f);
    Out = ((2.0f * CosThetaIn) * Normal) - In;
    if (Out.z <= 0.0f) return false;
    Probability = V4(1.0f);
    if (Rough) {
        float Gm = GGXSmithG1(In, Alpha);
        float D = GGXDistribution(Normal, Alpha);
        Probability *= V4((Gm * D) / (4.0f * In.z));
    }
    float Gs = GGXSmithG1(Out, Alpha);
    vec4 F = SchlickFresnelMetal(Base, Specular, CosThetaIn);
    Throughput = (Probability * Gs) * F;
    return true;
}

// ---------------- BasicTranslucent (basic_translucent.glsl.inc) ----------------
static const uint BASIC_TRANSLUCENT_IOR = 1, BASIC_TRANSLUCENT_ABBE_NUMBER = 2;        // :1-8
static const uint BASIC_TRANSLUCENT_ROUGHNESS = 3, BASIC_TRANSLUCENT_ROUGHNESS_ANISOTROPY = 5;
static const uint BASIC_TRANSLUCENT_TRANSMISSION_SPECTRUM = 7, BASIC_TRANSLUCENT_TRANSMISSION_DEPTH = 10;
static const uint BASIC_TRANSLUCENT_SCATTERING_SPECTRUM = 11, BASIC_TRANSLUCENT_SCATTERING_ANISOTROPY = 14;

// :10-48
inline void BasicTranslucent_GetParameters(const scene_view& S, const bsdf_parameters& P, vec3 In, vec4& RelativeIOR, vec2& Alpha, bool& Rough) {
    vec4 InteriorIOR = CauchyEmpiricalIOR(MaterialFloat(S, P.MaterialIndex, BASIC_TRANSLUCENT_IOR),
                                          MaterialFloat(S, P.MaterialIndex, BASIC_TRANSLUCENT_ABBE_NUMBER), P.Lambda);
    if (In.z < 0.0f) RelativeIOR = InteriorIOR / P.ExteriorIOR;
    else RelativeIOR = P.ExteriorIOR / InteriorIOR;
    float R = MaterialTexturableValue(S, P.MaterialIndex, BASIC_TRANSLUCENT_ROUGHNESS, P.TextureUV);
    float A = MaterialTexturableValue(S, P.MaterialIndex, BASIC_TRANSLUCENT_ROUGHNESS_ANISOTROPY, P.TextureUV);
    Alpha = GGXRoughnessAlpha(R, A);
    Rough = Alpha.x * Alpha.y > K_EPSILON;
}

// :55-82
inline bool BasicTranslucent_LoadMedium(const scene_view& S, uint MaterialIndex, vec4 Lambda, medium& M) {
    M.IOR = CauchyEmpiricalIOR(MaterialFloat(S, MaterialIndex, BASIC_TRANSLUCENT_IOR),
                               MaterialFloat(S, MaterialIndex, BASIC_TRANSLUCENT_ABBE_NUMBER), Lambda);
    float Depth = MaterialFloat(S, MaterialIndex, BASIC_TRANSLUCENT_TRANSMISSION_DEPTH);
    if (Depth > 0.0f) {
        vec4 Extinction = (-vlog(SampleParametricSpectrum(MaterialVec3(S, MaterialIndex, BASIC_TRANSLUCENT_TRANSMISSION_SPECTRUM), Lambda))) / Depth;
        vec4 Scattering = SampleParametricSpectrum(MaterialVec3(S, MaterialIndex, BASIC_TRANSLUCENT_SCATTERING_SPECTRUM), Lambda) / Depth;
        M.AbsorptionRate = vmax(Extinction - Scattering, 0.0f);
        M.ScatteringRate = Scattering;
        M.ScatteringAnisotropy = MaterialFloat(S, MaterialIndex, BASIC_TRANSLUCENT_SCATTERING_ANISOTROPY);
    } else {
        M.AbsorptionRate = V4(0.0f);
        M.ScatteringRate = V4(0.0f);
        M.ScatteringAnisotropy = 0.0f;
    }
    return true;
}
// :84-87
inline bool BasicTranslucent_HasDiracBSDF(const scene_view& S, const bsdf_parameters& P) {
    return MaterialTexturableValue(S, P.MaterialIndex, BASIC_TRANSLUCENT_ROUGHNESS, P.TextureUV) < 1e-3f;
}

// :90-169
inline bool BasicTranslucent_EvaluateBSDF(const scene_view& S, const bsdf_parameters& P, vec3 In, vec3 Out, vec4& Throughput, vec4& Probability) {
    vec4 Eta; vec2 Alpha; bool Rough;
    BasicTranslucent_GetParameters(S, P, In, Eta, Alpha, Rough);
    if (!Rough) { Probability = V4(0.0f); Throughput = V4(0.0f); return true; }
    float Gm = GGXSmithG1(In, Alpha);
    if (In.z * Out.z > 0.0f) {
        vec3 Half = SafeNormalize(Out + In);
        float CosThetaIn = dot(Half, In);
        vec4 F = FresnelDielectric(Eta, V4(CosThetaIn));
        float D = GGXDistribution(Half, Alpha);
        Probability = ((F * Gm) * D) / (4.0f * In.z);
    } else {
        vec3 H1 = SafeNormalize(Out + In * Eta.x);
        vec3 H2 = SafeNormalize(Out + In * Eta.y);
        vec3 H3 = SafeNormalize(Out + In * Eta.z);
        vec3 H4 = SafeNormalize(Out + In * Eta.w);
        vec4 Ci = V4(dot(In, H1), dot(In, H2), dot(In, H3), dot(In, H4));
        vec4 Co = V4(dot(Out, H1), dot(Out, H2), dot(Out, H3), dot(Out, H4));
        vec4 F = FresnelDielectric(Eta, Ci, Co);
        vec4 D = V4(0.0f);
        if (Ci.x * Co.x < 0.0f) D.x = GGXDistribution(H1, Alpha);
        if (Ci.y * Co.y < 0.0f) D.y = GGXDistribution(H2, Alpha);
        if (Ci.z * Co.z < 0.0f) D.z = GGXDistribution(H3, Alpha);
        if (Ci.w * Co.w < 0.0f) D.w = GGXDistribution(H4, Alpha);
        vec4 J = vabs(Co) / Sqr(Ci * Eta + Co);
        Probability = (((D * (1.0f - F)) * Gm) * J) * vabs(Ci / In.z);
    }
    float Gs = GGXSmithG1(Out, Alpha);
    Throughput = Probability * Gs;
    return true;
}

// :172-339
inline bool BasicTranslucent_SampleBSDF(const scene_view& S, const bsdf_parameters& P, vec3 In, rng& G, vec3& Out, vec4& Throughput, vec4& Probability) {
    vec4 Eta; vec2 Alpha; bool Rough;
    BasicTranslucent_GetParameters(S, P, In, Eta, Alpha, Rough);
    float U1 = G.Uniform();
    float U2 = G.Uniform();
    vec3 Normal = GGXVisibleNormal(In * gl_sign(In.z), Alpha, U1, U2);
    float CosThetaIn = gl_clamp(dot(Normal, In), -1.0f, 1.0f);
    float CosThetaRefracted = ComputeCosThetaRefracted(Eta.x, CosThetaIn);
    float Reflectance = FresnelDielectric(Eta.x, CosThetaIn, CosThetaRefracted);
    if (G.Uniform() < Reflectance) {
        Out = ((2.0f * CosThetaIn) * Normal) - In;
        if (Out.z * In.z <= 0.0f) return false;
        vec4 F = FresnelDielectric(Eta, V4(CosThetaIn));
        Probability = F;
        if (Rough) {
            float Gm = GGXSmithG1(In, Alpha);
            float D = GGXDistribution(Normal, Alpha);
            Probability *= (Gm * D) / (4.0f * fabsf(In.z));
        }
        float Gs = GGXSmithG1(Out, Alpha);
        Throughput = Probability * Gs;
        return true;
    }
    Out = ((CosThetaRefracted + Eta.x * CosThetaIn) * Normal) - (Eta.x * In);
    if (Out.z * In.z >= 0.0f) return false;
    if (Rough) {
        vec3 N2 = SafeNormalize(Out + In * Eta.y);
        vec3 N3 = SafeNormalize(Out + In * Eta.z);
        vec3 N4 = SafeNormalize(Out + In * Eta.w);
        vec4 Ci = V4(CosThetaIn, dot(In, N2), dot(In, N3), dot(In, N4));
        vec4 Co = V4(CosThetaRefracted, dot(Out, N2), dot(Out, N3), dot(Out, N4));
        vec4 F = FresnelDielectric(Eta, Ci, Co);
        vec4 D = V4(0.0f);
        D.x = GGXDistribution(Normal, Alpha);
        if (Ci.y * Co.y < 0.0f) D.y = GGXDistribution(N2, Alpha);
        if (Ci.z * Co.z < 0.0f) D.z = GGXDistribution(N3, Alpha);
        if (Ci.w * Co.w < 0.0f) D.w = GGXDistribution(N4, Alpha);
        float Gm = GGXSmithG1(In, Alpha);
        vec4 J = vabs(Co) / Sqr(Ci * Eta + Co);
        Probability = (((D * (1.0f - F)) * Gm) * J) * vabs(Ci / In.z);
    } else {
        Probability = V4(1.0f - Reflectance, 0.0f, 0.0f, 0.0f);   // quirk Q22
    }
    float Gs = GGXSmithG1(Out, Alpha);
    Throughput = Probability * Gs;
    return true;
}

// ---------------- dispatch: scene.glsl.inc:687-762 ----------------
// Type 3 (OpenPBR) and unknown types fall through: shader include is disabled
// at scene.glsl.inc:685, so such materials render black / terminate the path.
inline bool MaterialLoadMedium(const scene_view& S, uint MaterialIndex, vec4 Lambda, medium& M) {
    uint Type = MaterialType(S, MaterialIndex);
    if (Type == MATERIAL_TYPE_BASIC_TRANSLUCENT) return BasicTranslucent_LoadMedium(S, MaterialIndex, Lambda, M);
    return false;   // diffuse / metal return false and leave Medium untouched
}
inline bool MaterialHasDiracBSDF(const scene_view& S, const bsdf_parameters& P) {
    uint Type = MaterialType(S, P.MaterialIndex);
    if (Type == MATERIAL_TYPE_BASIC_DIFFUSE) return false;
    if (Type == MATERIAL_TYPE_BASIC_METAL) return BasicMetal_HasDiracBSDF(S, P);
    if (Type == MATERIAL_TYPE_BASIC_TRANSLUCENT) return BasicTranslucent_HasDiracBSDF(S, P);
    return false;
}
inline bool MaterialEvaluateBSDF(const scene_view& S, const bsdf_parameters& P, vec3 In, vec3 Out, vec4& T, vec4& Pr) {
    uint Type = MaterialType(S, P.MaterialIndex);
    if (Type == MATERIAL_TYPE_BASIC_DIFFUSE) return BasicDiffuse_EvaluateBSDF(S, P, In, Out, T, Pr);
    if (Type == MATERIAL_TYPE_BASIC_METAL) return BasicMetal_EvaluateBSDF(S, P, In, Out, T, Pr);
    if (Type == MATERIAL_TYPE_BASIC_TRANSLUCENT) return BasicTranslucent_EvaluateBSDF(S, P, In, Out, T, Pr);
    return false;
}
inline bool MaterialSampleBSDF(const scene_view& S, const bsdf_parameters& P, vec3 In, rng& G, vec3& Out, vec4& T, vec4& Pr) {
    uint Type = MaterialType(S, P.MaterialIndex);
    if (Type == MATERIAL_TYPE_BASIC_DIFFUSE) return BasicDiffuse_SampleBSDF(S, P, In, G, Out, T, Pr);
    if (Type == MATERIAL_TYPE_BASIC_METAL) return BasicMetal_SampleBSDF(S, P, In, G, Out, T, Pr);
    if (Type == MATERIAL_TYPE_BASIC_TRANSLUCENT) return BasicTranslucent_SampleBSDF(S, P, In, G, Out, T, Pr);
    return false;
}
// scene.glsl.inc:696-701 (preview base colour)
inline vec3 MaterialBaseColor(const scene_view& S, uint MaterialIndex, vec2 UV) {
    uint Type = MaterialType(S, MaterialIndex);
    if (Type == MATERIAL_TYPE_BASIC_DIFFUSE) return MaterialTexturableColor(S, MaterialIndex, BASIC_DIFFUSE_BASE_SPECTRUM, UV);
    if (Type == MATERIAL_TYPE_BASIC_METAL) return MaterialTexturableColor(S, MaterialIndex, BASIC_METAL_BASE_SPECTRUM, UV);
    if (Type == MATERIAL_TYPE_BASIC_TRANSLUCENT) return MaterialColor(S, MaterialIndex, BASIC_TRANSLUCENT_TRANSMISSION_SPECTRUM);
    return V3(0, 0, 0);
}

}  // namespace orc
