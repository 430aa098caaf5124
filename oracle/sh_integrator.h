// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Restatement of the integrator shaders:
//   /root/reference/src/integrator/basic.glsl.inc      (wavefront state, quantisation)
//   /root/reference/src/integrator/basic_trace.glsl    (K1)
//   /root/reference/src/integrator/basic_scatter.glsl  (K2/K3)
//   /root/reference/src/integrator/resolve.glsl        (K4)
//
// Generalisations over the reference, all identity when the
// reference's own preconditions hold (W,H multiples of 16, PathLengthLimit 0):
//  * slot index: tiles_x = ceil(W/16) instead of W/16; slots of out-of-image
//    pixels in partial edge tiles are never touched;
//  * PathLengthLimit > 0 ends a path after that many scatter vertices
//    (the reference plumbs the value but never reads it, D5);
//  * the resolve samples the accumulator 1:1 (no bilinear stretch) and applies
//    the sRGB OETF that the reference gets from its _SRGB swapchain format.
#pragma once
#include <vector>
#include "sh_materials.h"

namespace orc {

static const int ACTIVE_SHAPE_LIMIT = 4;   // basic.glsl.inc:11

// basic.glsl.inc:13-21
struct path {
    float NormalizedLambda0;
    vec4 Throughput;
    vec4 Probability;
    vec3 Sample;
    int ImageX, ImageY;
    uint ActiveShapeIndex[4];
    uint Length;   // generalisation: vertices processed so far (PathLengthLimit)
};

// basic.glsl.inc:76-85 == basic.cpp:17-25
struct push_constants {
    uint CameraIndex;
    uint RenderFlags;
    uint PathLengthLimit;
    float PathTerminationProbability;
    uint RandomSeed;
    uint Restart;
};

// basic.glsl.inc:23-59: SoA arrays, one 4-byte lane per slot.
struct wavefront_state {
    uint Width = 0, Height = 0, TilesX = 0, TilesY = 0, SlotCount = 0;
    // trace_buffer
    std::vector<float> OriginX, OriginY, OriginZ, Duration;
    std::vector<uint> PackedVelocity;
    std::vector<float> Time, TextureU, TextureV;
    std::vector<uint> ShapeAndMaterialIndex, PackedNormal, PackedTangentX;
    // path_buffer
    std::vector<int> ImagePosition;
    std::vector<float> NormalizedLambda0, Throughput0, Throughput1, Throughput2, Throughput3;
    std::vector<float> Probability0, Probability1, Probability2, Probability3, SampleR, SampleG, SampleB;
    std::vector<uint> ActiveShapeIndex01, ActiveShapeIndex23;
    std::vector<uint> PathLength;       // generalisation
    std::vector<uint> PrimitiveIndex;   // oracle-only: closest-hit primitive id per slot (not in reference state)
    // accumulator image, RGBA32F (integrator.cpp:47-60)
    std::vector<float> Accum;
    // counters
    uint64_t RaysTraced = 0, PathsFinished = 0;
    uint64_t NodeFetches = 0, FaceTests = 0, ShapeVisits = 0, MeshHits = 0;

    void Resize(uint W, uint H) {
        Width = W; Height = H;
        TilesX = (W + 15) / 16; TilesY = (H + 15) / 16;
        SlotCount = TilesX * TilesY * 256;
        size_t n = SlotCount;
        OriginX.assign(n, 0); OriginY.assign(n, 0); OriginZ.assign(n, 0); Duration.assign(n, 0);
        PackedVelocity.assign(n, 0); Time.assign(n, 0); TextureU.assign(n, 0); TextureV.assign(n, 0);
        ShapeAndMaterialIndex.assign(n, 0xFFFFFFFFu); PackedNormal.assign(n, 0); PackedTangentX.assign(n, 0);
        ImagePosition.assign(n, 0); NormalizedLambda0.assign(n, 0);
        Throughput0.assign(n, 0); Throughput1.assign(n, 0); Throughput2.assign(n, 0); Throughput3.assign(n, 0);
        Probability0.assign(n, 0); Probability1.assign(n, 0); Probability2.assign(n, 0); Probability3.assign(n, 0);
        SampleR.assign(n, 0); SampleG.assign(n, 0); SampleB.assign(n, 0);
        ActiveShapeIndex01.assign(n, 0xFFFFFFFFu); ActiveShapeIndex23.assign(n, 0xFFFFFFFFu);
        PathLength.assign(n, 0); PrimitiveIndex.assign(n, 0xFFFFFFFFu);
        Accum.assign((size_t)W * H * 4, 0.0f);
    }
    // basic_scatter.glsl:324-328 with tiles_x generalised (identity when W%16==0).
    uint SlotOf(uint X, uint Y) const { return ((Y / 16) * TilesX + (X / 16)) * 256 + (Y % 16) * 16 + (X % 16); }
};

// ---- basic.glsl.inc:87-225 load/store helpers ----
inline ray LoadTraceRay(const wavefront_state& W, uint I) {
    ray R;
    R.Origin = V3(W.OriginX[I], W.OriginY[I], W.OriginZ[I]);
    R.Velocity = UnpackUnitVector(W.PackedVelocity[I]);
    R.Duration = W.Duration[I];
    return R;
}
inline void StoreTraceRay(wavefront_state& W, uint I, const ray& R) {
    W.OriginX[I] = R.Origin.x; W.OriginY[I] = R.Origin.y; W.OriginZ[I] = R.Origin.z;
    W.PackedVelocity[I] = PackUnitVector(R.Velocity);
    W.Duration[I] = R.Duration;
}
// basic.glsl.inc:142-157: miss writes only the sentinel (Q6); ids truncate to 16 bits (Q10).
inline void StoreTraceHit(wavefront_state& W, uint I, const hit& H) {
    if (H.ShapeIndex == SHAPE_INDEX_NONE) { W.ShapeAndMaterialIndex[I] = 0xFFFFFFFFu; W.PrimitiveIndex[I] = 0xFFFFFFFFu; return; }
    W.ShapeAndMaterialIndex[I] = (H.ShapeIndex << 16) | H.MaterialIndex;
    W.Time[I] = H.Time;
    W.PackedNormal[I] = PackUnitVector(H.Normal);
    W.PackedTangentX[I] = PackUnitVector(H.TangentX);
    W.TextureU[I] = H.UV.x; W.TextureV[I] = H.UV.y;
    W.PrimitiveIndex[I] = H.PrimitiveIndex;
}
// basic.glsl.inc:99-131
inline void LoadTraceResult(const wavefront_state& W, uint I, ray& R, hit& H) {
    R = LoadTraceRay(W, I);
    memset(&H, 0, sizeof(H));
    uint SM = W.ShapeAndMaterialIndex[I];
    if (SM == 0xFFFFFFFFu) { H.ShapeIndex = SHAPE_INDEX_NONE; H.Time = HIT_TIME_LIMIT; return; }
    H.ShapeIndex = SM >> 16;
    H.MaterialIndex = SM & 0xFFFFu;
    H.Time = W.Time[I];
    H.Normal = UnpackUnitVector(W.PackedNormal[I]);
    H.TangentX = UnpackUnitVector(W.PackedTangentX[I]);
    H.TangentY = cross(H.Normal, H.TangentX);
    H.UV = V2(W.TextureU[I], W.TextureV[I]);
    H.Position = R.Origin + H.Time * R.Velocity;
}
// basic.glsl.inc:159-198 (Probability .z/.w swap on load, Q3)
inline path LoadPath(const wavefront_state& W, uint I) {
    path P;
    int IW = (int)W.Width;
    P.ImageX = W.ImagePosition[I] % IW;
    P.ImageY = W.ImagePosition[I] / IW;
    P.NormalizedLambda0 = W.NormalizedLambda0[I];
    P.Throughput = V4(W.Throughput0[I], W.Throughput1[I], W.Throughput2[I], W.Throughput3[I]);
    P.Probability.x = W.Probability0[I];
    P.Probability.y = W.Probability1[I];
    P.Probability.w = W.Probability2[I];
    P.Probability.z = W.Probability3[I];
    P.Sample = V3(W.SampleR[I], W.SampleG[I], W.SampleB[I]);
    uint A01 = W.ActiveShapeIndex01[I], A23 = W.ActiveShapeIndex23[I];
    P.ActiveShapeIndex[0] = A01 & 0xFFFFu; P.ActiveShapeIndex[1] = A01 >> 16;
    P.ActiveShapeIndex[2] = A23 & 0xFFFFu; P.ActiveShapeIndex[3] = A23 >> 16;
    for (int k = 0; k < 4; k++) if (P.ActiveShapeIndex[k] == 0xFFFFu) P.ActiveShapeIndex[k] = SHAPE_INDEX_NONE;
    P.Length = W.PathLength[I];
    return P;
}
// basic.glsl.inc:200-216.  NB the packing ORs a 32-bit SHAPE_INDEX_NONE into the
// low half, so a NONE in entry [0]/[2] wipes entry [1]/[3] — reproduced literally.
inline void StorePathVertexData(wavefront_state& W, uint I, const path& P) {
    W.Throughput0[I] = P.Throughput.x; W.Throughput1[I] = P.Throughput.y;
    W.Throughput2[I] = P.Throughput.z; W.Throughput3[I] = P.Throughput.w;
    W.Probability0[I] = P.Probability.x; W.Probability1[I] = P.Probability.y;
    W.Probability2[I] = P.Probability.w; W.Probability3[I] = P.Probability.z;
    W.SampleR[I] = P.Sample.x; W.SampleG[I] = P.Sample.y; W.SampleB[I] = P.Sample.z;
    W.ActiveShapeIndex01[I] = (P.ActiveShapeIndex[1] << 16) | P.ActiveShapeIndex[0];
    W.ActiveShapeIndex23[I] = (P.ActiveShapeIndex[3] << 16) | P.ActiveShapeIndex[2];
    W.PathLength[I] = P.Length;
}
// basic.glsl.inc:218-225
inline void StorePath(wavefront_state& W, uint I, const path& P) {
    W.ImagePosition[I] = P.ImageY * (int)W.Width + P.ImageX;
    W.NormalizedLambda0[I] = P.NormalizedLambda0;
    StorePathVertexData(W, I, P);
}

// Everything one scatter invocation needs (descriptor sets + push constants).
struct shade_ctx {
    const scene_view* S;
    push_constants PC;
    rng G;
};

// basic_scatter.glsl:7-42
inline void GenerateNewPath(shade_ctx& C, wavefront_state& W, uint Index, int X, int Y) {
    vec2 SamplePosition = V2((float)X, (float)Y);
    if ((C.PC.RenderFlags & RENDER_FLAG_SAMPLE_JITTER) != 0) {
        float Jx = C.G.Uniform();
        float Jy = C.G.Uniform();
        SamplePosition = SamplePosition + V2(Jx, Jy);
    } else {
        SamplePosition = SamplePosition + V2(0.5f, 0.5f);
    }
    vec2 NSP = SamplePosition / V2((float)W.Width, (float)W.Height);
    ray Ray = GenerateCameraRay(C.S->Cameras[C.PC.CameraIndex], NSP, C.G);
    StoreTraceRay(W, Index, Ray);
    path P;
    P.ImageX = X; P.ImageY = Y;
    P.NormalizedLambda0 = C.G.Uniform();
    P.Throughput = V4(1.0f);
    P.Probability = V4(1.0f);
    P.Sample = V3(0.0f);
    for (int k = 0; k < 4; k++) P.ActiveShapeIndex[k] = SHAPE_INDEX_NONE;
    P.Length = 0;
    StorePath(W, Index, P);
}

// basic_scatter.glsl:44-64.  For a shape whose material has no medium
// (diffuse/metal) the reference leaves `Medium` uninitialised; the oracle pins
// that case to vacuum (IOR 1, no absorption/scattering).
inline medium ResolveMedium(const shade_ctx& C, uint ShapeIndex, vec4 Lambda) {
    medium M;
    M.IOR = V4(1.0f); M.AbsorptionRate = V4(0.0f); M.ScatteringRate = V4(0.0f); M.ScatteringAnisotropy = 0.0f;
    if (ShapeIndex == SHAPE_INDEX_NONE) {
        M.Priority = 0xFFFFFFFFu;
        M.ScatteringRate = V4(C.S->Globals->SceneScatterRate);
    } else {
        MaterialLoadMedium(*C.S, C.S->Shapes[ShapeIndex].MaterialIndex, Lambda, M);
        M.Priority = ShapeIndex;
    }
    return M;
}

// basic_scatter.glsl:68-109   (integrator naming: Out = toward viewer, In = sampled)
inline bool SampleSurfaceIntegrand(shade_ctx& C, const hit& Hit, const bsdf_parameters& P, vec3 Out, vec3& In, vec4& Throughput, vec4& Probability) {
    const packed_scene_globals& G = *C.S->Globals;
    float LightProbability = MaterialHasDiracBSDF(*C.S, P) ? 0.0f : G.SkyboxSamplingProbability;
    vec4 MaterialPDF = V4(0.0f);
    vec3 SkyMean = V3(dot(G.SkyboxMeanDirection, Hit.TangentX), dot(G.SkyboxMeanDirection, Hit.TangentY), dot(G.SkyboxMeanDirection, Hit.Normal));
    if (C.G.Uniform() < LightProbability) {
        In = RandomVonMisesFisher(C.G, G.SkyboxConcentration, SkyMean);
        if (In.z < 0.0f) return false;
        if (!MaterialEvaluateBSDF(*C.S, P, Out, In, Throughput, MaterialPDF)) return false;
    } else {
        if (!MaterialSampleBSDF(*C.S, P, Out, C.G, In, Throughput, MaterialPDF)) return false;
    }
    vec4 SkyboxPDF = V4(VonMisesFisherPDF(G.SkyboxConcentration, SkyMean, In));
    Probability = LightProbability * SkyboxPDF + (1.0f - LightProbability) * MaterialPDF;
    return true;
}

// basic_scatter.glsl:114-310
inline bool Scatter(shade_ctx& C, path& Path, ray& Ray, const hit& Hit) {
    float L0 = Path.NormalizedLambda0;
    vec4 Lambda = V4(gl_mix(CIE_LAMBDA_MIN, CIE_LAMBDA_MAX, L0),
                     gl_mix(CIE_LAMBDA_MIN, CIE_LAMBDA_MAX, gl_fract(L0 + 0.25f)),
                     gl_mix(CIE_LAMBDA_MIN, CIE_LAMBDA_MAX, gl_fract(L0 + 0.50f)),
                     gl_mix(CIE_LAMBDA_MIN, CIE_LAMBDA_MAX, gl_fract(L0 + 0.75f)));

    uint ActiveShapeIndex = SHAPE_INDEX_NONE;
    for (int k = 0; k < ACTIVE_SHAPE_LIMIT; k++)
        if (Path.ActiveShapeIndex[k] < ActiveShapeIndex) ActiveShapeIndex = Path.ActiveShapeIndex[k];

    medium Medium = ResolveMedium(C, ActiveShapeIndex, Lambda);

    Path.Throughput *= vexp((-Medium.AbsorptionRate) * Hit.Time);

    float ScatteringTime = HIT_TIME_LIMIT;
    if (Medium.ScatteringRate.x > 0.0f)
        ScatteringTime = (-logf(C.G.Uniform())) / Medium.ScatteringRate.x;

    if (Hit.Time >= ScatteringTime) {
        if (ScatteringTime < HIT_TIME_LIMIT) {
            Ray.Origin += Ray.Velocity * ScatteringTime;
            vec3 X, Y, Z = Ray.Velocity;
            ComputeCoordinateFrame(Z, X, Y);
            float U1 = C.G.Uniform();
            float U2 = C.G.Uniform();
            vec3 Scattered = SampleDirectionHG(Medium.ScatteringAnisotropy, U1, U2);
            vec4 Density = Medium.ScatteringRate * vexp((-Medium.ScatteringRate) * ScatteringTime);
            Density /= fmaxf(K_EPSILON, max4(Density));
            Path.Throughput *= Density;
            Path.Probability *= Density;
            Ray.Velocity = normalize((X * Scattered.x + Y * Scattered.y) + Z * Scattered.z);
            Ray.Duration = HIT_TIME_LIMIT;
        } else {
            vec4 Emission = SampleSkyboxRadiance(*C.S, Ray.Velocity, Lambda);
            float ClusterPDF = ((Path.Probability.x + Path.Probability.y) + Path.Probability.z) + Path.Probability.w;
            Path.Sample += ObserverTimes(Lambda, Emission * Path.Throughput) / ClusterPDF;
            Path.Probability = V4(0.0f);
        }
        return max4(Path.Probability) > K_EPSILON;
    }

    vec3 In;
    vec3 Out = -V3(dot(Ray.Velocity, Hit.TangentX), dot(Ray.Velocity, Hit.TangentY), dot(Ray.Velocity, Hit.Normal));

    bool IsRealSurface = true;
    vec4 ExteriorIOR = V4(1.0f);
    uint ShapePriority = Hit.ShapeIndex;

    if (Out.z > 0.0f) {
        IsRealSurface = Medium.Priority > ShapePriority;
        if (IsRealSurface) ExteriorIOR = Medium.IOR;
    } else {
        IsRealSurface = Medium.Priority == ShapePriority;
        if (IsRealSurface) {
            uint ExteriorShapeIndex = SHAPE_INDEX_NONE;
            for (int k = 0; k < ACTIVE_SHAPE_LIMIT; k++) {
                if (Path.ActiveShapeIndex[k] == ActiveShapeIndex) continue;
                if (Path.ActiveShapeIndex[k] < ExteriorShapeIndex) ExteriorShapeIndex = Path.ActiveShapeIndex[k];
            }
            medium Exterior = ResolveMedium(C, ExteriorShapeIndex, Lambda);
            ExteriorIOR = Exterior.IOR;
        }
    }

    if (IsRealSurface) {
        bsdf_parameters Parameters;
        Parameters.MaterialIndex = Hit.MaterialIndex;
        Parameters.TextureUV = Hit.UV;
        Parameters.Lambda = Lambda;
        Parameters.ExteriorIOR = ExteriorIOR;
        vec4 Throughput = V4(0.0f), Probability = V4(0.0f);
        if (!SampleSurfaceIntegrand(C, Hit, Parameters, Out, In, Throughput, Probability)) return false;
        float Scale = 1.0f / fmaxf(K_EPSILON, max4(Probability));
        Path.Throughput *= Throughput * Scale;
        Path.Probability *= Probability * Scale;
    } else {
        In = -Out;
    }

    if (In.z * Out.z < 0.0f) {
        if (Out.z > 0.0f) {
            for (int k = 0; k < ACTIVE_SHAPE_LIMIT; k++)
                if (Path.ActiveShapeIndex[k] == SHAPE_INDEX_NONE) { Path.ActiveShapeIndex[k] = Hit.ShapeIndex; break; }
        } else {
            for (int k = 0; k < ACTIVE_SHAPE_LIMIT; k++)
                if (Path.ActiveShapeIndex[k] == Hit.ShapeIndex) { Path.ActiveShapeIndex[k] = SHAPE_INDEX_NONE; break; }
        }
    }

    if (C.G.Uniform() < C.PC.PathTerminationProbability) return false;
    Path.Probability *= 1.0f - C.PC.PathTerminationProbability;

    Ray.Velocity = (In.x * Hit.TangentX + In.y * Hit.TangentY) + In.z * Hit.Normal;
    Ray.Origin = Hit.Position + 1e-3f * Ray.Velocity;
    Ray.Duration = HIT_TIME_LIMIT;
    return max4(Path.Probability) > K_EPSILON;
}

// basic_scatter.glsl:312-360 for one pixel (X,Y) — gl_GlobalInvocationID.
inline void ScatterMain(const scene_view& S, const push_constants& PC, wavefront_state& W, uint X, uint Y) {
    shade_ctx C;
    C.S = &S; C.PC = PC;
    C.G.State = Y * 65537u + X + PC.RandomSeed * 277803737u;
    if (X >= W.Width || Y >= W.Height) return;
    uint Index = W.SlotOf(X, Y);
    if (PC.Restart != 0) {
        GenerateNewPath(C, W, Index, (int)X, (int)Y);
        size_t o = ((size_t)Y * W.Width + X) * 4;
        W.Accum[o] = W.Accum[o + 1] = W.Accum[o + 2] = W.Accum[o + 3] = 0.0f;
        return;
    }
    path Path = LoadPath(W, Index);
    ray Ray; hit Hit;
    LoadTraceResult(W, Index, Ray, Hit);
    bool Alive = Scatter(C, Path, Ray, Hit);
    Path.Length += 1;
    if (Alive && PC.PathLengthLimit != 0 && Path.Length >= PC.PathLengthLimit) Alive = false;   // generalisation (D5)
    if (Alive) {
        StoreTraceRay(W, Index, Ray);
        StorePathVertexData(W, Index, Path);
    } else {
        size_t o = ((size_t)Path.ImageY * W.Width + Path.ImageX) * 4;
        vec4 V = V4(Path.Sample, 1.0f);
        if ((PC.RenderFlags & RENDER_FLAG_ACCUMULATE) != 0)
            V = V + V4(W.Accum[o], W.Accum[o + 1], W.Accum[o + 2], W.Accum[o + 3]);
        W.Accum[o] = V.x; W.Accum[o + 1] = V.y; W.Accum[o + 2] = V.z; W.Accum[o + 3] = V.w;
        GenerateNewPath(C, W, Index, Path.ImageX, Path.ImageY);
    }
}

// basic_trace.glsl:7-16 for one slot.
inline void TraceMain(const scene_view& S, wavefront_state& W, uint Index, tie_mode Tie, hit* OutHit = nullptr) {
    ray Ray = LoadTraceRay(W, Index);
    hit Hit = Trace(S, Ray, Tie);
    StoreTraceHit(W, Index, Hit);
    if (OutHit) *OutHit = Hit;
}

// ---- resolve.glsl:60-130 ----
static const uint TONE_MAPPING_MODE_CLAMP = 0, TONE_MAPPING_MODE_REINHARD = 1, TONE_MAPPING_MODE_HABLE = 2, TONE_MAPPING_MODE_ACES = 3;
struct resolve_parameters { float Brightness; uint ToneMappingMode; float ToneMappingWhiteLevel; };

inline float Luminance(vec3 C) { return dot(C, V3(0.2126f, 0.7152f, 0.0722f)); }
// resolve.glsl:65-72 (no zero-luminance guard, Q15)
inline vec3 ToneMapReinhard(vec3 Color, float WhiteLevel) {
    float OldL = Luminance(Color);
    float MaxL = WhiteLevel;
    float N = OldL * (1.0f + (OldL / (MaxL * MaxL)));
    float NewL = N / (1.0f + OldL);
    return (Color * NewL) / OldL;
}
// resolve.glsl:74-88
inline vec3 ToneMapHablePartial(vec3 X) {
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    vec3 Num = X * (A * X + V3(C * B)) + V3(D * E);
    vec3 Den = X * (A * X + V3(B)) + V3(D * F);
    return (Num / Den) - V3(E / F);
}
inline vec3 ToneMapHable(vec3 Color) {
    vec3 Current = ToneMapHablePartial(Color * 2.0f);
    vec3 WhiteScale = V3(1.0f) / ToneMapHablePartial(V3(11.2f));
    return Current * WhiteScale;
}
// resolve.glsl:90-110 (column-major literals)
inline vec3 ToneMapACES(vec3 Color) {
    const vec3 I0 = {0.59719f, 0.07600f, 0.02840f}, I1 = {0.35458f, 0.90834f, 0.13383f}, I2 = {0.04823f, 0.01566f, 0.83777f};
    const vec3 O0 = {1.60475f, -0.10208f, -0.00327f}, O1 = {-0.53108f, 1.10813f, -0.07276f}, O2 = {-0.07367f, -0.00605f, 1.07602f};
    vec3 V = (I0 * Color.x + I1 * Color.y) + I2 * Color.z;
    vec3 A = V * (V + V3(0.0245786f)) - V3(0.000090537f);
    vec3 B = V * (0.983729f * V + V3(0.4329510f)) + V3(0.238081f);
    vec3 R = A / B;
    return (O0 * R.x + O1 * R.y) + O2 * R.z;
}
// resolve.glsl:112-130: linear (pre-OETF) colour for one accumulator texel.
inline vec3 ResolvePixel(vec4 Value, const resolve_parameters& P) {
    vec3 Color = V3(0, 0, 0);
    if (Value.w > 0.0f) Color = XYZToSRGB((P.Brightness * xyz(Value)) / Value.w);
    if (P.ToneMappingMode == TONE_MAPPING_MODE_CLAMP) Color = V3(gl_clamp(Color.x, 0, 1), gl_clamp(Color.y, 0, 1), gl_clamp(Color.z, 0, 1));
    if (P.ToneMappingMode == TONE_MAPPING_MODE_REINHARD) Color = ToneMapReinhard(Color, P.ToneMappingWhiteLevel);
    if (P.ToneMappingMode == TONE_MAPPING_MODE_HABLE) Color = ToneMapHable(Color);
    if (P.ToneMappingMode == TONE_MAPPING_MODE_ACES) Color = ToneMapACES(Color);
    return Color;
}
// sRGB OETF + UNORM8 conversion the reference obtains from its B8G8R8A8_SRGB
// swapchain (vulkan.cpp:1407); Vulkan spec "sRGB EOTF^-1" then round(x*255).
inline uint8_t LinearToSRGB8(float L) {
    if (!(L > 0.0f)) L = 0.0f;       // also maps NaN to 0
    if (L > 1.0f) L = 1.0f;
    float E = L <= 0.0031308f ? L * 12.92f : 1.055f * powf(L, 1.0f / 2.4f) - 0.055f;
    return (uint8_t)(int)(E * 255.0f + 0.5f);
}

}  // namespace orc
