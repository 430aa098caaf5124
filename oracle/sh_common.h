// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// Restatement of /root/reference/src/core/common.glsl.inc as host C++.
// Each function cites the reference lines it follows.  Random-number draws are
// sequenced in explicit statements (GLSL evaluates arguments left to right;
// C++ does not) — SURVEY.md Appendix C.
#pragma once
#include "layouts.h"

namespace orc {

// common.glsl.inc:4-12  (PI/TAU are the reference's truncated literals, quirk Q1)
static const float K_INFINITY = 1e30f;
static const float K_EPSILON = 1e-9f;
static const float K_PI = 3.141592653f;
static const float K_TAU = 6.283185306f;
static const float HIT_TIME_LIMIT = 1048576.0f;
static const float CIE_LAMBDA_MIN = 360.0f;
static const float CIE_LAMBDA_MAX = 830.0f;
// common.glsl.inc:14-15
static const uint RENDER_FLAG_ACCUMULATE = 1u << 0;
static const uint RENDER_FLAG_SAMPLE_JITTER = 1u << 1;

// common.glsl.inc:23-28
struct ray { vec3 Origin; vec3 Velocity; float Duration; };

// common.glsl.inc:31-38
struct medium {
    uint Priority;
    vec4 IOR;
    vec4 AbsorptionRate;
    vec4 ScatteringRate;
    float ScatteringAnisotropy;
};

// common.glsl.inc:40-58
inline vec3 TransformPosition(vec3 P, const packed_transform& T) { return xyz(mul(T.To, V4(P, 1.0f))); }
inline vec3 TransformVector(vec3 V, const packed_transform& T) { return xyz(mul(T.To, V4(V, 0.0f))); }
inline vec3 TransformNormal(vec3 N, const packed_transform& T) { return normalize(xyz(mul(V4(N, 0.0f), T.From))); }
inline vec3 TransformDirection(vec3 D, const packed_transform& T) { return normalize(TransformVector(D, T)); }

// common.glsl.inc:60-67
inline ray TransformRay(ray R, const packed_transform& T) {
    ray O;
    O.Origin = TransformPosition(R.Origin, T);
    O.Velocity = TransformVector(R.Velocity, T);
    O.Duration = R.Duration;
    return O;
}

// common.glsl.inc:69-91  (direction NOT renormalised: object-space t == world t)
inline ray InverseTransformRay(ray R, const packed_transform& T) {
    ray O;
    O.Origin = xyz(mul(T.From, V4(R.Origin, 1.0f)));
    O.Velocity = xyz(mul(T.From, V4(R.Velocity, 0.0f)));
    O.Duration = R.Duration;
    return O;
}

// common.glsl.inc:93-100
inline vec3 SafeNormalize(vec3 V) {
    float LenSq = dot(V, V);
    if (LenSq < 1e-12f) return V3(0, 0, 1);
    return V / sqrtf(LenSq);
}

// common.glsl.inc:102-110
inline vec4 Sqr(vec4 X) { return X * X; }
inline float max4(vec4 V) { return fmaxf(fmaxf(V.x, V.y), fmaxf(V.z, V.w)); }

// common.glsl.inc:113-117
inline vec3 ComputeTangentVector(vec3 Normal) {
    vec3 V = fabsf(Normal.x) < 0.9f ? V3(1, 0, 0) : V3(0, 1, 0);
    return normalize(cross(V, Normal));
}

// common.glsl.inc:120-125   (Y = cross(X, Z), quirk Q18)
inline void ComputeCoordinateFrame(vec3 Z, vec3& X, vec3& Y) {
    vec3 V = fabsf(Z.x) < 0.9f ? V3(1, 0, 0) : V3(0, 1, 0);
    X = normalize(cross(V, Z));
    Y = cross(X, Z);
}

// common.glsl.inc:127-134
inline vec2 SignNotZero(vec2 V) { return V2(V.x >= 0.0f ? 1.0f : -1.0f, V.y >= 0.0f ? 1.0f : -1.0f); }

// common.glsl.inc:137-142  (octahedral encode, snorm16x2)
inline uint PackUnitVector(vec3 V) {
    float S = 1.0f / ((fabsf(V.x) + fabsf(V.y)) + fabsf(V.z));
    vec2 P = V2(V.x * S, V.y * S);
    if (V.z <= 0.0f) {
        vec2 Q = V2(1.0f - fabsf(P.y), 1.0f - fabsf(P.x));
        P = Q * SignNotZero(P);
    }
    return packSnorm2x16(P);
}

// common.glsl.inc:145-151
inline vec3 UnpackUnitVector(uint PackedV) {
    vec2 P = unpackSnorm2x16(PackedV);
    float Z = (1.0f - fabsf(P.x)) - fabsf(P.y);
    if (Z < 0.0f) {
        vec2 Q = V2(1.0f - fabsf(P.y), 1.0f - fabsf(P.x));
        P = Q * SignNotZero(P);
    }
    return normalize(V3(P.x, P.y, Z));
}

// common.glsl.inc:153-185  (true divisions; fminf/fmaxf NaN semantics, Q19).
// `strict` selects the canonical-tie variant: a box whose entry time equals the
// current reach is still visited, so a lower-ID primitive at equal t is found.
inline float IntersectBoundingBox(const ray& Ray, float Reach, vec3 Min, vec3 Max, bool strict = false) {
    vec3 MinT = (Min - Ray.Origin) / Ray.Velocity;
    vec3 MaxT = (Max - Ray.Origin) / Ray.Velocity;
    vec3 EarlierT = vmin(MinT, MaxT);
    vec3 LaterT = vmax(MinT, MaxT);
    float EntryT = fmaxf(fmaxf(EarlierT.x, EarlierT.y), EarlierT.z);
    float ExitT = fminf(fminf(LaterT.x, LaterT.y), LaterT.z);
    if (ExitT < EntryT) return K_INFINITY;
    if (ExitT <= 0.0f) return K_INFINITY;
    if (strict ? (EntryT > Reach) : (EntryT >= Reach)) return K_INFINITY;
    return EntryT;
}

// ---- RNG: common.glsl.inc:189-203 (32-bit LCG state, PCG RXS-M-XS output) ----
struct rng {
    uint State;
    uint Next() {
        State = State * 747796405u + 2891336453u;
        uint s = State;
        uint w = ((s >> ((s >> 28u) + 4u)) ^ s) * 277803737u;
        return (w >> 22u) ^ w;
    }
    // uint -> float rounds to nearest, so this can return exactly 1.0 (quirk Q2).
    float Uniform() { return (float)Next() / 4294967296.0f; }
};

// common.glsl.inc:205-210
inline vec2 RandomPointOnDisk(rng& G) {
    float R = sqrtf(G.Uniform());
    float Theta = G.Uniform() * K_TAU;
    return V2(R * cosf(Theta), R * sinf(Theta));
}

// common.glsl.inc:212-218
inline vec3 RandomDirection(rng& G) {
    float Z = 2.0f * G.Uniform() - 1.0f;
    float R = sqrtf(1.0f - Z * Z);
    float Phi = K_TAU * G.Uniform();
    return V3(R * cosf(Phi), R * sinf(Phi), Z);
}

// common.glsl.inc:228-239
inline vec3 RandomVonMisesFisher(rng& G, float Kappa) {
    float Xi = G.Uniform();
    float Z = 1.0f + (1.0f / Kappa) * logf(Xi + (1.0f - Xi) * expf(-2.0f * Kappa));
    float R = sqrtf(1.0f - Z * Z);
    float Phi = G.Uniform() * K_TAU;
    return V3(R * cosf(Phi), R * sinf(Phi), Z);
}

// common.glsl.inc:241-247
inline vec3 RandomVonMisesFisher(rng& G, float Kappa, vec3 Mu) {
    vec3 V = RandomVonMisesFisher(G, Kappa);
    vec3 MuX, MuY;
    ComputeCoordinateFrame(Mu, MuX, MuY);
    return SafeNormalize((V.x * MuX + V.y * MuY) + V.z * Mu);
}

// common.glsl.inc:249-254
inline float VonMisesFisherPDF(float Kappa, vec3 Mu, vec3 Direction) {
    if (Kappa < K_EPSILON) return 1.0f / (4.0f * K_PI);
    float C = Kappa / ((2.0f * K_PI) * (1.0f - expf(-2.0f * Kappa)));
    return C * expf(Kappa * (dot(Mu, Direction) - 1.0f));
}

// common.glsl.inc:259-276
inline vec3 SampleDirectionHG(float Anisotropy, float U1, float U2) {
    float Z;
    if (fabsf(Anisotropy) < 1e-3f) {
        Z = 1.0f - 2.0f * U1;
    } else {
        float G = Anisotropy;
        float S = (1.0f - G * G) / ((1.0f + G) - (2.0f * G) * U1);
        Z = -((1.0f + G * G) - S * S) / (2.0f * G);
    }
    float R = sqrtf(1.0f - Z * Z);
    float Phi = U2 * K_TAU;
    return V3(R * cosf(Phi), R * sinf(Phi), Z);
}

// common.glsl.inc:281-288
inline vec2 GGXRoughnessAlpha(float Roughness, float Anisotropy) {
    float R = Roughness;
    float S = 1.0f - Anisotropy;
    float AlphaX = (R * R) * sqrtf(2.0f / (1.0f + S * S));
    float AlphaY = S * AlphaX;
    return V2(AlphaX, AlphaY);
}

// common.glsl.inc:294-301
inline float GGXSmithG1(vec3 Direction, vec2 RoughnessAlpha) {
    vec3 DirectionSq = Direction * Direction;
    if (DirectionSq.z < K_EPSILON) return 0.0f;
    vec2 RoughnessAlphaSq = RoughnessAlpha * RoughnessAlpha;
    float AlphaSqByTanThetaSq = dot(RoughnessAlphaSq, V2(DirectionSq.x, DirectionSq.y)) / DirectionSq.z;
    return 2.0f / (1.0f + sqrtf(1.0f + AlphaSqByTanThetaSq));
}

// common.glsl.inc:306-346
inline vec3 GGXVisibleNormal(vec3 Direction, vec2 RoughnessAlpha, float U1, float U2) {
    vec3 Vz = SafeNormalize(V3(RoughnessAlpha.x * Direction.x, RoughnessAlpha.y * Direction.y, Direction.z));
    float LengthSq = Vz.x * Vz.x + Vz.y * Vz.y;
    vec3 Vx = LengthSq > 0.0f ? V3(-Vz.y, Vz.x, 0.0f) / sqrtf(LengthSq) : V3(1, 0, 0);
    vec3 Vy = cross(Vz, Vx);
    float R = sqrtf(U1);
    float Phi = K_TAU * U2;
    float S = 0.5f * (1.0f + Vz.z);
    float Tx = R * cosf(Phi);
    float Ty = (1.0f - S) * sqrtf(1.0f - Tx * Tx) + (S * R) * sinf(Phi);
    float Tz = sqrtf(fmaxf(0.0f, (1.0f - Tx * Tx) - Ty * Ty));
    vec3 Normal = (Tx * Vx + Ty * Vy) + Tz * Vz;
    return SafeNormalize(V3(RoughnessAlpha.x * Normal.x, RoughnessAlpha.y * Normal.y, fmaxf(0.0f, Normal.z)));
}

// common.glsl.inc:349-354
inline float GGXDistribution(vec3 Normal, vec2 RoughnessAlpha) {
    vec2 A = 1.0f / RoughnessAlpha;
    float B = dot(Normal * Normal, V3(A.x * A.x, A.y * A.y, 1.0f));
    return 1.0f / ((((K_PI * RoughnessAlpha.x) * RoughnessAlpha.y) * B) * B);
}

// common.glsl.inc:360-371
inline vec4 CauchyEmpiricalIOR(float BaseIOR, float AbbeNumber, vec4 Lambda) {
    const float LC = 656.3f, Ld = 587.6f, LF = 486.1f;
    float B = (BaseIOR - 1.0f) / (AbbeNumber * (1.0f / (LF * LF) - 1.0f / (LC * LC)));
    float A = BaseIOR - B / (Ld * Ld);
    return A + B / (Lambda * Lambda);
}

// common.glsl.inc:379-390
inline float ComputeCosThetaRefracted(float Eta, float CosTheta) {
    float C2 = 1.0f - (Eta * Eta) * (1.0f - CosTheta * CosTheta);
    return -gl_sign(CosTheta) * sqrtf(fmaxf(C2, 0.0f));
}
inline vec4 ComputeCosThetaRefracted(vec4 Eta, vec4 CosTheta) {
    vec4 C2 = 1.0f - (Eta * Eta) * (1.0f - CosTheta * CosTheta);
    return (-gl_sign(CosTheta)) * vsqrt(vmax(C2, 0.0f));
}

// common.glsl.inc:396-420
inline float FresnelDielectric(float Eta, float CosTheta1, float CosTheta2) {
    float Ks = Eta * CosTheta1;
    float SqrtRs = (Ks + CosTheta2) / (Ks - CosTheta2);
    float Kp = Eta * CosTheta2;
    float SqrtRp = (Kp + CosTheta1) / (Kp - CosTheta1);
    return 0.5f * (SqrtRs * SqrtRs + SqrtRp * SqrtRp);
}
inline vec4 FresnelDielectric(vec4 Eta, vec4 CosTheta1, vec4 CosTheta2) {
    vec4 Ks = Eta * CosTheta1;
    vec4 SqrtRs = (Ks + CosTheta2) / (Ks - CosTheta2);
    vec4 Kp = Eta * CosTheta2;
    vec4 SqrtRp = (Kp + CosTheta1) / (Kp - CosTheta1);
    return 0.5f * (SqrtRs * SqrtRs + SqrtRp * SqrtRp);
}
inline vec4 FresnelDielectric(vec4 Eta, vec4 CosTheta1) {
    return FresnelDielectric(Eta, CosTheta1, ComputeCosThetaRefracted(Eta, CosTheta1));
}

// common.glsl.inc:425-436  (F82-tint)
inline vec4 SchlickFresnelMetal(vec4 Base, vec4 Specular, float CosTheta) {
    const float CosThetaMax = 1.0f / 7.0f;
    vec4 FSchlick = Base + (1.0f - Base) * powf(1.0f - CosTheta, 5.0f);
    vec4 FSchlickMax = Base + (1.0f - Base) * powf(1.0f - CosThetaMax, 5.0f);
    vec4 FMax = Specular * FSchlickMax;
    const float Denominator = CosThetaMax * powf(1.0f - CosThetaMax, 6.0f);
    float Nominator = CosTheta * powf(1.0f - CosTheta, 6.0f);
    return FSchlick - (Nominator / Denominator) * (FSchlickMax - FMax);
}

}  // namespace orc
