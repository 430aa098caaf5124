// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
// Restatement of /root/reference/src/scene/scene.glsl.inc:181-655
// (textures, skybox, material attribute fetch, intersection, Trace, cameras).
#pragma once
#include "sh_spectrum.h"

namespace orc {

// scene.glsl.inc:7-28
static const uint SHAPE_INDEX_NONE = 0xFFFFFFFFu;
static const uint TEXTURE_INDEX_NONE = 0xFFFFFFFFu;
static const uint SHAPE_TYPE_MESH_INSTANCE = 0, SHAPE_TYPE_PLANE = 1, SHAPE_TYPE_SPHERE = 2, SHAPE_TYPE_CUBE = 3;
static const uint TEXTURE_FLAG_FILTER_NEAREST = 1u << 0;
static const uint MATERIAL_TYPE_BASIC_DIFFUSE = 0, MATERIAL_TYPE_BASIC_METAL = 1, MATERIAL_TYPE_BASIC_TRANSLUCENT = 2;
static const uint CAMERA_MODEL_PINHOLE = 0, CAMERA_MODEL_THIN_LENS = 1, CAMERA_MODEL_360 = 2;

// Tie handling: FAITHFUL keeps the reference's traversal-order
// dependent rules; CANONICAL picks the lexicographically lowest
// (t, ShapeIndex, PrimitiveIndex) and visits boxes whose entry equals the reach.
enum tie_mode { TIE_FAITHFUL = 0, TIE_CANONICAL = 1 };

// scene.glsl.inc:102-119 (+ byte-model counters of SURVEY.md §8(d))
struct hit {
    float Time;
    uint ShapeIndex;
    vec3 Position;
    vec3 Normal;
    vec3 TangentX;
    uint MaterialIndex;
    vec2 UV;
    vec3 TangentY;
    uint ShapeType;
    vec3 PrimitiveCoordinates;
    uint PrimitiveIndex;
    uint SceneComplexity;
    uint MeshComplexity;
    // oracle-only counters: 32-B node records fetched, faces tested, shapes visited
    uint NodeFetches;
    uint FaceTests;
    uint ShapeVisits;
};

// ---- textures: scene.glsl.inc:181-205 + Vulkan 1.0 §15 texel addressing ----
// REPEAT addressing, LOD 0, unnormalised coordinate u = s*size; NEAREST takes
// floor(u); LINEAR takes texels floor(u-0.5), +1 with weights frac(u-0.5).
// Weights are kept in fp32 (hardware filters use >= 8 fractional bits; unpinned).
inline int wrap_repeat(int i, int n) { int m = i % n; return m < 0 ? m + n : m; }

inline vec4 fetch_texel(const scene_view& S, uint layer, int x, int y) {
    const float* L = S.AtlasLayers[layer];
    size_t o = ((size_t)wrap_repeat(y, (int)S.AtlasHeight) * S.AtlasWidth + (size_t)wrap_repeat(x, (int)S.AtlasWidth)) * 4;
    return V4(L[o], L[o + 1], L[o + 2], L[o + 3]);
}

inline vec4 SampleTexture(const scene_view& S, uint Index, vec2 UV) {
    const packed_texture& T = S.Textures[Index];
    float U = gl_mix(T.AtlasPlacementMinimum.x, T.AtlasPlacementMaximum.x, gl_fract(UV.x));
    float V = gl_mix(T.AtlasPlacementMinimum.y, T.AtlasPlacementMaximum.y, gl_fract(UV.y));
    float u = U * (float)S.AtlasWidth;
    float v = V * (float)S.AtlasHeight;
    if ((T.Flags & TEXTURE_FLAG_FILTER_NEAREST) != 0) {
        return fetch_texel(S, T.AtlasImageIndex, (int)floorf(u), (int)floorf(v));
    }
    float ub = u - 0.5f, vb = v - 0.5f;
    float fu = floorf(ub), fv = floorf(vb);
    float a = ub - fu, b = vb - fv;
    int i0 = (int)fu, j0 = (int)fv;
    vec4 t00 = fetch_texel(S, T.AtlasImageIndex, i0, j0);
    vec4 t10 = fetch_texel(S, T.AtlasImageIndex, i0 + 1, j0);
    vec4 t01 = fetch_texel(S, T.AtlasImageIndex, i0, j0 + 1);
    vec4 t11 = fetch_texel(S, T.AtlasImageIndex, i0 + 1, j0 + 1);
    vec4 r0 = t00 * (1.0f - a) + t10 * a;
    vec4 r1 = t01 * (1.0f - a) + t11 * a;
    return r0 * (1.0f - b) + r1 * b;
}

// scene.glsl.inc:209-221
inline vec4 SampleSkyboxSpectrum(const scene_view& S, vec3 Direction) {
    if (S.Globals->SkyboxTextureIndex == TEXTURE_INDEX_NONE) return V4(0, 0, 100, 1);
    float Phi = atan2f(Direction.y, Direction.x);
    float Theta = asinf(Direction.z);
    float U = 0.5f + Phi / K_TAU;
    float V = 0.5f + Theta / K_PI;
    return SampleTexture(S, S.Globals->SkyboxTextureIndex, V2(U, V));
}

// scene.glsl.inc:225-229
inline vec4 SampleSkyboxRadiance(const scene_view& S, vec3 Direction, vec4 Lambda) {
    vec4 Spectrum = SampleSkyboxSpectrum(S, Direction);
    return SampleParametricSpectrum(Spectrum, Lambda) * S.Globals->SkyboxBrightness;
}

// ---- material attribute fetch: scene.glsl.inc:231-302 ----
inline uint MaterialType(const scene_view& S, uint Slot) { return S.MaterialData[32 * Slot]; }
inline uint MaterialUint(const scene_view& S, uint Slot, uint A) { return S.MaterialData[32 * Slot + A]; }
inline float MaterialFloat(const scene_view& S, uint Slot, uint A) { return uintBitsToFloat(MaterialUint(S, Slot, A)); }
inline vec3 MaterialVec3(const scene_view& S, uint Slot, uint A) {
    return V3(MaterialFloat(S, Slot, A), MaterialFloat(S, Slot, A + 1), MaterialFloat(S, Slot, A + 2));
}
inline vec3 MaterialColor(const scene_view& S, uint Slot, uint A) {
    return ObserveParametricSpectrumUnderD65(MaterialVec3(S, Slot, A));
}
inline vec3 MaterialTexturableColor(const scene_view& S, uint Slot, uint A, vec2 UV) {
    vec3 Color = ObserveParametricSpectrumUnderD65(MaterialVec3(S, Slot, A));
    uint Tex = MaterialUint(S, Slot, A + 3);
    if (Tex != TEXTURE_INDEX_NONE) Color = Color * ObserveParametricSpectrumUnderD65(xyz(SampleTexture(S, Tex, UV)));
    return Color;
}
inline vec4 MaterialTexturableReflectance(const scene_view& S, uint Slot, uint A, vec4 Lambda, vec2 UV) {
    vec4 Value = SampleParametricSpectrum(MaterialVec3(S, Slot, A), Lambda);
    uint Tex = MaterialUint(S, Slot, A + 3);
    if (Tex != TEXTURE_INDEX_NONE) Value *= SampleParametricSpectrum(xyz(SampleTexture(S, Tex, UV)), Lambda);
    return Value;
}
inline float MaterialTexturableValue(const scene_view& S, uint Slot, uint A, vec2 UV) {
    float Value = MaterialFloat(S, Slot, A);
    uint Tex = MaterialUint(S, Slot, A + 1);
    if (Tex != TEXTURE_INDEX_NONE) Value *= SampleTexture(S, Tex, UV).x;
    return Value;
}

// Lexicographic "(ShapeIndex, PrimitiveIndex) is lower than the stored hit's".
inline bool lower_id(uint Shape, uint Prim, const hit& H) {
    return Shape < H.ShapeIndex || (Shape == H.ShapeIndex && Prim < H.PrimitiveIndex);
}

// scene.glsl.inc:304-334  Möller–Trumbore, two-sided.
inline void IntersectMeshFace(const scene_view& S, const ray& Ray, uint FaceIndex, uint ShapeIndex, hit& Hit, tie_mode Tie) {
    Hit.FaceTests++;
    const packed_mesh_face& F = S.MeshFaces[FaceIndex];
    vec3 Edge1 = F.Position1 - F.Position0;
    vec3 Edge2 = F.Position2 - F.Position0;
    vec3 RayCrossEdge2 = cross(Ray.Velocity, Edge2);
    float Det = dot(Edge1, RayCrossEdge2);
    if (fabsf(Det) < K_EPSILON) return;
    float InvDet = 1.0f / Det;
    vec3 Sv = Ray.Origin - F.Position0;
    float U = InvDet * dot(Sv, RayCrossEdge2);
    if (U < 0.0f || U > 1.0f) return;
    vec3 SCrossEdge1 = cross(Sv, Edge1);
    float V = InvDet * dot(Ray.Velocity, SCrossEdge1);
    if (V < 0.0f || U + V > 1.0f) return;
    float T = InvDet * dot(Edge2, SCrossEdge1);
    if (T < 0.0f || T > Hit.Time) return;
    if (Tie == TIE_CANONICAL && T == Hit.Time && !lower_id(ShapeIndex, FaceIndex, Hit)) return;
    Hit.Time = T;
    Hit.ShapeType = SHAPE_TYPE_MESH_INSTANCE;
    Hit.ShapeIndex = ShapeIndex;  // reference tags 0xFFFFFFFE and patches at :410-411; same result
    Hit.PrimitiveIndex = FaceIndex;
    Hit.PrimitiveCoordinates = V3((1.0f - U) - V, U, V);
}

// scene.glsl.inc:336-399  per-mesh BVH2, near-first, 32-entry stack (Q20: no
// overflow guard in the reference; the oracle aborts the path loudly instead).
inline void IntersectMeshNode(const scene_view& S, const ray& Ray, uint MeshNodeIndex, uint ShapeIndex, hit& Hit, tie_mode Tie) {
    uint Stack[64];
    uint Depth = 0;
    const bool strict = (Tie == TIE_CANONICAL);
    packed_mesh_node Node = S.MeshNodes[MeshNodeIndex];
    Hit.NodeFetches++;
    while (true) {
        Hit.MeshComplexity++;
        if (Node.FaceEndIndex > 0) {
            for (uint Fi = Node.FaceBeginOrNodeIndex; Fi < Node.FaceEndIndex; Fi++)
                IntersectMeshFace(S, Ray, Fi, ShapeIndex, Hit, Tie);
        } else {
            uint Index = Node.FaceBeginOrNodeIndex;
            Node = S.MeshNodes[Index];
            float Time = IntersectBoundingBox(Ray, Hit.Time, Node.Minimum, Node.Maximum, strict);
            uint IndexB = Index + 1;
            packed_mesh_node NodeB = S.MeshNodes[IndexB];
            float TimeB = IntersectBoundingBox(Ray, Hit.Time, NodeB.Minimum, NodeB.Maximum, strict);
            Hit.NodeFetches += 2;
            if (Time > TimeB) {
                if (Time < K_INFINITY) { if (Depth < 64) Stack[Depth++] = Index; }
                Node = NodeB;
                continue;
            }
            if (TimeB < K_INFINITY) {
                if (Depth < 64) Stack[Depth++] = IndexB;
                continue;
            }
            if (Time < K_INFINITY) continue;
        }
        if (Depth == 0) break;
        Node = S.MeshNodes[Stack[--Depth]];
        Hit.NodeFetches++;
    }
}

// scene.glsl.inc:401-466
inline void IntersectShape(const scene_view& S, const ray& WorldRay, uint ShapeIndex, hit& Hit, tie_mode Tie) {
    Hit.ShapeVisits++;
    const packed_shape& Shape = S.Shapes[ShapeIndex];
    ray Ray = InverseTransformRay(WorldRay, Shape.Transform);
    if (Shape.Type == SHAPE_TYPE_MESH_INSTANCE) {
        IntersectMeshNode(S, Ray, Shape.MeshRootNodeIndex, ShapeIndex, Hit, Tie);
    } else if (Shape.Type == SHAPE_TYPE_PLANE) {
        float T = -Ray.Origin.z / Ray.Velocity.z;
        if (T < 0.0f || T > Hit.Time) return;
        if (Tie == TIE_CANONICAL && T == Hit.Time && !lower_id(ShapeIndex, 0, Hit)) return;
        Hit.Time = T;
        Hit.ShapeType = SHAPE_TYPE_PLANE;
        Hit.ShapeIndex = ShapeIndex;
        Hit.PrimitiveIndex = 0;
        Hit.PrimitiveCoordinates = Ray.Origin + Ray.Velocity * T;
    } else if (Shape.Type == SHAPE_TYPE_SPHERE) {
        float V = dot(Ray.Velocity, Ray.Velocity);
        float P = dot(Ray.Origin, Ray.Velocity);
        float Q = dot(Ray.Origin, Ray.Origin) - 1.0f;
        float D2 = P * P - Q * V;
        if (D2 < 0.0f) return;
        float D = sqrtf(D2);
        if (D < P) return;
        float S0 = -P - D;
        float S1 = -P + D;
        float Sx = S0 < 0.0f ? S1 : S0;
        if (Sx < 0.0f || Sx > V * Hit.Time) return;
        float T = Sx / V;
        if (Tie == TIE_CANONICAL && T == Hit.Time && !lower_id(ShapeIndex, 0, Hit)) return;
        Hit.Time = T;
        Hit.ShapeType = SHAPE_TYPE_SPHERE;
        Hit.ShapeIndex = ShapeIndex;
        Hit.PrimitiveIndex = 0;
        Hit.PrimitiveCoordinates = Ray.Origin + Ray.Velocity * Hit.Time;
    } else if (Shape.Type == SHAPE_TYPE_CUBE) {
        vec3 Minimum = (V3(-1, -1, -1) - Ray.Origin) / Ray.Velocity;
        vec3 Maximum = (V3(+1, +1, +1) - Ray.Origin) / Ray.Velocity;
        vec3 Earlier = vmin(Minimum, Maximum);
        vec3 Later = vmax(Minimum, Maximum);
        float T0 = fmaxf(fmaxf(Earlier.x, Earlier.y), Earlier.z);
        float T1 = fminf(fminf(Later.x, Later.y), Later.z);
        if (T1 < T0) return;
        if (T1 <= 0.0f) return;
        float T = T0 < 0.0f ? T1 : T0;
        if (Tie == TIE_FAITHFUL) { if (T >= Hit.Time) return; }
        else { if (T > Hit.Time) return; if (T == Hit.Time && !lower_id(ShapeIndex, 0, Hit)) return; }
        Hit.Time = T;
        Hit.ShapeType = SHAPE_TYPE_CUBE;
        Hit.ShapeIndex = ShapeIndex;
        Hit.PrimitiveIndex = 0;
        Hit.PrimitiveCoordinates = Ray.Origin + Ray.Velocity * T;
    }
}

// scene.glsl.inc:468-520  top-level BVH2 over shapes (16-bit child indices, Q9).
inline void Intersect(const scene_view& S, const ray& Ray, hit& Hit, tie_mode Tie) {
    if (S.Globals->ShapeCount == 0) return;
    uint Stack[64];
    uint Depth = 0;
    const bool strict = (Tie == TIE_CANONICAL);
    packed_shape_node NodeA = S.ShapeNodes[0];
    packed_shape_node NodeB;
    Hit.NodeFetches++;
    while (true) {
        Hit.SceneComplexity++;
        if (NodeA.ChildNodeIndices == 0) {
            IntersectShape(S, Ray, NodeA.ShapeIndex, Hit, Tie);
        } else {
            uint IndexA = NodeA.ChildNodeIndices & 0xFFFFu;
            uint IndexB = NodeA.ChildNodeIndices >> 16;
            NodeA = S.ShapeNodes[IndexA];
            NodeB = S.ShapeNodes[IndexB];
            Hit.NodeFetches += 2;
            float TimeA = IntersectBoundingBox(Ray, Hit.Time, NodeA.Minimum, NodeA.Maximum, strict);
            float TimeB = IntersectBoundingBox(Ray, Hit.Time, NodeB.Minimum, NodeB.Maximum, strict);
            if (TimeA > TimeB) {
                if (TimeA < K_INFINITY) { if (Depth < 64) Stack[Depth++] = IndexA; }
                NodeA = NodeB;
                continue;
            }
            if (TimeB < K_INFINITY) {
                if (Depth < 64) Stack[Depth++] = IndexB;
                continue;
            }
            if (TimeA < K_INFINITY) continue;
        }
        if (Depth == 0) break;
        NodeA = S.ShapeNodes[Stack[--Depth]];
        Hit.NodeFetches++;
    }
}

// scene.glsl.inc:522-611
inline hit Trace(const scene_view& S, const ray& Ray, tie_mode Tie) {
    hit Hit;
    memset(&Hit, 0, sizeof(Hit));
    Hit.ShapeIndex = SHAPE_INDEX_NONE;
    Hit.PrimitiveIndex = 0xFFFFFFFFu;
    Hit.Time = Ray.Duration;
    Intersect(S, Ray, Hit, Tie);
    if (Hit.ShapeIndex == SHAPE_INDEX_NONE) return Hit;

    const packed_shape& Shape = S.Shapes[Hit.ShapeIndex];
    Hit.ShapeVisits++;
    Hit.MaterialIndex = Shape.MaterialIndex;

    if (Hit.ShapeType == SHAPE_TYPE_MESH_INSTANCE) {
        const packed_mesh_face& F = S.MeshFaces[Hit.PrimitiveIndex];
        const packed_mesh_vertex& V0 = S.MeshVertices[F.VertexIndex0];
        const packed_mesh_vertex& V1 = S.MeshVertices[F.VertexIndex1];
        const packed_mesh_vertex& V2_ = S.MeshVertices[F.VertexIndex2];
        vec3 C = Hit.PrimitiveCoordinates;
        vec3 Normal = SafeNormalize((UnpackUnitVector(V0.PackedNormal) * C.x + UnpackUnitVector(V1.PackedNormal) * C.y) +
                                    UnpackUnitVector(V2_.PackedNormal) * C.z);
        Hit.Normal = TransformNormal(Normal, Shape.Transform);
        Hit.TangentX = ComputeTangentVector(Hit.Normal);
        Hit.UV = (unpackHalf2x16(V0.PackedUV) * C.x + unpackHalf2x16(V1.PackedUV) * C.y) + unpackHalf2x16(V2_.PackedUV) * C.z;
    } else if (Hit.ShapeType == SHAPE_TYPE_PLANE) {
        Hit.Normal = TransformNormal(V3(0, 0, 1), Shape.Transform);
        Hit.TangentX = TransformDirection(V3(1, 0, 0), Shape.Transform);
        Hit.UV = V2(gl_fract(Hit.PrimitiveCoordinates.x), gl_fract(Hit.PrimitiveCoordinates.y));
    } else if (Hit.ShapeType == SHAPE_TYPE_SPHERE) {
        vec3 P = Hit.PrimitiveCoordinates;
        float U = (atan2f(P.y, P.x) + K_PI) / K_TAU;
        float V = (P.z + 1.0f) / 2.0f;
        Hit.Normal = TransformNormal(P, Shape.Transform);
        Hit.TangentX = TransformDirection(cross(P, V3(-P.y, P.x, 0.0f)), Shape.Transform);
        Hit.UV = V2(U, V);
    } else if (Hit.ShapeType == SHAPE_TYPE_CUBE) {
        vec3 P = Hit.PrimitiveCoordinates;
        vec3 Q = vabs(P);
        vec3 Normal, TangentX;
        if (Q.x >= Q.y && Q.x >= Q.z) {
            float Sg = gl_sign(P.x);
            Normal = V3(Sg, 0, 0); TangentX = V3(0, Sg, 0);
            Hit.UV = V2(0.5f * (1.0f + P.y), 0.5f * (1.0f + P.z));
        } else if (Q.y >= Q.x && Q.y >= Q.z) {
            float Sg = gl_sign(P.y);
            Normal = V3(0, Sg, 0); TangentX = V3(0, 0, Sg);
            Hit.UV = V2(0.5f * (1.0f + P.x), 0.5f * (1.0f + P.z));
        } else {
            float Sg = gl_sign(P.z);
            Normal = V3(0, 0, Sg); TangentX = V3(Sg, 0, 0);
            Hit.UV = V2(0.5f * (1.0f + P.x), 0.5f * (1.0f + P.y));
        }
        Hit.Normal = TransformNormal(Normal, Shape.Transform);
        Hit.TangentX = TransformDirection(TangentX, Shape.Transform);
    }
    return Hit;
}

// scene.glsl.inc:613-655
inline ray GenerateCameraRay(const packed_camera& Camera, vec2 NSP, rng& G) {
    ray Ray;
    Ray.Origin = V3(0, 0, 0);
    Ray.Velocity = V3(0, 0, 0);
    Ray.Duration = HIT_TIME_LIMIT;
    if (Camera.Model == CAMERA_MODEL_PINHOLE) {
        vec3 SensorPosition = V3(-Camera.SensorSize.x * (NSP.x - 0.5f), -Camera.SensorSize.y * (0.5f - NSP.y), Camera.SensorDistance);
        vec2 D = RandomPointOnDisk(G);
        Ray.Origin = V3(Camera.ApertureRadius * D.x, Camera.ApertureRadius * D.y, 0.0f);
        Ray.Velocity = normalize(Ray.Origin - SensorPosition);
    } else if (Camera.Model == CAMERA_MODEL_THIN_LENS) {
        vec3 SensorPosition = V3(-Camera.SensorSize.x * (NSP.x - 0.5f), -Camera.SensorSize.y * (0.5f - NSP.y), Camera.SensorDistance);
        vec3 ObjectPosition = ((-SensorPosition) * Camera.FocalLength) / (SensorPosition.z - Camera.FocalLength);
        vec2 D = RandomPointOnDisk(G);
        Ray.Origin = V3(Camera.ApertureRadius * D.x, Camera.ApertureRadius * D.y, 0.0f);
        Ray.Velocity = normalize(ObjectPosition - Ray.Origin);
    } else if (Camera.Model == CAMERA_MODEL_360) {
        float Phi = (NSP.x - 0.5f) * K_TAU;
        float Theta = (0.5f - NSP.y) * K_PI;
        Ray.Origin = V3(0, 0, 0);
        Ray.Velocity = V3(cosf(Theta) * sinf(Phi), sinf(Theta), -cosf(Theta) * cosf(Phi));
    }
    return TransformRay(Ray, Camera.Transform);
}

}  // namespace orc
