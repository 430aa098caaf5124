// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// Byte layouts of the reference's packed scene arrays (std430), restated from
// /root/reference/src/scene/scene.hpp:82-173 and scene.glsl.inc:30-99.
// These are the *input contract* of the hot path; the static_asserts pin the
// sizes/offsets listed in SURVEY.md Appendix A.
#pragma once
#include <cstddef>
#include "glsl_shim.h"

namespace orc {

// scene.hpp:82-86 / common.glsl.inc:17-21
struct packed_transform { mat4 To; mat4 From; };
static_assert(sizeof(packed_transform) == 128, "packed_transform");

// scene.hpp:90-98 / scene.glsl.inc:30-38
struct packed_texture {
    vec2 AtlasPlacementMinimum;
    vec2 AtlasPlacementMaximum;
    uint AtlasImageIndex;
    uint Type;
    uint Flags;
    uint Unused0;
};
static_assert(sizeof(packed_texture) == 32, "packed_texture");

// scene.hpp:102-108 / scene.glsl.inc:40-46
struct packed_shape {
    uint Type;
    uint MaterialIndex;
    uint MeshRootNodeIndex;
    uint Pad0;
    packed_transform Transform;
};
static_assert(sizeof(packed_shape) == 144 && offsetof(packed_shape, Transform) == 16, "packed_shape");

// scene.hpp:112-118 / scene.glsl.inc:48-54
struct packed_shape_node {
    vec3 Minimum; uint ChildNodeIndices;
    vec3 Maximum; uint ShapeIndex;
};
static_assert(sizeof(packed_shape_node) == 32, "packed_shape_node");

// scene.hpp:122-130 / scene.glsl.inc:56-64
struct packed_mesh_face {
    vec3 Position0; uint VertexIndex0;
    vec3 Position1; uint VertexIndex1;
    vec3 Position2; uint VertexIndex2;
};
static_assert(sizeof(packed_mesh_face) == 48, "packed_mesh_face");

// scene.hpp:134-138 / scene.glsl.inc:66-70
struct packed_mesh_vertex { uint PackedNormal; uint PackedUV; };
static_assert(sizeof(packed_mesh_vertex) == 8, "packed_mesh_vertex");

// scene.hpp:142-148 / scene.glsl.inc:72-78
struct packed_mesh_node {
    vec3 Minimum; uint FaceBeginOrNodeIndex;
    vec3 Maximum; uint FaceEndIndex;
};
static_assert(sizeof(packed_mesh_node) == 32, "packed_mesh_node");

// scene.hpp:152-161 / scene.glsl.inc:90-99
struct packed_scene_globals {
    vec3  SkyboxMeanDirection;
    float SkyboxConcentration;
    float SkyboxSamplingProbability;
    float SkyboxBrightness;
    uint  SkyboxTextureIndex;
    uint  ShapeCount;
    float SceneScatterRate;
    uint  Pad[3];
};
static_assert(sizeof(packed_scene_globals) == 48 && offsetof(packed_scene_globals, SceneScatterRate) == 32, "globals");

// scene.hpp:165-173 / scene.glsl.inc:80-88
struct packed_camera {
    uint  Model;
    float FocalLength;
    float ApertureRadius;
    float SensorDistance;
    vec2  SensorSize;
    uint  Pad[2];
    packed_transform Transform;
};
static_assert(sizeof(packed_camera) == 160 && offsetof(packed_camera, Transform) == 32, "packed_camera");

// What the shaders see bound at set=BIND_SCENE (scene.glsl.inc:121-179) as plain
// arrays.  Field order/meaning equals include/ptb.h's ptb_scene_view, which is
// the same bytes viewed through the product's own declarations.
struct scene_view {
    const packed_texture*     Textures;      uint TextureCount;    uint _p0;
    const uint*               MaterialData;  uint MaterialWordCount; uint _p1;
    const packed_shape*       Shapes;        uint ShapeCount;      uint _p2;
    const packed_shape_node*  ShapeNodes;    uint ShapeNodeCount;  uint _p3;
    const packed_mesh_face*   MeshFaces;     uint MeshFaceCount;   uint _p4;
    const packed_mesh_vertex* MeshVertices;  uint MeshVertexCount; uint _p5;
    const packed_mesh_node*   MeshNodes;     uint MeshNodeCount;   uint _p6;
    const packed_camera*      Cameras;       uint CameraCount;     uint _p7;
    const float* const*       AtlasLayers;   uint AtlasLayerCount; // each layer AtlasWidth*AtlasHeight RGBA32F
    uint                      AtlasWidth;
    uint                      AtlasHeight;   uint _p8;
    const packed_scene_globals* Globals;
};

}  // namespace orc
