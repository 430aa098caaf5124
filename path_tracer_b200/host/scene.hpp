// Host-side scene layer: the source description (entities, meshes, materials,
// textures) and PackSceneData(), which flattens it into the packed arrays that
// are the input contract of the hot path (include/ptb.h: ptb_scene_view).
//
// Mirrors the operator-level interface of /root/reference/src/scene/scene.hpp
// :410-442 — same function names, argument meaning and dirty-flag protocol — so
// tests and the bench drive the integrator the way the reference application
// does.  It is NOT part of the C ABI: in a drop-in the reference keeps its own
// scene.cpp and hands its packed vectors to ptb_scene_update().
#pragma once
#include <string>
#include <vector>
#include "../../include/ptb.h"
#include "hmath.hpp"
#include "spectrum.hpp"

namespace ptb { namespace host {

constexpr uint32_t SHAPE_INDEX_NONE = 0xFFFFFFFFu;     // scene.hpp:7
constexpr uint32_t TEXTURE_INDEX_NONE = 0xFFFFFFFFu;   // scene.hpp:8

enum texture_type { TEXTURE_TYPE_RAW = 0, TEXTURE_TYPE_REFLECTANCE_WITH_ALPHA = 1, TEXTURE_TYPE_RADIANCE = 2 };   // scene.hpp:10-16
enum texture_flag : uint32_t { TEXTURE_FLAG_FILTER_NEAREST = 1u << 0 };                                           // scene.hpp:18-21
enum material_type { MATERIAL_TYPE_BASIC_DIFFUSE = 0, MATERIAL_TYPE_BASIC_METAL = 1,
                     MATERIAL_TYPE_BASIC_TRANSLUCENT = 2, MATERIAL_TYPE_OPENPBR = 3 };                            // scene.hpp:23-30
enum shape_type : int32_t { SHAPE_TYPE_MESH_INSTANCE = 0, SHAPE_TYPE_PLANE = 1, SHAPE_TYPE_SPHERE = 2, SHAPE_TYPE_CUBE = 3 };  // :32-38
enum camera_model : int32_t { CAMERA_MODEL_PINHOLE = 0, CAMERA_MODEL_THIN_LENS = 1, CAMERA_MODEL_360 = 2 };       // scene.hpp:40-46
enum entity_type { ENTITY_TYPE_ROOT = 0, ENTITY_TYPE_CONTAINER = 1, ENTITY_TYPE_CAMERA = 2, ENTITY_TYPE_MESH_INSTANCE = 3,
                   ENTITY_TYPE_PLANE = 4, ENTITY_TYPE_SPHERE = 5, ENTITY_TYPE_CUBE = 6 };                         // scene.hpp:229-240
enum scene_dirty_flag : uint32_t {                                                                                // scene.hpp:323-333
    SCENE_DIRTY_GLOBALS = 1u << 0, SCENE_DIRTY_TEXTURES = 1u << 1, SCENE_DIRTY_MATERIALS = 1u << 2,
    SCENE_DIRTY_SHAPES = 1u << 3, SCENE_DIRTY_MESHES = 1u << 4, SCENE_DIRTY_CAMERAS = 1u << 5,
    SCENE_DIRTY_SKYBOX_TEXTURE = 1u << 6, SCENE_DIRTY_ALL = 0xFFFFFFFFu,
};

struct transform {                 // common.hpp:48-54
    vec3 Position{0, 0, 0};
    vec3 Rotation{0, 0, 0};
    vec3 Scale{1, 1, 1};
};

struct bounds {                    // common.hpp:56-60
    vec3 Minimum{+INFINITY, +INFINITY, +INFINITY};
    vec3 Maximum{-INFINITY, -INFINITY, -INFINITY};
};

struct texture {                   // scene.hpp:177-186
    std::string Name = "New Texture";
    texture_type Type = TEXTURE_TYPE_RAW;
    bool EnableNearestFiltering = false;
    uint32_t Width = 0, Height = 0;
    std::vector<vec4> Pixels;
    uint32_t PackedTextureIndex = 0;
};

// scene.hpp:188-197 plus the three active material structs
// (basic_diffuse.hpp:3-9, basic_metal.hpp:3-15, basic_translucent.hpp:3-17) as one tagged record.
struct material {
    material_type Type = MATERIAL_TYPE_BASIC_DIFFUSE;
    std::string Name = "New Material";
    uint32_t PackedMaterialIndex = 0;
    // diffuse + metal
    vec3 BaseColor{1, 1, 1};
    texture* BaseTexture = nullptr;
    // metal
    vec3 SpecularColor{1, 1, 1};
    texture* SpecularTexture = nullptr;
    // metal + translucent
    float Roughness = 0.3f;
    texture* RoughnessTexture = nullptr;
    float RoughnessAnisotropy = 0.0f;
    texture* RoughnessAnisotropyTexture = nullptr;
    // translucent
    float IOR = 1.5f;
    float AbbeNumber = 20.0f;
    vec3 TransmissionColor{1, 1, 1};
    float TransmissionDepth = 0.0f;
    vec3 ScatteringColor{1, 1, 1};
    float ScatteringAnisotropy = 0.0f;
};

struct mesh_face { uint32_t VertexIndex[3]; };                      // scene.hpp:199-202
struct mesh_vertex { vec3 Position; vec3 Normal; vec2 UV; };        // scene.hpp:204-209
struct mesh_node { bounds Bounds; uint32_t FaceBeginIndex = 0, FaceEndIndex = 0, ChildNodeIndex = 0; };  // scene.hpp:211-217
struct mesh {                                                       // scene.hpp:219-227
    std::string Name;
    std::vector<mesh_vertex> Vertices;
    std::vector<mesh_face> Faces;
    std::vector<mesh_node> Nodes;
    uint32_t Depth = 0;
    uint32_t PackedRootNodeIndex = 0;
};

struct camera_pinhole { float FieldOfViewInDegrees = 90.0f; float ApertureDiameterInMM = 0.0f; };           // scene.hpp:271-275
struct camera_thin_lens { vec2 SensorSizeInMM{32.0f, 18.0f}; float FocalLengthInMM = 20.0f;
                          float ApertureDiameterInMM = 10.0f; float FocusDistance = 1.0f; };                 // scene.hpp:277-283

// scene.hpp:242-316 as one tagged record.
struct entity {
    std::string Name = "Entity";
    entity_type Type = ENTITY_TYPE_ROOT;
    bool Active = true;
    transform Transform;
    entity* Parent = nullptr;
    std::vector<entity*> Children;
    material* Material = nullptr;
    uint32_t PackedShapeIndex = SHAPE_INDEX_NONE;
    // root (scene.hpp:256-264)
    float ScatterRate = 0.0f;
    float SkyboxBrightness = 1.0f;
    float SkyboxSamplingProbability = 0.0f;
    texture* SkyboxTexture = nullptr;
    // camera (scene.hpp:285-294)
    camera_model CameraModel = CAMERA_MODEL_PINHOLE;
    camera_pinhole Pinhole;
    camera_thin_lens ThinLens;
    uint32_t PackedCameraIndex = 0;
    // mesh instance (scene.hpp:296-301)
    mesh* Mesh = nullptr;
};

struct prefab { entity* Entity = nullptr; };            // scene.hpp:318-321

struct load_model_options {        // scene.hpp:381-389
    char const* Name = nullptr;
    material*   DefaultMaterial = nullptr;
    std::string DirectoryPath;                          // empty: the directory of the model file
    mat4        VertexTransform = mat4::identity();
    mat4        NormalTransform = mat4::identity();
    float       TextureCoordinateTransform[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};   // column-major mat3
};

struct scene {                     // scene.hpp:335-362
    entity Root;
    std::vector<prefab*> Prefabs;
    std::vector<mesh*> Meshes;
    std::vector<material*> Materials;
    std::vector<texture*> Textures;
    parametric_spectrum_table const* RGBSpectrumTable = nullptr;   // shared, not owned

    uint32_t AtlasWidth = 4096, AtlasHeight = 4096;                // scene.cpp:1122-1123 (configurable here)
    std::vector<std::vector<float>> Images;                        // atlas layers, RGBA32F
    std::vector<ptb_packed_texture> TexturePack;
    std::vector<ptb_packed_shape> ShapePack;
    std::vector<ptb_packed_shape_node> ShapeNodePack;
    std::vector<uint32_t> MaterialAttributePack;
    std::vector<ptb_packed_mesh_face> MeshFacePack;
    std::vector<ptb_packed_mesh_vertex> MeshVertexPack;
    std::vector<ptb_packed_mesh_node> MeshNodePack;
    std::vector<ptb_packed_camera> CameraPack;
    ptb_packed_scene_globals Globals;

    uint32_t DirtyFlags = SCENE_DIRTY_ALL;

    // view over the packs (valid until the next PackSceneData / scene edit)
    std::vector<const float*> ImagePointers;
    ptb_scene_view View() ;
};

// ---- scene.hpp:410-442 ----
entity* CreateEntity(scene* Scene, entity_type Type, entity* Parent = nullptr);
void DestroyEntity(scene* Scene, entity* Entity);
material* CreateMaterial(scene* Scene, material_type Type, char const* Name);
texture* CreateCheckerTexture(scene* Scene, char const* Name, texture_type Type, vec4 const& ColorA, vec4 const& ColorB);
// Programmatic twin of LoadTexture (scene.cpp:294-318): pixels are linear RGBA32F, row 0 first.
texture* CreateTexture(scene* Scene, char const* Name, texture_type Type, uint32_t Width, uint32_t Height, vec4 const* Pixels);
// Programmatic twin of the tail of LoadModelAsPrefab (scene.cpp:851-866): takes
// ownership of the vertex/face lists, builds the mesh BVH2 (BuildMeshNode).
mesh* CreateMesh(scene* Scene, char const* Name, std::vector<mesh_vertex>&& Vertices, std::vector<mesh_face>&& Faces);
// CreateMesh in two steps, so that independent meshes can be built on several threads and attached in order.
mesh* BuildMesh(char const* Name, std::vector<mesh_vertex>&& Vertices, std::vector<mesh_face>&& Faces);
mesh* AttachMesh(scene* Scene, mesh* M);
// Model import and prefabs (scene.hpp:415,428-429; model.cpp).  Wavefront OBJ + MTL; textures through DecodeImageFile.
// DecodeImageFile (image.cpp) = stbi_loadf(Path, &W, &H, &N, 4) for PNG, JPEG and Radiance HDR (plus PPM / PFM).
bool DecodeImageFile(char const* Path, int& Width, int& Height, std::vector<vec4>& Pixels);
texture* LoadTexture(scene* Scene, char const* Path, texture_type Type, char const* Name = nullptr);
prefab* LoadModelAsPrefab(scene* Scene, char const* Path, load_model_options* Options = nullptr);
entity* CreateEntity(scene* Scene, prefab* Prefab, entity* Parent = nullptr);
void DestroyPrefab(scene* Scene, prefab* Prefab);
// Image files for the resolved film (the reference presents to a swapchain and has no writer).
bool WriteImagePNG(char const* Path, uint32_t Width, uint32_t Height, uint8_t const* RGBA);
bool WriteImagePFM(char const* Path, uint32_t Width, uint32_t Height, float const* RGB);
// Scene files (scene.hpp:432-433; serializer.cpp): the reference's JSON + zlib side-file format.
scene* LoadScene(char const* Path, std::string* Error = nullptr);
bool SaveScene(char const* Path, scene* Scene, std::string* Error = nullptr);
// CreateScene (scene.cpp:912-943).  WithDefaults adds the reference's default
// checker plane + camera.  The RGB->spectrum table is process-wide (see below).
scene* CreateScene(bool WithDefaults = true);
void DestroyScene(scene* Scene);
uint32_t PackSceneData(scene* Scene);
entity* FindEntityByPackedShapeIndex(scene* Scene, uint32_t PackedShapeIndex);

// Process-wide sRGB spectrum table: loaded from `Path` (reference raw format) or
// built with all host threads and saved there (scene.cpp:918-927).
parametric_spectrum_table const* SharedSpectrumTable(char const* Path = nullptr);
void SetSpectrumTablePath(char const* Path);

}}  // namespace ptb::host
