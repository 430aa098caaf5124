// ORACLE - TEST INFRASTRUCTURE ONLY.
//
// oracle/_ref/libref_host.so: the reference's OWN host-side scene code compiled where it lies -
//   src/scene/scene.cpp up to its Vulkan half (CreateScene, CreateEntity, CreateMaterial, CreateCheckerTexture,
//   LoadTexture, LoadModelAsPrefab, BuildMeshNode, PackSceneData: texture atlas with stb_rect_pack, material
//   attribute slots, mesh packs, shape BVH, cameras, globals), src/core/spectrum.cpp (table fit + lookup),
//   the material packers (scene/basic_*.hpp), and the vendored src/core/stb_image.h, stb_rect_pack.h,
//   tiny_obj_loader.h - against oracle/ref/hostshim (a stand-in for the un-vendored glm 1.0.1 and a stub for
//   core/vulkan.hpp).  Nothing of the reference is copied: oracle/ref/hostgen.py writes the cut scene.cpp into the
//   git-ignored oracle/_ref/gen/, everything else is #included from /root/reference/src.
//
// The exported C functions carry the names and signatures of path_tracer_b200/host/capi.cpp (ptbh_*), so that the
// Python Scene class can drive the reference's scene API and this repository's restatement of it with the same
// calls; tests/test_reference_host.py then compares every packed array byte for byte and the
// image decoders against stbi_loadf (row N4).
#include <cassert>
#include <cstdio>
#include <cstring>

// MSVC-only calls used by the reference (spectrum.cpp:429,442)
static inline int fopen_s(FILE** f, const char* path, const char* mode) { *f = std::fopen(path, mode); return *f ? 0 : 1; }

#include "scene_host.cpp"                 // generated: reference src/scene/scene.cpp, non-Vulkan half
#include "core/spectrum.cpp"              // reference, where it lies
#define STB_IMAGE_IMPLEMENTATION
#include "core/stb_image.h"
#define STB_RECT_PACK_IMPLEMENTATION
#include "core/stb_rect_pack.h"
#define TINYOBJLOADER_IMPLEMENTATION
#include "core/tiny_obj_loader.h"
// LoadScene / SaveScene come from ref_serializer.cpp (the reference's serializer.cpp wants scene.hpp with
// SERIALIZER_IMPLEMENTATION defined: its own translation unit), miniz from the reference's miniz.c compiled as C.

#include "ptb.h"

static_assert(sizeof(packed_shape) == sizeof(ptb_packed_shape), "packed_shape");
static_assert(sizeof(packed_shape_node) == sizeof(ptb_packed_shape_node), "packed_shape_node");
static_assert(sizeof(packed_mesh_face) == sizeof(ptb_packed_mesh_face), "packed_mesh_face");
static_assert(sizeof(packed_mesh_vertex) == sizeof(ptb_packed_mesh_vertex), "packed_mesh_vertex");
static_assert(sizeof(packed_mesh_node) == sizeof(ptb_packed_mesh_node), "packed_mesh_node");
static_assert(sizeof(packed_texture) == sizeof(ptb_packed_texture), "packed_texture");
static_assert(sizeof(packed_camera) == sizeof(ptb_packed_camera), "packed_camera");
static_assert(sizeof(packed_scene_globals) == sizeof(ptb_packed_scene_globals), "packed_scene_globals");

namespace {
parametric_spectrum_table* g_table = nullptr;
std::string g_table_path;
struct view_store { std::vector<const float*> layers; };
std::unordered_map<scene*, view_store> g_views;

// CreateScene (scene.cpp:912-943) minus the file lookup in the working directory: the table is the shared one
// (loaded from the path given, in the reference's raw format), the default plane + camera only on request.
scene* create_scene(bool with_defaults) {
    if (!g_table) {
        g_table = new parametric_spectrum_table;
        if (g_table_path.empty() || !LoadParametricSpectrumTable(g_table, g_table_path.c_str())) { delete g_table; g_table = nullptr; return nullptr; }
    }
    auto Scene = new scene;
    Scene->Root.Name = "Scene";
    Scene->RGBSpectrumTable = g_table;
    if (with_defaults) {
        auto PlaneMaterial = (basic_diffuse_material*)CreateMaterial(Scene, MATERIAL_TYPE_BASIC_DIFFUSE, "Plane Material");
        PlaneMaterial->BaseTexture = CreateCheckerTexture(Scene, "Plane Texture", TEXTURE_TYPE_REFLECTANCE_WITH_ALPHA, glm::vec4(1, 1, 1, 1), glm::vec4(0.5, 0.5, 0.5, 1));
        PlaneMaterial->BaseTexture->EnableNearestFiltering = true;
        plane_entity* Plane = (plane_entity*)CreateEntity(Scene, ENTITY_TYPE_PLANE);
        Plane->Name = "Plane";
        Plane->Material = PlaneMaterial;
        camera_entity* Camera = (camera_entity*)CreateEntity(Scene, ENTITY_TYPE_CAMERA);
        Camera->Name = "Camera";
        Camera->Transform.Position = {0, 0, 1};
    }
    Scene->DirtyFlags = SCENE_DIRTY_ALL;
    return Scene;
}
}  // namespace

extern "C" {

void ptbh_set_spectrum_table_path(const char* path) { g_table_path = path; }
int ptbh_ensure_spectrum_table() {
    if (g_table) return 1;
    scene* s = create_scene(false);
    if (!s) return 0;
    delete s;
    return 1;
}
void ptbh_rgb_to_spectrum(const float* rgb, float* beta) {
    vec3 b = GetParametricSpectrumCoefficients(g_table, vec3(rgb[0], rgb[1], rgb[2]));
    beta[0] = b.x; beta[1] = b.y; beta[2] = b.z;
}
// One column of BuildParametricSpectrumTableForSRGB (spectrum.cpp:365-411): the 64 warm-started fits of (I, J, L),
// written as 64 x 3 floats in K order.  (The whole table takes the reference minutes; tests sample columns.)
void refh_spectrum_column(int I, int J, int L, float* out) {
    constexpr int M = parametric_spectrum_table::SCALE_BINS;
    auto Denormalize = [](dvec3 const& B) -> vec3 {
        constexpr float C0 = CIE_LAMBDA_MIN;
        constexpr float C1 = 1.f / (CIE_LAMBDA_MAX - CIE_LAMBDA_MIN);
        return { B[0] * C1 * C1, B[1] * C1 - 2 * B[0] * C0 * C1 * C1, B[2] - B[1] * C0 * C1 + B[0] * C0 * C0 * C1 * C1 };
    };
    dvec3 NormalizedBeta = {};
    for (int K = M / 5; K < M; K++) {
        NormalizedBeta = OptimizeSpectrum(NormalizedBeta, SRGBToXYZ(IndexToColor(I, J, K, L)), 15);
        vec3 b = Denormalize(NormalizedBeta);
        out[3 * K] = b.x; out[3 * K + 1] = b.y; out[3 * K + 2] = b.z;
    }
    NormalizedBeta = {};
    for (int K = M / 5; K >= 0; K--) {
        NormalizedBeta = OptimizeSpectrum(NormalizedBeta, SRGBToXYZ(IndexToColor(I, J, K, L)), 15);
        vec3 b = Denormalize(NormalizedBeta);
        out[3 * K] = b.x; out[3 * K + 1] = b.y; out[3 * K + 2] = b.z;
    }
}
// CIELAB distance between an sRGB colour and what the spectrum `beta` (denormalised coefficients) looks like under D65.
double refh_spectrum_lab_error(const float* rgb, const float* beta) {
    const double C0 = CIE_LAMBDA_MIN, C1 = 1.0 / (CIE_LAMBDA_MAX - CIE_LAMBDA_MIN);
    // invert DenormalizeBeta: b0 = B0 C1^2 ; b1 = B1 C1 - 2 B0 C0 C1^2 ; b2 = B2 - B1 C0 C1 + B0 C0^2 C1^2
    const double B0 = beta[0] / (C1 * C1);
    const double B1 = (beta[1] + 2 * B0 * C0 * C1 * C1) / C1;
    const double B2 = beta[2] + B1 * C0 * C1 - B0 * C0 * C0 * C1 * C1;
    dvec3 lab0 = XYZToLab(ObserveSpectrumUnderD65(dvec3(B0, B1, B2)));
    dvec3 lab1 = XYZToLab(dvec3(SRGBToXYZ(vec3(rgb[0], rgb[1], rgb[2]))));
    return glm::length(lab0 - lab1);
}

// LoadScene / SaveScene of the reference (serializer.cpp:501-529).  A loaded scene brings its own spectrum table (spectrum.dat
// beside the scene file); everything else here shares g_table, so the loaded copy replaces nothing and is simply kept.
void* ptbh_load_scene(const char* path) {
    try { return LoadScene(path); } catch (...) { return nullptr; }
}
int ptbh_save_scene(void* s, const char* path) {
    try { SaveScene(path, (scene*)s); return 0; } catch (...) { return -1; }
}

void* ptbh_scene_create(int with_defaults) { return create_scene(with_defaults != 0); }
// The table is shared, so not DestroyScene; but the atlas layers (256 MiB each, never freed by the reference: scene.cpp:1140
// clears the vector of a previous pack without deleting) go, or a fuzz over hundreds of scenes runs out of memory.
void ptbh_scene_destroy(void* s) {
    scene* S = (scene*)s;
    g_views.erase(S);
    for (image const& I : S->Images) delete[] (glm::vec4*)I.Pixels;
    S->Images.clear();
}
uint32_t ptbh_scene_pack(void* s) { return PackSceneData((scene*)s); }
void ptbh_scene_mark_dirty(void* s, uint32_t flags) { ((scene*)s)->DirtyFlags |= flags; }
void ptbh_scene_view(void* s, ptb_scene_view* out) {
    scene* S = (scene*)s;
    view_store& vs = g_views[S];
    vs.layers.clear();
    for (image const& I : S->Images) vs.layers.push_back((const float*)I.Pixels);
    std::memset(out, 0, sizeof *out);
    out->Textures = (const ptb_packed_texture*)S->TexturePack.data(); out->TextureCount = (uint32_t)S->TexturePack.size();
    out->MaterialData = S->MaterialAttributePack.data(); out->MaterialWordCount = (uint32_t)S->MaterialAttributePack.size();
    out->Shapes = (const ptb_packed_shape*)S->ShapePack.data(); out->ShapeCount = (uint32_t)S->ShapePack.size();
    out->ShapeNodes = (const ptb_packed_shape_node*)S->ShapeNodePack.data(); out->ShapeNodeCount = (uint32_t)S->ShapeNodePack.size();
    out->MeshFaces = (const ptb_packed_mesh_face*)S->MeshFacePack.data(); out->MeshFaceCount = (uint32_t)S->MeshFacePack.size();
    out->MeshVertices = (const ptb_packed_mesh_vertex*)S->MeshVertexPack.data(); out->MeshVertexCount = (uint32_t)S->MeshVertexPack.size();
    out->MeshNodes = (const ptb_packed_mesh_node*)S->MeshNodePack.data(); out->MeshNodeCount = (uint32_t)S->MeshNodePack.size();
    out->Cameras = (const ptb_packed_camera*)S->CameraPack.data(); out->CameraCount = (uint32_t)S->CameraPack.size();
    out->AtlasLayers = vs.layers.data(); out->AtlasLayerCount = (uint32_t)vs.layers.size();
    out->AtlasWidth = 4096; out->AtlasHeight = 4096;                       // scene.cpp:1122-1123
    out->Globals = (const ptb_packed_scene_globals*)&S->Globals;
}
void ptbh_root_set(void* s, float scatter_rate, float sky_brightness, float sky_sampling_probability, void* sky_texture) {
    scene* S = (scene*)s;
    S->Root.ScatterRate = scatter_rate; S->Root.SkyboxBrightness = sky_brightness;
    S->Root.SkyboxSamplingProbability = sky_sampling_probability; S->Root.SkyboxTexture = (texture*)sky_texture;
    S->DirtyFlags |= SCENE_DIRTY_GLOBALS | SCENE_DIRTY_SKYBOX_TEXTURE;
}

void* ptbh_create_entity(void* s, int type, void* parent) { return CreateEntity((scene*)s, (entity_type)type, (entity*)parent); }
// The reference's DestroyEntity leaves the dirty flags to its caller (imgui_main.cpp:525 sets SCENE_DIRTY_SHAPES); the product's sets
// shapes and cameras itself, and so does this wrapper.
void ptbh_destroy_entity(void* s, void* e) { DestroyEntity((scene*)s, (entity*)e); ((scene*)s)->DirtyFlags |= SCENE_DIRTY_SHAPES | SCENE_DIRTY_CAMERAS; }
void ptbh_entity_set_transform(void* s, void* e, const float* pos, const float* rot, const float* scale) {
    entity* E = (entity*)e;
    E->Transform.Position = {pos[0], pos[1], pos[2]};
    E->Transform.Rotation = {rot[0], rot[1], rot[2]};
    E->Transform.Scale = {scale[0], scale[1], scale[2]};
    ((scene*)s)->DirtyFlags |= SCENE_DIRTY_SHAPES | SCENE_DIRTY_CAMERAS;
}
void ptbh_entity_set_active(void* s, void* e, int active) { ((entity*)e)->Active = active != 0; ((scene*)s)->DirtyFlags |= SCENE_DIRTY_SHAPES | SCENE_DIRTY_CAMERAS; }
void ptbh_entity_set_material(void* s, void* e, void* m) { ((entity*)e)->Material = (material*)m; ((scene*)s)->DirtyFlags |= SCENE_DIRTY_SHAPES; }
void ptbh_entity_set_mesh(void* s, void* e, void* m) { ((mesh_entity*)(entity*)e)->Mesh = (mesh*)m; ((scene*)s)->DirtyFlags |= SCENE_DIRTY_SHAPES; }
uint32_t ptbh_entity_shape_index(void* e) { return ((entity*)e)->PackedShapeIndex; }
uint32_t ptbh_entity_camera_index(void* e) { return ((camera_entity*)(entity*)e)->PackedCameraIndex; }
void ptbh_camera_set(void* s, void* e, int model, const float* p) {
    camera_entity* E = (camera_entity*)(entity*)e;
    E->CameraModel = (camera_model)model;
    if (model == CAMERA_MODEL_PINHOLE) { E->Pinhole.FieldOfViewInDegrees = p[0]; E->Pinhole.ApertureDiameterInMM = p[1]; }
    if (model == CAMERA_MODEL_THIN_LENS) { E->ThinLens.SensorSizeInMM = {p[0], p[1]}; E->ThinLens.FocalLengthInMM = p[2]; E->ThinLens.ApertureDiameterInMM = p[3]; E->ThinLens.FocusDistance = p[4]; }
    ((scene*)s)->DirtyFlags |= SCENE_DIRTY_CAMERAS;
}

void* ptbh_create_material(void* s, int type, const char* name) { return CreateMaterial((scene*)s, (material_type)type, name); }
int ptbh_material_set(void* s, void* m, const char* field, const float* v) {
    material* M = (material*)m;
    ((scene*)s)->DirtyFlags |= SCENE_DIRTY_MATERIALS;
    auto is = [&](const char* n) { return !std::strcmp(field, n); };
    vec3 c = {v[0], is("BaseColor") || is("SpecularColor") || is("TransmissionColor") || is("ScatteringColor") ? v[1] : 0.0f,
              is("BaseColor") || is("SpecularColor") || is("TransmissionColor") || is("ScatteringColor") ? v[2] : 0.0f};
    switch (M->Type) {
        case MATERIAL_TYPE_BASIC_DIFFUSE: {
            auto D = (basic_diffuse_material*)M;
            if (is("BaseColor")) { D->BaseColor = c; return 0; }
            return -1;
        }
        case MATERIAL_TYPE_BASIC_METAL: {
            auto D = (basic_metal_material*)M;
            if (is("BaseColor")) { D->BaseColor = c; return 0; }
            if (is("SpecularColor")) { D->SpecularColor = c; return 0; }
            if (is("Roughness")) { D->Roughness = v[0]; return 0; }
            if (is("RoughnessAnisotropy")) { D->RoughnessAnisotropy = v[0]; return 0; }
            return -1;
        }
        case MATERIAL_TYPE_BASIC_TRANSLUCENT: {
            auto D = (basic_translucent_material*)M;
            if (is("IOR")) { D->IOR = v[0]; return 0; }
            if (is("AbbeNumber")) { D->AbbeNumber = v[0]; return 0; }
            if (is("Roughness")) { D->Roughness = v[0]; return 0; }
            if (is("RoughnessAnisotropy")) { D->RoughnessAnisotropy = v[0]; return 0; }
            if (is("TransmissionColor")) { D->TransmissionColor = c; return 0; }
            if (is("TransmissionDepth")) { D->TransmissionDepth = v[0]; return 0; }
            if (is("ScatteringColor")) { D->ScatteringColor = c; return 0; }
            if (is("ScatteringAnisotropy")) { D->ScatteringAnisotropy = v[0]; return 0; }
            return -1;
        }
        default: return -1;
    }
}
int ptbh_material_set_texture(void* s, void* m, const char* field, void* t) {
    material* M = (material*)m; texture* T = (texture*)t;
    ((scene*)s)->DirtyFlags |= SCENE_DIRTY_MATERIALS;
    auto is = [&](const char* n) { return !std::strcmp(field, n); };
    if (M->Type == MATERIAL_TYPE_BASIC_DIFFUSE && is("BaseTexture")) { ((basic_diffuse_material*)M)->BaseTexture = T; return 0; }
    if (M->Type == MATERIAL_TYPE_BASIC_METAL) {
        auto D = (basic_metal_material*)M;
        if (is("BaseTexture")) { D->BaseTexture = T; return 0; }
        if (is("SpecularTexture")) { D->SpecularTexture = T; return 0; }
        if (is("RoughnessTexture")) { D->RoughnessTexture = T; return 0; }
        if (is("RoughnessAnisotropyTexture")) { D->RoughnessAnisotropyTexture = T; return 0; }
    }
    if (M->Type == MATERIAL_TYPE_BASIC_TRANSLUCENT) {
        auto D = (basic_translucent_material*)M;
        if (is("RoughnessTexture")) { D->RoughnessTexture = T; return 0; }
        if (is("RoughnessAnisotropyTexture")) { D->RoughnessAnisotropyTexture = T; return 0; }
    }
    return -1;
}
uint32_t ptbh_material_slot(void* m) { return ((material*)m)->PackedMaterialIndex; }

// The reference has no programmatic texture constructor besides the checker: this is the tail of LoadTexture
// (scene.cpp:299-312) on caller pixels.
void* ptbh_create_texture(void* s, const char* name, int type, uint32_t w, uint32_t h, const float* rgba, int nearest) {
    scene* Scene = (scene*)s;
    glm::vec4* Pixels = (glm::vec4*)std::malloc(sizeof(glm::vec4) * w * h);
    std::memcpy((void*)Pixels, rgba, sizeof(glm::vec4) * w * h);
    auto Texture = new texture{ .Name = name, .Type = (texture_type)type, .Width = w, .Height = h, .Pixels = Pixels };
    Texture->EnableNearestFiltering = nearest != 0;
    Scene->Textures.push_back(Texture);
    Scene->DirtyFlags |= SCENE_DIRTY_TEXTURES;
    return Texture;
}
void* ptbh_create_checker_texture(void* s, const char* name, int type, const float* a, const float* b, int nearest) {
    texture* T = CreateCheckerTexture((scene*)s, name, (texture_type)type, glm::vec4(a[0], a[1], a[2], a[3]), glm::vec4(b[0], b[1], b[2], b[3]));
    T->EnableNearestFiltering = nearest != 0;
    return T;
}
void* ptbh_load_texture(void* s, const char* path, int type, const char* name) { return LoadTexture((scene*)s, path, (texture_type)type, name && name[0] ? name : nullptr); }

// The tail of LoadModelAsPrefab (scene.cpp:851-866) on caller vertices and faces: BuildMeshNode over the whole list.
void* ptbh_create_mesh(void* s, const char* name, const float* vertices, uint32_t nv, const uint32_t* faces, uint32_t nf) {
    scene* Scene = (scene*)s;
    mesh* Mesh = new mesh;
    Mesh->Name = name;
    Mesh->Depth = 0;                 // the reference never initialises it (scene.hpp:225); BuildMeshNode only raises it
    Mesh->Vertices.resize(nv);
    for (uint32_t i = 0; i < nv; i++) {
        const float* p = vertices + (size_t)i * 8;
        Mesh->Vertices[i].Position = {p[0], p[1], p[2]}; Mesh->Vertices[i].Normal = {p[3], p[4], p[5]}; Mesh->Vertices[i].UV = {p[6], p[7]};
    }
    Mesh->Faces.resize(nf);
    for (uint32_t i = 0; i < nf; i++) for (int k = 0; k < 3; k++) Mesh->Faces[i].VertexIndex[k] = faces[3 * i + k];
    Mesh->Nodes.reserve(2 * (size_t)nf);
    auto Root = mesh_node{ .FaceBeginIndex = 0, .FaceEndIndex = nf, .ChildNodeIndex = 0 };
    Mesh->Nodes.push_back(Root);
    BuildMeshNode(Mesh, 0, 0);
    Scene->Meshes.push_back(Mesh);
    Scene->DirtyFlags |= SCENE_DIRTY_MESHES;
    return Mesh;
}
uint32_t ptbh_mesh_depth(void* m) { return ((mesh*)m)->Depth; }
uint32_t ptbh_mesh_node_count(void* m) { return (uint32_t)((mesh*)m)->Nodes.size(); }
uint32_t ptbh_mesh_face_count(void* m) { return (uint32_t)((mesh*)m)->Faces.size(); }
uint32_t ptbh_scene_mesh_count(void* s) { return (uint32_t)((scene*)s)->Meshes.size(); }
void* ptbh_scene_mesh(void* s, uint32_t i) { return ((scene*)s)->Meshes[i]; }
void ptbh_mesh_faces(void* m, uint32_t* out) { mesh* M = (mesh*)m; std::memcpy(out, M->Faces.data(), M->Faces.size() * sizeof(mesh_face)); }
uint32_t ptbh_mesh_vertex_count(void* m) { return (uint32_t)((mesh*)m)->Vertices.size(); }
void ptbh_mesh_vertices(void* m, float* out) {
    mesh* M = (mesh*)m;
    for (size_t i = 0; i < M->Vertices.size(); i++) {
        const mesh_vertex& V = M->Vertices[i];
        float* o = out + 8 * i;
        o[0] = V.Position.x; o[1] = V.Position.y; o[2] = V.Position.z; o[3] = V.Normal.x; o[4] = V.Normal.y; o[5] = V.Normal.z; o[6] = V.UV.x; o[7] = V.UV.y;
    }
}

void* ptbh_load_model(void* s, const char* path, const char* name, void* default_material, const float* vertex_transform) {
    load_model_options O;
    O.Name = name && name[0] ? name : nullptr;
    O.DefaultMaterial = (material*)default_material;
    std::string p = path;
    size_t slash = p.find_last_of('/');
    O.DirectoryPath = slash == std::string::npos ? std::string(".") : p.substr(0, slash);
    if (vertex_transform) {
        std::memcpy((void*)&O.VertexTransform, vertex_transform, 64);
        mat4 N = O.VertexTransform;
        N[3] = glm::vec4(0, 0, 0, 1);
        O.NormalTransform = N;
    }
    return LoadModelAsPrefab((scene*)s, path, &O);
}
void* ptbh_create_prefab_instance(void* s, void* prefab_, void* parent) { return CreateEntity((scene*)s, (prefab*)prefab_, (entity*)parent); }
uint32_t ptbh_prefab_mesh_count(void* prefab_) {
    entity* E = ((prefab*)prefab_)->Entity;
    return E->Type == ENTITY_TYPE_MESH_INSTANCE ? 1u : (uint32_t)E->Children.size();
}

// ---- the reference's small helpers, for known-answer comparisons with path_tracer_b200/host/hmath.hpp ----
void ptbh_make_transform(const float* pos, const float* rot, const float* scale, float* out16) {
    mat4 m = MakeTransformMatrix(vec3(pos[0], pos[1], pos[2]), vec3(rot[0], rot[1], rot[2]), vec3(scale[0], scale[1], scale[2]));
    std::memcpy(out16, &m, 64);
}
void ptbh_inverse(const float* in16, float* out16) { mat4 m; std::memcpy((void*)&m, in16, 64); mat4 r = glm::inverse(m); std::memcpy(out16, &r, 64); }
uint32_t ptbh_pack_unit_vector(float x, float y, float z) { return PackUnitVector(vec3(x, y, z)); }

// ---- stb_image (vendored by the reference): what LoadTexture hands to the scene ----
// Returns 0 and fills w, h; pixels (w*h*4 floats) are copied to `out` when it is not NULL and holds `capacity` floats.
int refh_stbi_loadf(const char* path, int* w, int* h, float* out, size_t capacity) {
    int n = 0;
    float* px = stbi_loadf(path, w, h, &n, 4);
    if (!px) return -1;
    const size_t count = (size_t)*w * *h * 4;
    if (out && capacity >= count) std::memcpy(out, px, count * sizeof(float));
    stbi_image_free(px);
    return 0;
}

}  // extern "C"
