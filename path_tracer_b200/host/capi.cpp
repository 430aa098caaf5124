// Flat C surface over the host scene layer for Python (ctypes): tests and
// bench.py build scenes with the same calls the reference application makes
// (CreateScene / CreateEntity / CreateMaterial / PackSceneData) and obtain the
// ptb_scene_view that ptb_scene_update() consumes.  CPU only; no CUDA here.
#include "procedural.hpp"
#include <cstring>

using namespace ptb::host;

extern "C" {

void ptbh_set_spectrum_table_path(const char* path) { SetSpectrumTablePath(path); }
// Forces the table to exist (loads or builds+saves); returns 1 when ready.
int ptbh_ensure_spectrum_table() { return SharedSpectrumTable() != nullptr; }
void ptbh_rgb_to_spectrum(const float* rgb, float* beta) {
    vec3 b = GetParametricSpectrumCoefficients(SharedSpectrumTable(), {rgb[0], rgb[1], rgb[2]});
    beta[0] = b.x; beta[1] = b.y; beta[2] = b.z;
}
void ptbh_fit_spectrum(const float* rgb, float* beta) {
    vec3 b = FitParametricSpectrum({rgb[0], rgb[1], rgb[2]});
    beta[0] = b.x; beta[1] = b.y; beta[2] = b.z;
}
float ptbh_sample_spectrum(const float* beta, float lambda) { return SampleParametricSpectrum({beta[0], beta[1], beta[2]}, lambda); }

// ---- named procedural scenes (already packed) ----
void* ptbh_make_scene(const char* name, const double* p, int np) {
    auto P = [&](int i, double d) { return i < np ? p[i] : d; };
    if (!std::strcmp(name, "cornell")) return MakeCornellScene();
    if (!std::strcmp(name, "terrain")) return MakeTerrainScene((uint32_t)P(0, 708));
    if (!std::strcmp(name, "room")) return MakeRoomScene((uint32_t)P(0, 1), (uint32_t)P(1, 4096));
    if (!std::strcmp(name, "spherefield")) return MakeSphereFieldScene((uint32_t)P(0, 488), (uint32_t)P(1, 5), (int)P(2, 2));
    return nullptr;
}

// ---- generic builder API ----
void* ptbh_scene_create(int with_defaults) { return CreateScene(with_defaults != 0); }
void ptbh_scene_destroy(void* s) { DestroyScene((scene*)s); }
void ptbh_scene_set_atlas_size(void* s, uint32_t w, uint32_t h) { ((scene*)s)->AtlasWidth = w; ((scene*)s)->AtlasHeight = h; ((scene*)s)->DirtyFlags |= SCENE_DIRTY_TEXTURES; }
uint32_t ptbh_scene_pack(void* s) { return PackSceneData((scene*)s); }
void ptbh_scene_mark_dirty(void* s, uint32_t flags) { ((scene*)s)->DirtyFlags |= flags; }
void ptbh_scene_view(void* s, ptb_scene_view* out) { *out = ((scene*)s)->View(); }
void* ptbh_scene_root(void* s) { return &((scene*)s)->Root; }
void ptbh_root_set(void* s, float scatter_rate, float sky_brightness, float sky_sampling_probability, void* sky_texture) {
    scene* S = (scene*)s;
    S->Root.ScatterRate = scatter_rate; S->Root.SkyboxBrightness = sky_brightness;
    S->Root.SkyboxSamplingProbability = sky_sampling_probability; S->Root.SkyboxTexture = (texture*)sky_texture;
    S->DirtyFlags |= SCENE_DIRTY_GLOBALS | SCENE_DIRTY_SKYBOX_TEXTURE;
}

void* ptbh_create_entity(void* s, int type, void* parent) { return CreateEntity((scene*)s, (entity_type)type, (entity*)parent); }
void ptbh_destroy_entity(void* s, void* e) { DestroyEntity((scene*)s, (entity*)e); }
void ptbh_entity_set_transform(void* s, void* e, const float* pos, const float* rot, const float* scale) {
    entity* E = (entity*)e;
    E->Transform.Position = {pos[0], pos[1], pos[2]};
    E->Transform.Rotation = {rot[0], rot[1], rot[2]};
    E->Transform.Scale = {scale[0], scale[1], scale[2]};
    ((scene*)s)->DirtyFlags |= SCENE_DIRTY_SHAPES | SCENE_DIRTY_CAMERAS;
}
void ptbh_entity_look_at(void* s, void* e, const float* from, const float* to) {
    entity* E = (entity*)e;
    E->Transform.Position = {from[0], from[1], from[2]};
    E->Transform.Rotation = LookAtRotation({from[0], from[1], from[2]}, {to[0], to[1], to[2]});
    ((scene*)s)->DirtyFlags |= SCENE_DIRTY_SHAPES | SCENE_DIRTY_CAMERAS;
}
void ptbh_entity_set_active(void* s, void* e, int active) { ((entity*)e)->Active = active != 0; ((scene*)s)->DirtyFlags |= SCENE_DIRTY_SHAPES | SCENE_DIRTY_CAMERAS; }
void ptbh_entity_set_material(void* s, void* e, void* m) { ((entity*)e)->Material = (material*)m; ((scene*)s)->DirtyFlags |= SCENE_DIRTY_SHAPES; }
void ptbh_entity_set_mesh(void* s, void* e, void* m) { ((entity*)e)->Mesh = (mesh*)m; ((scene*)s)->DirtyFlags |= SCENE_DIRTY_SHAPES; }
uint32_t ptbh_entity_shape_index(void* e) { return ((entity*)e)->PackedShapeIndex; }
uint32_t ptbh_entity_camera_index(void* e) { return ((entity*)e)->PackedCameraIndex; }
// model: 0 pinhole {fov_deg, aperture_mm}; 1 thin lens {sensor_w_mm, sensor_h_mm, focal_mm, aperture_mm, focus_m}; 2 = 360 {}
void ptbh_camera_set(void* s, void* e, int model, const float* p) {
    entity* E = (entity*)e;
    E->CameraModel = (camera_model)model;
    if (model == CAMERA_MODEL_PINHOLE) { E->Pinhole.FieldOfViewInDegrees = p[0]; E->Pinhole.ApertureDiameterInMM = p[1]; }
    if (model == CAMERA_MODEL_THIN_LENS) { E->ThinLens.SensorSizeInMM = {p[0], p[1]}; E->ThinLens.FocalLengthInMM = p[2]; E->ThinLens.ApertureDiameterInMM = p[3]; E->ThinLens.FocusDistance = p[4]; }
    ((scene*)s)->DirtyFlags |= SCENE_DIRTY_CAMERAS;
}

void* ptbh_create_material(void* s, int type, const char* name) { return CreateMaterial((scene*)s, (material_type)type, name); }
// Fields by name keep the binding small; colours are 3 floats, scalars 1.
int ptbh_material_set(void* s, void* m, const char* field, const float* v) {
    material* M = (material*)m;
    ((scene*)s)->DirtyFlags |= SCENE_DIRTY_MATERIALS;
    auto c3 = [&](vec3& d) { d = {v[0], v[1], v[2]}; return 0; };
    if (!std::strcmp(field, "BaseColor")) return c3(M->BaseColor);
    if (!std::strcmp(field, "SpecularColor")) return c3(M->SpecularColor);
    if (!std::strcmp(field, "TransmissionColor")) return c3(M->TransmissionColor);
    if (!std::strcmp(field, "ScatteringColor")) return c3(M->ScatteringColor);
    if (!std::strcmp(field, "Roughness")) { M->Roughness = v[0]; return 0; }
    if (!std::strcmp(field, "RoughnessAnisotropy")) { M->RoughnessAnisotropy = v[0]; return 0; }
    if (!std::strcmp(field, "IOR")) { M->IOR = v[0]; return 0; }
    if (!std::strcmp(field, "AbbeNumber")) { M->AbbeNumber = v[0]; return 0; }
    if (!std::strcmp(field, "TransmissionDepth")) { M->TransmissionDepth = v[0]; return 0; }
    if (!std::strcmp(field, "ScatteringAnisotropy")) { M->ScatteringAnisotropy = v[0]; return 0; }
    return -1;
}
int ptbh_material_set_texture(void* s, void* m, const char* field, void* t) {
    material* M = (material*)m; texture* T = (texture*)t;
    ((scene*)s)->DirtyFlags |= SCENE_DIRTY_MATERIALS;
    if (!std::strcmp(field, "BaseTexture")) { M->BaseTexture = T; return 0; }
    if (!std::strcmp(field, "SpecularTexture")) { M->SpecularTexture = T; return 0; }
    if (!std::strcmp(field, "RoughnessTexture")) { M->RoughnessTexture = T; return 0; }
    if (!std::strcmp(field, "RoughnessAnisotropyTexture")) { M->RoughnessAnisotropyTexture = T; return 0; }
    return -1;
}
uint32_t ptbh_material_slot(void* m) { return ((material*)m)->PackedMaterialIndex; }

void* ptbh_create_texture(void* s, const char* name, int type, uint32_t w, uint32_t h, const float* rgba, int nearest) {
    texture* T = CreateTexture((scene*)s, name, (texture_type)type, w, h, (vec4 const*)rgba);
    T->EnableNearestFiltering = nearest != 0;
    return T;
}
void* ptbh_create_checker_texture(void* s, const char* name, int type, const float* a, const float* b, int nearest) {
    texture* T = CreateCheckerTexture((scene*)s, name, (texture_type)type, {a[0], a[1], a[2], a[3]}, {b[0], b[1], b[2], b[3]});
    T->EnableNearestFiltering = nearest != 0;
    return T;
}

// vertices: nv x 8 floats (px,py,pz, nx,ny,nz, u,v); faces: nf x 3 uint32.
void* ptbh_create_mesh(void* s, const char* name, const float* vertices, uint32_t nv, const uint32_t* faces, uint32_t nf) {
    std::vector<mesh_vertex> V(nv);
    for (uint32_t i = 0; i < nv; i++) {
        const float* p = vertices + (size_t)i * 8;
        V[i].Position = {p[0], p[1], p[2]}; V[i].Normal = {p[3], p[4], p[5]}; V[i].UV = {p[6], p[7]};
    }
    std::vector<mesh_face> F(nf);
    for (uint32_t i = 0; i < nf; i++) F[i] = {{faces[3 * i], faces[3 * i + 1], faces[3 * i + 2]}};
    return CreateMesh((scene*)s, name, std::move(V), std::move(F));
}
uint32_t ptbh_mesh_depth(void* m) { return ((mesh*)m)->Depth; }
uint32_t ptbh_mesh_node_count(void* m) { return (uint32_t)((mesh*)m)->Nodes.size(); }
uint32_t ptbh_mesh_face_count(void* m) { return (uint32_t)((mesh*)m)->Faces.size(); }
uint32_t ptbh_scene_mesh_count(void* s) { return (uint32_t)((scene*)s)->Meshes.size(); }
void* ptbh_scene_mesh(void* s, uint32_t i) { return ((scene*)s)->Meshes[i]; }
// Face vertex indices after the builder reordered them (nf x 3).
void ptbh_mesh_faces(void* m, uint32_t* out) { mesh* M = (mesh*)m; std::memcpy(out, M->Faces.data(), M->Faces.size() * sizeof(mesh_face)); }

// ---- model import, prefabs, image files (model.cpp) ----
// vertex_transform: 16 floats column-major or NULL (identity; also used for normals with w = 0).
void* ptbh_load_model(void* s, const char* path, const char* name, void* default_material, const float* vertex_transform) {
    load_model_options O;
    O.Name = name && name[0] ? name : nullptr;
    O.DefaultMaterial = (material*)default_material;
    if (vertex_transform) {
        std::memcpy((void*)&O.VertexTransform, vertex_transform, 64);
        mat4 N = O.VertexTransform;                       // rotation/scale part for directions
        N.c[3] = {0, 0, 0, 1};
        O.NormalTransform = N;
    }
    return LoadModelAsPrefab((scene*)s, path, &O);
}
void* ptbh_create_prefab_instance(void* s, void* prefab_, void* parent) { return CreateEntity((scene*)s, (prefab*)prefab_, (entity*)parent); }
void ptbh_destroy_prefab(void* s, void* prefab_) { DestroyPrefab((scene*)s, (prefab*)prefab_); }
uint32_t ptbh_prefab_mesh_count(void* prefab_) {
    entity* E = ((prefab*)prefab_)->Entity;
    return E->Type == ENTITY_TYPE_MESH_INSTANCE ? 1u : (uint32_t)E->Children.size();
}
void* ptbh_load_texture(void* s, const char* path, int type, const char* name) { return LoadTexture((scene*)s, path, (texture_type)type, name && name[0] ? name : nullptr); }
// DecodeImageFile (image.cpp): 0 and the size on success; pixels (w*h*4 floats) are copied when `out` holds `capacity` floats.
int ptbh_decode_image(const char* path, int* w, int* h, float* out, size_t capacity) {
    std::vector<vec4> px;
    if (!DecodeImageFile(path, *w, *h, px)) return -1;
    if (out && capacity >= px.size() * 4) std::memcpy(out, px.data(), px.size() * sizeof(vec4));
    return 0;
}
int ptbh_write_png(const char* path, uint32_t w, uint32_t h, const uint8_t* rgba) { return WriteImagePNG(path, w, h, rgba) ? 0 : -1; }
int ptbh_write_pfm(const char* path, uint32_t w, uint32_t h, const float* rgb) { return WriteImagePFM(path, w, h, rgb) ? 0 : -1; }
uint32_t ptbh_mesh_vertex_count(void* m) { return (uint32_t)((mesh*)m)->Vertices.size(); }
// nv x 8 floats (px,py,pz, nx,ny,nz, u,v)
void ptbh_mesh_vertices(void* m, float* out) {
    mesh* M = (mesh*)m;
    for (size_t i = 0; i < M->Vertices.size(); i++) {
        const mesh_vertex& V = M->Vertices[i];
        float* o = out + 8 * i;
        o[0] = V.Position.x; o[1] = V.Position.y; o[2] = V.Position.z; o[3] = V.Normal.x; o[4] = V.Normal.y; o[5] = V.Normal.z; o[6] = V.UV.x; o[7] = V.UV.y;
    }
}

// ---- scene files (serializer.cpp) ----
static thread_local std::string g_io_error;
const char* ptbh_last_error() { return g_io_error.c_str(); }
int ptbh_save_scene(void* s, const char* path) { return SaveScene(path, (scene*)s, &g_io_error) ? 0 : -1; }
void* ptbh_load_scene(const char* path) { return LoadScene(path, &g_io_error); }

// 4x4 helpers exposed for unit tests of the matrix kit.
void ptbh_make_transform(const float* pos, const float* rot, const float* scale, float* out16) {
    mat4 m = make_transform({pos[0], pos[1], pos[2]}, {rot[0], rot[1], rot[2]}, {scale[0], scale[1], scale[2]});
    std::memcpy(out16, &m, 64);
}
void ptbh_inverse(const float* in16, float* out16) { mat4 m; std::memcpy((void*)&m, in16, 64); mat4 r = inverse(m); std::memcpy(out16, &r, 64); }
uint32_t ptbh_pack_unit_vector(float x, float y, float z) { return pack_unit_vector({x, y, z}); }
uint32_t ptbh_pack_half2(float x, float y) { return pack_half2x16(x, y); }

}  // extern "C"
