// Procedural scene generators (see procedural.hpp).  All randomness comes from
// SplitMix64 with the seeds named in SURVEY.md §8(d).
#include "procedural.hpp"
#include <atomic>
#include <map>
#include <thread>

namespace ptb { namespace host {

namespace {

struct splitmix {
    uint64_t s;
    uint64_t next() { uint64_t z = (s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }
    float unit() { return (float)(next() >> 40) * (1.0f / 16777216.0f); }
};

inline float hash_lattice(uint64_t seed, int x, int y, int z = 0) {
    splitmix r{seed ^ ((uint64_t)(uint32_t)x * 0x9E3779B1ull) ^ ((uint64_t)(uint32_t)y * 0x85EBCA77ull << 1) ^ ((uint64_t)(uint32_t)z * 0xC2B2AE3Dull << 2)};
    return r.unit();
}
inline float smooth(float t) { return t * t * (3.0f - 2.0f * t); }
float value_noise2(uint64_t seed, float x, float y) {
    int ix = (int)std::floor(x), iy = (int)std::floor(y);
    float fx = smooth(x - ix), fy = smooth(y - iy);
    float a = hash_lattice(seed, ix, iy), b = hash_lattice(seed, ix + 1, iy);
    float c = hash_lattice(seed, ix, iy + 1), d = hash_lattice(seed, ix + 1, iy + 1);
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy;
}
float value_noise3(uint64_t seed, vec3 p) {
    int ix = (int)std::floor(p.x), iy = (int)std::floor(p.y), iz = (int)std::floor(p.z);
    float fx = smooth(p.x - ix), fy = smooth(p.y - iy), fz = smooth(p.z - iz);
    float v[2];
    for (int k = 0; k < 2; k++) {
        float a = hash_lattice(seed, ix, iy, iz + k), b = hash_lattice(seed, ix + 1, iy, iz + k);
        float c = hash_lattice(seed, ix, iy + 1, iz + k), d = hash_lattice(seed, ix + 1, iy + 1, iz + k);
        v[k] = (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy;
    }
    return v[0] * (1 - fz) + v[1] * fz;
}
float fbm2(uint64_t seed, float x, float y, int octaves) {
    float sum = 0, amp = 0.5f, f = 1.0f;
    for (int o = 0; o < octaves; o++) { sum += amp * value_noise2(seed + o, x * f, y * f); amp *= 0.5f; f *= 2.0f; }
    return sum;
}
float fbm3(uint64_t seed, vec3 p, int octaves) {
    float sum = 0, amp = 0.5f, f = 1.0f;
    for (int o = 0; o < octaves; o++) { sum += amp * value_noise3(seed + o, p * f); amp *= 0.5f; f *= 2.0f; }
    return sum;
}

entity* add_camera(scene* S, vec3 from, vec3 to) {
    entity* C = CreateEntity(S, ENTITY_TYPE_CAMERA);
    C->Name = "Camera";
    C->Transform.Position = from;
    C->Transform.Rotation = LookAtRotation(from, to);
    return C;
}
entity* add_shape(scene* S, entity_type T, material* M, vec3 pos, vec3 scale, vec3 rot = {0, 0, 0}, entity* parent = nullptr) {
    entity* E = CreateEntity(S, T, parent);
    E->Material = M;
    E->Transform.Position = pos; E->Transform.Scale = scale; E->Transform.Rotation = rot;
    return E;
}
material* diffuse(scene* S, const char* name, vec3 c) { material* M = CreateMaterial(S, MATERIAL_TYPE_BASIC_DIFFUSE, name); M->BaseColor = c; return M; }
material* metal(scene* S, const char* name, vec3 c, float rough) { material* M = CreateMaterial(S, MATERIAL_TYPE_BASIC_METAL, name); M->BaseColor = c; M->Roughness = rough; return M; }
material* glass(scene* S, const char* name, float ior, float abbe, float rough) {
    material* M = CreateMaterial(S, MATERIAL_TYPE_BASIC_TRANSLUCENT, name); M->IOR = ior; M->AbbeNumber = abbe; M->Roughness = rough; return M;
}

// Icosphere of subdivision `level` (20*4^level faces), shared vertices.
void icosphere(uint32_t level, std::vector<vec3>& P, std::vector<mesh_face>& F) {
    const float t = (1.0f + std::sqrt(5.0f)) / 2.0f;
    P = {{-1, t, 0}, {1, t, 0}, {-1, -t, 0}, {1, -t, 0}, {0, -1, t}, {0, 1, t}, {0, -1, -t}, {0, 1, -t}, {t, 0, -1}, {t, 0, 1}, {-t, 0, -1}, {-t, 0, 1}};
    for (auto& p : P) p = normalize3(p);
    const uint32_t I[20][3] = {{0, 11, 5}, {0, 5, 1}, {0, 1, 7}, {0, 7, 10}, {0, 10, 11}, {1, 5, 9}, {5, 11, 4}, {11, 10, 2}, {10, 7, 6}, {7, 1, 8},
                               {3, 9, 4}, {3, 4, 2}, {3, 2, 6}, {3, 6, 8}, {3, 8, 9}, {4, 9, 5}, {2, 4, 11}, {6, 2, 10}, {8, 6, 7}, {9, 8, 1}};
    F.clear();
    for (auto& f : I) F.push_back({{f[0], f[1], f[2]}});
    for (uint32_t l = 0; l < level; l++) {
        std::map<uint64_t, uint32_t> mid;
        auto midpoint = [&](uint32_t a, uint32_t b) {
            uint64_t key = a < b ? ((uint64_t)a << 32) | b : ((uint64_t)b << 32) | a;
            auto it = mid.find(key);
            if (it != mid.end()) return it->second;
            P.push_back(normalize3((P[a] + P[b]) * 0.5f));
            return mid[key] = (uint32_t)P.size() - 1;
        };
        std::vector<mesh_face> G;
        G.reserve(F.size() * 4);
        for (auto& f : F) {
            uint32_t a = f.VertexIndex[0], b = f.VertexIndex[1], c = f.VertexIndex[2];
            uint32_t ab = midpoint(a, b), bc = midpoint(b, c), ca = midpoint(c, a);
            G.push_back({{a, ab, ca}}); G.push_back({{b, bc, ab}}); G.push_back({{c, ca, bc}}); G.push_back({{ab, bc, ca}});
        }
        F.swap(G);
    }
}

// Displaced icosphere mesh with area-weighted smooth normals and spherical UVs.
mesh* displaced_sphere_mesh(scene* S, const char* name, uint32_t level, uint64_t seed, float amplitude) {      // S == nullptr: build only
    std::vector<vec3> P; std::vector<mesh_face> F;
    icosphere(level, P, F);
    for (auto& p : P) { float d = 1.0f + amplitude * (fbm3(seed, p * 2.5f + vec3{7.1f, 3.3f, 5.7f}, 4) - 0.5f); p = p * d; }
    std::vector<vec3> N(P.size());
    for (auto& f : F) {
        vec3 n = cross3(P[f.VertexIndex[1]] - P[f.VertexIndex[0]], P[f.VertexIndex[2]] - P[f.VertexIndex[0]]);
        for (int k = 0; k < 3; k++) N[f.VertexIndex[k]] = N[f.VertexIndex[k]] + n;
    }
    std::vector<mesh_vertex> V(P.size());
    for (size_t i = 0; i < P.size(); i++) {
        vec3 d = normalize3(P[i]);
        V[i].Position = P[i];
        V[i].Normal = normalize3(N[i]);
        V[i].UV = {0.5f + std::atan2(d.y, d.x) / 6.2831853f, 0.5f + std::asin(std::max(-1.0f, std::min(1.0f, d.z))) / 3.14159265f};
    }
    return S ? CreateMesh(S, name, std::move(V), std::move(F)) : BuildMesh(name, std::move(V), std::move(F));
}

// Regular grid patch in the local xy plane, displaced along +z by h(u,v); normals by central differences.
template <class H>
mesh* grid_mesh(scene* S, const char* name, uint32_t nx, uint32_t ny, float sx, float sy, float uv_scale, H&& h) {
    std::vector<mesh_vertex> V((size_t)(nx + 1) * (ny + 1));
    auto at = [&](uint32_t i, uint32_t j) -> mesh_vertex& { return V[(size_t)j * (nx + 1) + i]; };
    for (uint32_t j = 0; j <= ny; j++) for (uint32_t i = 0; i <= nx; i++) {
        float u = (float)i / nx, v = (float)j / ny;
        at(i, j).Position = {(u - 0.5f) * sx, (v - 0.5f) * sy, h(u, v)};
        at(i, j).UV = {u * uv_scale, v * uv_scale};
    }
    for (uint32_t j = 0; j <= ny; j++) for (uint32_t i = 0; i <= nx; i++) {
        uint32_t i0 = i > 0 ? i - 1 : i, i1 = i < nx ? i + 1 : i, j0 = j > 0 ? j - 1 : j, j1 = j < ny ? j + 1 : j;
        vec3 dx = at(i1, j).Position - at(i0, j).Position, dy = at(i, j1).Position - at(i, j0).Position;
        at(i, j).Normal = normalize3(cross3(dx, dy));
    }
    std::vector<mesh_face> F;
    F.reserve((size_t)nx * ny * 2);
    for (uint32_t j = 0; j < ny; j++) for (uint32_t i = 0; i < nx; i++) {
        uint32_t a = j * (nx + 1) + i, b = a + 1, c = a + (nx + 1), d = c + 1;
        F.push_back({{a, b, d}}); F.push_back({{a, d, c}});
    }
    return CreateMesh(S, name, std::move(V), std::move(F));
}

}  // namespace

vec3 LookAtRotation(vec3 From, vec3 To) {
    vec3 f = normalize3(To - From);
    float pitch = std::asin(std::max(-1.0f, std::min(1.0f, f.z)));
    float yaw = std::atan2(-f.x, f.y);
    return {1.57079632679f + pitch, 0.0f, yaw};
}

// ---- C1 ------------------------------------------------------------------------
scene* MakeCornellScene() {
    scene* S = CreateScene(false);
    S->AtlasWidth = S->AtlasHeight = 64;
    material* White = diffuse(S, "White", {0.73f, 0.73f, 0.73f});
    material* Red = diffuse(S, "Red", {0.65f, 0.05f, 0.05f});
    material* Green = diffuse(S, "Green", {0.12f, 0.45f, 0.15f});
    material* Mirror = metal(S, "Mirror", {0.95f, 0.95f, 0.95f}, 0.0f);
    material* Brushed = metal(S, "Brushed", {0.9f, 0.6f, 0.3f}, 0.35f);
    material* Glass = glass(S, "Glass", 1.5f, 1e4f, 0.0f);
    // Room: floor plane, three slab walls, a ceiling with a skylight gap; the
    // camera-facing side is open (the sky is the only light, quirk Q16).
    add_shape(S, ENTITY_TYPE_PLANE, White, {0, 0, 0}, {1, 1, 1});
    add_shape(S, ENTITY_TYPE_CUBE, Red, {-2.05f, 0, 2}, {0.05f, 2, 2});
    add_shape(S, ENTITY_TYPE_CUBE, Green, {+2.05f, 0, 2}, {0.05f, 2, 2});
    add_shape(S, ENTITY_TYPE_CUBE, White, {0, 2.05f, 2}, {2.1f, 0.05f, 2});
    add_shape(S, ENTITY_TYPE_CUBE, White, {-1.3f, 0, 4.05f}, {0.8f, 2, 0.05f});
    add_shape(S, ENTITY_TYPE_CUBE, White, {+1.3f, 0, 4.05f}, {0.8f, 2, 0.05f});
    // Contents.
    add_shape(S, ENTITY_TYPE_CUBE, White, {-0.7f, 0.6f, 1.2f}, {0.6f, 0.6f, 1.2f}, {0, 0, 0.3f});
    add_shape(S, ENTITY_TYPE_CUBE, Brushed, {0.8f, -0.4f, 0.6f}, {0.6f, 0.6f, 0.6f}, {0, 0, -0.3f});
    add_shape(S, ENTITY_TYPE_SPHERE, Mirror, {0.8f, -0.4f, 1.7f}, {0.5f, 0.5f, 0.5f});
    add_shape(S, ENTITY_TYPE_SPHERE, Glass, {-0.6f, -0.9f, 0.5f}, {0.5f, 0.5f, 0.5f});
    add_shape(S, ENTITY_TYPE_SPHERE, Red, {0.1f, 1.0f, 0.35f}, {0.35f, 0.35f, 0.35f});
    entity* Cam = add_camera(S, {0, -7.5f, 2.0f}, {0, 0, 1.9f});
    Cam->Pinhole.FieldOfViewInDegrees = 40.0f;
    PackSceneData(S);
    return S;
}

// ---- C2 ------------------------------------------------------------------------
scene* MakeTerrainScene(uint32_t Grid) {
    scene* S = CreateScene(false);
    S->AtlasWidth = S->AtlasHeight = 64;
    material* Grey = diffuse(S, "Terrain", {0.7f, 0.7f, 0.7f});
    material* Ground = diffuse(S, "Ground", {0.5f, 0.5f, 0.5f});
    const uint64_t Seed = 0x5EED0001ull;
    mesh* M = grid_mesh(S, "Terrain", Grid, Grid, 10.0f, 10.0f, 8.0f, [&](float u, float v) {
        return 2.5f * fbm2(Seed, u * 6.0f, v * 6.0f, 5);
    });
    entity* T = CreateEntity(S, ENTITY_TYPE_MESH_INSTANCE);
    T->Name = "Terrain"; T->Mesh = M; T->Material = Grey;
    add_shape(S, ENTITY_TYPE_PLANE, Ground, {0, 0, 0}, {1, 1, 1});
    entity* Cam = add_camera(S, {0, -12.0f, 6.0f}, {0, 0, 0.8f});
    Cam->Pinhole.FieldOfViewInDegrees = 45.0f;
    PackSceneData(S);
    return S;
}

// ---- C3 ------------------------------------------------------------------------
scene* MakeRoomScene(uint32_t Detail, uint32_t AtlasSize) {
    scene* S = CreateScene(false);
    S->AtlasWidth = S->AtlasHeight = AtlasSize;
    const uint64_t Seed = 0x5EED0003ull;
    const uint32_t TexN = std::min(1024u, AtlasSize / 4);
    // Textures: checker, two noise albedos, a roughness map, and the HDR sky.
    auto make_tex = [&](const char* name, texture_type type, uint32_t w, uint32_t h, auto&& fn) {
        std::vector<vec4> px((size_t)w * h);
        for (uint32_t y = 0; y < h; y++) for (uint32_t x = 0; x < w; x++) px[(size_t)y * w + x] = fn((x + 0.5f) / w, (y + 0.5f) / h);
        return CreateTexture(S, name, type, w, h, px.data());
    };
    texture* Checker = make_tex("Checker", TEXTURE_TYPE_REFLECTANCE_WITH_ALPHA, TexN, TexN, [&](float u, float v) {
        bool c = (((int)(u * 8) + (int)(v * 8)) & 1) != 0;
        return c ? vec4{0.8f, 0.8f, 0.75f, 1} : vec4{0.25f, 0.22f, 0.2f, 1};
    });
    texture* Wood = make_tex("Wood", TEXTURE_TYPE_REFLECTANCE_WITH_ALPHA, TexN, TexN, [&](float u, float v) {
        float n = fbm2(Seed + 11, u * 3.0f, v * 24.0f, 4);
        return vec4{0.45f + 0.3f * n, 0.28f + 0.2f * n, 0.12f + 0.1f * n, 1};
    });
    texture* Plaster = make_tex("Plaster", TEXTURE_TYPE_REFLECTANCE_WITH_ALPHA, TexN, TexN, [&](float u, float v) {
        float n = fbm2(Seed + 23, u * 16.0f, v * 16.0f, 4);
        return vec4{0.7f + 0.2f * n, 0.68f + 0.2f * n, 0.62f + 0.2f * n, 1};
    });
    texture* RoughMap = make_tex("Roughness", TEXTURE_TYPE_RAW, TexN, TexN, [&](float u, float v) {
        float n = fbm2(Seed + 37, u * 10.0f, v * 10.0f, 3);
        return vec4{0.3f + 0.7f * n, 0, 0, 1};
    });
    const uint32_t SkyW = std::min(2048u, AtlasSize / 2), SkyH = SkyW / 2;
    const vec3 SunDir = normalize3({0.35f, -0.55f, 0.75f});
    texture* Sky = make_tex("Sky", TEXTURE_TYPE_RADIANCE, SkyW, SkyH, [&](float u, float v) {
        // equirect: u -> phi in [-pi,pi]; memory row 0 is the zenith (scene.glsl.inc:214-218 with the
        // V flip of the atlas placement, scene.cpp:1168-1177; same convention as scene.cpp:1573-1576)
        float phi = (u - 0.5f) * 6.2831853f, theta = (0.5f - v) * 3.14159265f;
        vec3 d{std::cos(theta) * std::cos(phi), std::cos(theta) * std::sin(phi), std::sin(theta)};
        float up = std::max(0.0f, d.z);
        vec3 c = vec3{0.25f, 0.35f, 0.6f} * (0.3f + 0.7f * up) + vec3{0.5f, 0.45f, 0.4f} * (1.0f - up) * 0.4f;
        float s = std::max(0.0f, dot3(d, SunDir));
        c = c + vec3{1.0f, 0.9f, 0.7f} * (40.0f * std::pow(s, 600.0f) + 1.5f * std::pow(s, 20.0f));
        if (d.z < 0) c = c * 0.2f;
        return vec4{c.x, c.y, c.z, 1};
    });
    S->Root.SkyboxTexture = Sky;
    S->Root.SkyboxSamplingProbability = 0.5f;
    S->Root.SkyboxBrightness = 1.0f;
    S->DirtyFlags |= SCENE_DIRTY_SKYBOX_TEXTURE;

    material* Floor = diffuse(S, "Floor", {1, 1, 1}); Floor->BaseTexture = Wood;
    material* Wall = diffuse(S, "Wall", {1, 1, 1}); Wall->BaseTexture = Plaster;
    material* Tile = diffuse(S, "Tile", {1, 1, 1}); Tile->BaseTexture = Checker;
    material* Bronze = metal(S, "Bronze", {0.8f, 0.5f, 0.25f}, 0.6f); Bronze->RoughnessTexture = RoughMap;
    material* Clear = glass(S, "Clear", 1.5f, 20.0f, 0.0f);
    material* Frosted = glass(S, "Frosted", 1.5f, 20.0f, 0.2f);
    material* Amber = glass(S, "Amber", 1.45f, 40.0f, 0.0f); Amber->TransmissionColor = {0.9f, 0.6f, 0.2f}; Amber->TransmissionDepth = 0.5f;

    // Room shell: subdivided, gently displaced panels (floor, back, left, right); open top and front.
    const uint32_t n = 80 * Detail;      // 4 panels x 2*n*n triangles
    auto bump = [&](uint64_t s) { return [=](float u, float v) { return 0.03f * (fbm2(s, u * 12.0f, v * 12.0f, 3) - 0.5f); }; };
    mesh* PanelA = grid_mesh(S, "FloorPanel", n, n, 8.0f, 8.0f, 4.0f, bump(Seed + 1));
    mesh* PanelB = grid_mesh(S, "WallPanel", n, n, 8.0f, 4.0f, 2.0f, bump(Seed + 2));
    auto inst = [&](mesh* M, material* Mt, vec3 pos, vec3 rot) { entity* E = CreateEntity(S, ENTITY_TYPE_MESH_INSTANCE); E->Mesh = M; E->Material = Mt; E->Transform.Position = pos; E->Transform.Rotation = rot; return E; };
    inst(PanelA, Floor, {0, 0, 0}, {0, 0, 0});
    inst(PanelB, Wall, {0, 4, 2}, {1.57079632679f, 0, 0});                 // back wall, normal -y
    inst(PanelB, Wall, {-4, 0, 2}, {1.57079632679f, 0, -1.57079632679f});  // left wall (instanced mesh, shared BVH)
    inst(PanelB, Tile, {+4, 0, 2}, {1.57079632679f, 0, +1.57079632679f});  // right wall
    // 20 displaced icospheres (level 3 at Detail 1 -> 1280 tris each) scattered on the floor.
    splitmix R{Seed};
    const uint32_t lvl = Detail >= 2 ? 4 : 3;
    for (int k = 0; k < 20; k++) {
        char nm[32]; std::snprintf(nm, sizeof nm, "Rock %d", k);
        mesh* M = displaced_sphere_mesh(S, nm, lvl, Seed + 100 + k, 0.35f);
        float r = 0.25f + 0.25f * R.unit();
        entity* E = inst(M, k % 3 == 0 ? Bronze : (k % 3 == 1 ? Tile : Wall), {-3.2f + 6.4f * R.unit(), -1.0f + 4.5f * R.unit(), r * 0.9f}, {R.unit(), R.unit(), 6.28f * R.unit()});
        E->Transform.Scale = {r, r, r};
    }
    add_shape(S, ENTITY_TYPE_SPHERE, Clear, {-1.2f, -1.6f, 0.6f}, {0.6f, 0.6f, 0.6f});
    add_shape(S, ENTITY_TYPE_SPHERE, Frosted, {0.3f, -1.9f, 0.45f}, {0.45f, 0.45f, 0.45f});
    add_shape(S, ENTITY_TYPE_SPHERE, Amber, {1.5f, -1.4f, 0.55f}, {0.55f, 0.55f, 0.55f});
    entity* Cam = add_camera(S, {0.2f, -6.5f, 1.6f}, {0, 0, 0.8f});
    Cam->CameraModel = CAMERA_MODEL_THIN_LENS;
    Cam->ThinLens.SensorSizeInMM = {36.0f, 20.25f};
    Cam->ThinLens.FocalLengthInMM = 35.0f;
    Cam->ThinLens.ApertureDiameterInMM = 8.0f;
    Cam->ThinLens.FocusDistance = 5.0f;
    PackSceneData(S);
    return S;
}

// ---- C4 / C5 -------------------------------------------------------------------
scene* MakeSphereFieldScene(uint32_t Copies, uint32_t Level, int CameraModel) {
    scene* S = CreateScene(false);
    S->AtlasWidth = S->AtlasHeight = 64;
    const uint64_t Seed = 0x5EED0004ull;
    material* Mats[6] = {
        diffuse(S, "D0", {0.75f, 0.3f, 0.25f}), metal(S, "M0", {0.9f, 0.85f, 0.7f}, 0.3f), glass(S, "T0", 1.5f, 30.0f, 0.1f),
        diffuse(S, "D1", {0.3f, 0.55f, 0.75f}), metal(S, "M1", {0.7f, 0.75f, 0.9f}, 0.3f), diffuse(S, "D2", {0.6f, 0.7f, 0.35f}),
    };
    material* Ground = diffuse(S, "Ground", {0.45f, 0.45f, 0.45f});
    add_shape(S, ENTITY_TYPE_PLANE, Ground, {0, 0, -1.2f}, {1, 1, 1});
    // Jittered grid around the origin; the cell at the centre is left empty for the 360 camera.
    uint32_t side = (uint32_t)std::ceil(std::sqrt((double)Copies + 1));
    splitmix R{Seed};
    // Placement first (one RNG stream, consumed in the order of the sequential recipe), then the meshes — copied, not
    // instanced: every blob is its own mesh with its own BVH — on all host threads, then attached
    // in order: the packed arrays are identical to a sequential build.
    struct blob { vec3 Position, Scale, Rotation; };
    std::vector<blob> Blobs;
    for (uint32_t j = 0; j < side && Blobs.size() < Copies; j++) for (uint32_t i = 0; i < side && Blobs.size() < Copies; i++) {
        float cx = ((float)i - 0.5f * (side - 1)) * 2.6f, cy = ((float)j - 0.5f * (side - 1)) * 2.6f;
        float jx = (R.unit() - 0.5f) * 0.8f, jy = (R.unit() - 0.5f) * 0.8f, jz = (R.unit() - 0.5f) * 1.0f, rs = 0.8f + 0.4f * R.unit();
        if (std::fabs(cx) < 1.4f && std::fabs(cy) < 1.4f) continue;
        blob B;
        B.Position = {cx + jx, cy + jy, jz};
        B.Scale = {rs, rs, rs};
        float rx = R.unit() * 6.28f, ry = R.unit() * 6.28f, rz = R.unit() * 6.28f;
        B.Rotation = {rx, ry, rz};
        Blobs.push_back(B);
    }
    std::vector<mesh*> Built(Blobs.size(), nullptr);
    {
        std::atomic<size_t> Next(0);
        auto work = [&]() {
            for (size_t k = Next.fetch_add(1); k < Blobs.size(); k = Next.fetch_add(1)) {
                char nm[32]; std::snprintf(nm, sizeof nm, "Blob %u", (unsigned)k);
                Built[k] = displaced_sphere_mesh(nullptr, nm, Level, Seed + 1000 + k, 0.4f);
            }
        };
        unsigned T = std::max(1u, std::min<unsigned>(std::thread::hardware_concurrency(), (unsigned)Blobs.size()));
        std::vector<std::thread> Pool;
        for (unsigned t = 1; t < T; t++) Pool.emplace_back(work);
        work();
        for (auto& th : Pool) th.join();
    }
    for (size_t k = 0; k < Blobs.size(); k++) {
        mesh* M = AttachMesh(S, Built[k]);
        entity* E = CreateEntity(S, ENTITY_TYPE_MESH_INSTANCE);
        E->Mesh = M; E->Material = Mats[k % 6];
        E->Transform.Position = Blobs[k].Position;
        E->Transform.Scale = Blobs[k].Scale;
        E->Transform.Rotation = Blobs[k].Rotation;
    }
    entity* Cam;
    if (CameraModel == CAMERA_MODEL_360) {
        Cam = add_camera(S, {0, 0, 0.2f}, {0, 1, 0.2f});
        Cam->CameraModel = CAMERA_MODEL_360;
    } else {
        float ext = side * 1.3f;
        Cam = add_camera(S, {0, -ext * 1.6f, ext * 0.9f}, {0, 0, 0});
        Cam->Pinhole.FieldOfViewInDegrees = 50.0f;
    }
    PackSceneData(S);
    return S;
}

}}  // namespace ptb::host
