// Host scene layer, part 4: model import, prefabs and image files.
//
// LoadModelAsPrefab follows /root/reference/src/scene/scene.cpp:601-903 step by step
// (missing normals accumulated from unit face normals, one mesh per (shape, material)
// pair, shapes recentred on their bounding-box centre, exact-match vertex welding,
// BuildMeshNode per mesh, single mesh entity or container prefab).  The reference reads
// the file with tinyobjloader (un-vendored dependency semantics: 1-based / negative
// indices, `v/vt/vn` corners, polygons triangulated as tinyobjloader does (triangulate() below), `o`/`g` start a shape, `usemtl`
// selects the material, `mtllib` names the material file); the subset of that format
// meshes actually use is parsed here directly.  Differences, all deliberate:
//   * materials become BasicDiffuse (BaseColor = Kd, BaseTexture = map_Kd): the reference
//     creates OpenPBR materials, which its own build has disabled;
//   * a mesh's child entity is placed at the origin of ITS shape (the reference indexes
//     the origin list by mesh number, scene.cpp:889, which is wrong as soon as a shape is
//     empty or uses two materials);
//   * material groups of a shape are ordered by first use (the reference iterates an
//     unordered_set, scene.cpp:773-778);
//   * textures are decoded by host/image.cpp (PNG, JPEG, Radiance HDR with stb_image's conversions; PPM, PFM).
// The reference has no image writer; the headless replacement for its swapchain is a file:
// PNG (8-bit sRGB, stored deflate blocks) and PFM (linear float).
#include "scene.hpp"

#include <algorithm>
#include <atomic>
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <thread>
#include <unordered_map>

namespace ptb { namespace host {

namespace {

struct corner { int v = -1, vt = -1, vn = -1; };
struct obj_shape { std::string name; std::vector<corner> corners; std::vector<int> material_ids; };   // 3 corners per face
struct obj_material { std::string name; float kd[3] = {0, 0, 0}; float ke[3] = {0, 0, 0}; std::string map_kd, map_ke; };   // tinyobjloader's InitMaterial: black

// Polygons become triangles the way the reference's loader makes them (tinyobjloader 2.0.0rc, "simple" method,
// tiny_obj_loader.h:1509-1964; float arithmetic): which triangles come out decides face numbering and shading, so a fan is
// not good enough.  Found by tools/fuzz_obj.py against the reference's own LoadModelAsPrefab.
//  * a quad is cut along its SHORTER diagonal: (0,1,2)(0,2,3) if |p2-p0|^2 < |p3-p1|^2, else (0,1,3)(1,2,3);
//  * a larger polygon is ear-clipped in the plane of the two axes other than the dominant one of its first non-degenerate
//    corner: starting at vertex 0, a corner (g, g+1, g+2) is an ear if its turn has the sign of the "area" term
//    x0*y1 - y0*x1 (the loader's test, kept literally) and no other remaining vertex is inside it (crossing-number test);
//    an ear is emitted and its middle vertex removed, otherwise g advances; the search gives up after one fruitless
//    round, so a polygon can yield FEWER than n - 2 triangles; the last three vertices are emitted as they are.
int crossing_number(const float* vx, const float* vy, float tx, float ty) {
    int c = 0;
    for (int i = 0, j = 2; i < 3; j = i++)
        if (((vy[i] > ty) != (vy[j] > ty)) && (tx < (vx[j] - vx[i]) * (ty - vy[i]) / (vy[j] - vy[i]) + vx[i])) c = !c;
    return c;
}

void triangulate(const std::vector<corner>& poly, const std::vector<float>& P, std::vector<corner>& out) {
    const size_t n = poly.size();
    auto pos = [&](const corner& c, int a) { return P[3 * (size_t)c.v + a]; };
    if (n == 3) { out.insert(out.end(), poly.begin(), poly.end()); return; }
    if (n == 4) {
        float d02 = 0.0f, d13 = 0.0f;
        for (int a = 0; a < 3; a++) {
            const float e02 = pos(poly[2], a) - pos(poly[0], a), e13 = pos(poly[3], a) - pos(poly[1], a);
            d02 = a == 0 ? e02 * e02 : d02 + e02 * e02;
            d13 = a == 0 ? e13 * e13 : d13 + e13 * e13;
        }
        const int pick[2][6] = {{0, 1, 2, 0, 2, 3}, {0, 1, 3, 1, 2, 3}};
        for (int k = 0; k < 6; k++) out.push_back(poly[pick[d02 < d13 ? 0 : 1][k]]);
        return;
    }
    // the two working axes
    int ax0 = 1, ax1 = 2;
    for (size_t k = 0; k < n; k++) {
        const corner &a = poly[k % n], &b = poly[(k + 1) % n], &c = poly[(k + 2) % n];
        const float e0x = pos(b, 0) - pos(a, 0), e0y = pos(b, 1) - pos(a, 1), e0z = pos(b, 2) - pos(a, 2);
        const float e1x = pos(c, 0) - pos(b, 0), e1y = pos(c, 1) - pos(b, 1), e1z = pos(c, 2) - pos(b, 2);
        const float cx = std::fabs(e0y * e1z - e0z * e1y), cy = std::fabs(e0z * e1x - e0x * e1z), cz = std::fabs(e0x * e1y - e0y * e1x);
        const float eps = std::numeric_limits<float>::epsilon();
        if (cx > eps || cy > eps || cz > eps) {
            if (!(cx > cy && cx > cz)) {
                ax0 = 0;
                if (cz > cx && cz > cy) ax1 = 1;
            }
            break;
        }
    }
    std::vector<corner> rest = poly;
    size_t guess = 0, tries = n, last_count = n;
    while (rest.size() > 3 && tries > 0) {
        const size_t m = rest.size();
        if (guess >= m) guess -= m;
        if (last_count != m) { last_count = m; tries = m; } else tries--;
        corner tri[3];
        float vx[3], vy[3];
        for (int k = 0; k < 3; k++) {
            tri[k] = rest[(guess + k) % m];
            vx[k] = pos(tri[k], ax0);
            vy[k] = pos(tri[k], ax1);
        }
        const float e0x = vx[1] - vx[0], e0y = vy[1] - vy[0], e1x = vx[2] - vx[1], e1y = vy[2] - vy[1];
        const float turn = e0x * e1y - e0y * e1x;
        const float area = (vx[0] * vy[1] - vy[0] * vx[1]) * 0.5f;
        if (turn * area < 0.0f) { guess++; continue; }
        bool inside = false;
        for (size_t o = 3; o < m && !inside; o++) {
            const corner& q = rest[(guess + o) % m];
            inside = crossing_number(vx, vy, pos(q, ax0), pos(q, ax1)) != 0;
        }
        if (inside) { guess++; continue; }
        out.insert(out.end(), tri, tri + 3);
        rest.erase(rest.begin() + (std::ptrdiff_t)((guess + 1) % m));
    }
    if (rest.size() == 3) out.insert(out.end(), rest.begin(), rest.end());
}

std::string dir_of(const std::string& p) {
    size_t s = p.find_last_of("/\\");
    return s == std::string::npos ? std::string(".") : p.substr(0, s);
}
std::string stem_of(const std::string& p) {
    size_t s = p.find_last_of("/\\");
    std::string f = s == std::string::npos ? p : p.substr(s + 1);
    size_t d = f.find_last_of('.');
    return d == std::string::npos ? f : f.substr(0, d);
}
std::string rest_of(std::istringstream& ss) {
    std::string r;
    std::getline(ss, r);
    size_t a = r.find_first_not_of(" \t\r"), b = r.find_last_not_of(" \t\r");
    return a == std::string::npos ? std::string() : r.substr(a, b - a + 1);
}

// The file name of a `map_*` statement the way tinyobjloader finds it (ParseTextureNameAndOption, tiny_obj_loader.h): options
// are skipped with the number of blank-separated arguments each one CONSUMES there (three for -o / -s / -t whatever they are, two
// for -mm, one for the others), and the first thing that is not a known option starts the name, which runs to the end of the
// line, blanks included.
std::string texture_name_of(const std::string& rest) {
    static const struct { const char* name; int args; } options[] = {
        {"-blendu", 1}, {"-blendv", 1}, {"-clamp", 1}, {"-boost", 1}, {"-bm", 1}, {"-o", 3}, {"-s", 3}, {"-t", 3},
        {"-type", 1}, {"-texres", 1}, {"-imfchan", 1}, {"-mm", 2}, {"-colorspace", 1}};
    size_t at = 0;
    auto blanks = [&]() { while (at < rest.size() && (rest[at] == ' ' || rest[at] == '\t')) at++; };
    auto word = [&]() { while (at < rest.size() && rest[at] != ' ' && rest[at] != '\t' && rest[at] != '\r') at++; };
    for (;;) {
        blanks();
        if (at >= rest.size()) return std::string();
        bool option = false;
        for (auto const& o : options) {
            const size_t n = std::strlen(o.name);
            if (rest.compare(at, n, o.name) == 0 && at + n < rest.size() && (rest[at + n] == ' ' || rest[at + n] == '\t')) {
                at += n;
                for (int k = 0; k < o.args; k++) { blanks(); word(); }
                option = true;
                break;
            }
        }
        if (!option) return rest.substr(at);
    }
}

bool load_mtl(const std::string& path, std::vector<obj_material>& out, std::unordered_map<std::string, int>& index) {
    std::ifstream f(path);
    if (!f) return false;
    std::string line;
    obj_material* cur = nullptr;
    bool has_kd = false;              // tinyobjloader keeps this flag for the whole FILE (newmtl does not clear it): kept, it decides colours
    while (std::getline(f, line)) {
        std::istringstream ss(line);
        std::string key;
        if (!(ss >> key) || key[0] == '#') continue;
        if (key == "newmtl") {
            obj_material m;
            ss >> m.name;                                      // parseString: the first word (a name with blanks is cut there, in `usemtl` too)
            index[m.name] = (int)out.size();
            out.push_back(m);
            cur = &out.back();
        } else if (cur && key == "Kd") { ss >> cur->kd[0] >> cur->kd[1] >> cur->kd[2]; has_kd = true; }
        else if (cur && key == "Ke") ss >> cur->ke[0] >> cur->ke[1] >> cur->ke[2];
        else if (cur && key == "map_Kd") {
            cur->map_kd = texture_name_of(rest_of(ss));
            if (!has_kd) cur->kd[0] = cur->kd[1] = cur->kd[2] = 0.6f;          // "a decent default" when a texture comes without Kd
        } else if (cur && key == "map_Ke") cur->map_ke = texture_name_of(rest_of(ss));
    }
    return true;
}

// One `f` corner: v, v/vt, v//vn or v/vt/vn; 1-based, negative = relative to the end.
bool parse_corner(const std::string& tok, size_t nv, size_t nt, size_t nn, corner& c) {
    int idx[3] = {0, 0, 0};
    int field = 0;
    const char* p = tok.c_str();
    while (*p && field < 3) {
        if (*p == '/') { field++; p++; continue; }
        char* end;
        long v = std::strtol(p, &end, 10);
        if (end == p) return false;
        idx[field] = (int)v;
        p = end;
    }
    auto fix = [](int i, size_t n) { return i > 0 ? i - 1 : (i < 0 ? (int)n + i : -1); };
    c.v = fix(idx[0], nv); c.vt = fix(idx[1], nt); c.vn = fix(idx[2], nn);
    return c.v >= 0 && c.v < (int)nv && c.vt < (int)nt && c.vn < (int)nn;
}

struct vertex_key {
    mesh_vertex V;
    bool operator==(vertex_key const& o) const { return !std::memcmp(&V, &o.V, sizeof V); }
};
struct vertex_hash {
    size_t operator()(vertex_key const& k) const {
        const uint32_t* w = reinterpret_cast<const uint32_t*>(&k.V);
        uint64_t h = 1469598103934665603ull;
        for (size_t i = 0; i < sizeof(mesh_vertex) / 4; i++) { h ^= w[i]; h *= 1099511628211ull; }
        return (size_t)h;
    }
};

entity* clone_tree(scene* Scene, entity const* Source, entity* Parent) {     // scene.cpp:206-251
    entity* E = new entity(*Source);
    E->Parent = Parent;
    E->Children.clear();
    E->PackedShapeIndex = SHAPE_INDEX_NONE;
    if (Parent) Parent->Children.push_back(E);
    for (entity* C : Source->Children) clone_tree(Scene, C, E);
    return E;
}
void delete_tree(entity* E) {
    for (entity* C : E->Children) delete_tree(C);
    delete E;
}

}  // namespace

// LoadTexture (scene.cpp:294-318): stbi_loadf's contract (linear RGBA32F, row 0 first) through host/image.cpp
// (PNG, JPEG, Radiance HDR; plus PPM / PFM).  The default name is the file name, as in the reference.
texture* LoadTexture(scene* Scene, char const* Path, texture_type Type, char const* Name) {
    int w = 0, h = 0;
    std::vector<vec4> px;
    if (!DecodeImageFile(Path, w, h, px)) return nullptr;
    const std::string p = Path;
    const size_t slash = p.find_last_of("/\\");
    const std::string file_name = slash == std::string::npos ? p : p.substr(slash + 1);
    return CreateTexture(Scene, Name ? Name : file_name.c_str(), Type, (uint32_t)w, (uint32_t)h, px.data());
}

prefab* LoadModelAsPrefab(scene* Scene, char const* Path, load_model_options* Options) {
    load_model_options Default;
    if (!Options) Options = &Default;
    std::ifstream file(Path);
    if (!file) return nullptr;
    const std::string Directory = Options->DirectoryPath.empty() ? dir_of(Path) : Options->DirectoryPath;

    std::vector<float> Positions, Normals, TexCoords;
    std::vector<obj_shape> Shapes;
    std::vector<obj_material> FileMaterials;
    std::unordered_map<std::string, int> MaterialIndex;
    int CurrentMaterial = -1;
    Shapes.emplace_back();
    std::string line;
    std::vector<corner> poly;
    while (std::getline(file, line)) {
        std::istringstream ss(line);
        std::string key;
        if (!(ss >> key) || key[0] == '#') continue;
        if (key == "v") { float x = 0, y = 0, z = 0; ss >> x >> y >> z; Positions.insert(Positions.end(), {x, y, z}); }
        else if (key == "vn") { float x = 0, y = 0, z = 0; ss >> x >> y >> z; Normals.insert(Normals.end(), {x, y, z}); }
        else if (key == "vt") { float u = 0, v = 0; ss >> u >> v; TexCoords.insert(TexCoords.end(), {u, v}); }
        else if (key == "f") {
            poly.clear();
            std::string tok;
            bool good = true;
            while (ss >> tok) {
                corner c;
                if (!parse_corner(tok, Positions.size() / 3, TexCoords.size() / 2, Normals.size() / 3, c)) { good = false; break; }
                poly.push_back(c);
            }
            if (!good || poly.size() < 3) continue;
            obj_shape& S = Shapes.back();
            const size_t before = S.corners.size();
            triangulate(poly, Positions, S.corners);
            for (size_t k = before; k < S.corners.size(); k += 3) S.material_ids.push_back(CurrentMaterial);
        } else if (key == "o" || key == "g") {
            // tinyobjloader takes `g` / `o` as a statement only when a blank FOLLOWS the letter: a bare "g" (or "g\r") is skipped
            const size_t at = line.find_first_not_of(" \t");
            if (at + 1 >= line.size() || (line[at + 1] != ' ' && line[at + 1] != '\t')) continue;
            std::string name = rest_of(ss);
            if (Shapes.back().corners.empty()) Shapes.back().name = name;
            else { Shapes.emplace_back(); Shapes.back().name = name; }
        } else if (key == "usemtl") {
            std::string name;
            ss >> name;
            auto it = MaterialIndex.find(name);
            CurrentMaterial = it == MaterialIndex.end() ? -1 : it->second;
        } else if (key == "mtllib") {
            // tinyobjloader: several file names separated by blanks (a backslash escapes one), the first that loads wins
            const std::string all = rest_of(ss);
            std::vector<std::string> names;
            std::string cur_name;
            bool escaping = false;
            for (char ch : all) {
                if (escaping) escaping = false;
                else if (ch == '\\') { escaping = true; continue; }
                else if (ch == ' ') { if (!cur_name.empty()) names.push_back(cur_name); cur_name.clear(); continue; }
                cur_name += ch;
            }
            names.push_back(cur_name);
            for (std::string const& n : names)
                if (load_mtl(Directory + "/" + n, FileMaterials, MaterialIndex)) break;
        }
    }

    // scene.cpp:615-663: no normals in the file -> sum of unit face normals per position, normalised.
    if (Normals.empty()) {
        Normals.assign(Positions.size(), 0.0f);
        for (obj_shape& S : Shapes) {
            for (size_t i = 0; i + 2 < S.corners.size(); i += 3) {
                vec3 P[3];
                for (int j = 0; j < 3; j++) { const float* p = &Positions[3 * (size_t)S.corners[i + j].v]; P[j] = {p[0], p[1], p[2]}; }
                vec3 C = cross3(P[1] - P[0], P[2] - P[0]);
                vec3 N = C * (1.0f / std::sqrt(dot3(C, C)));           // glm::normalize is v * inversesqrt(dot(v, v)), not v / length: NaN for a
                                                                       // degenerate face, like the reference
                for (int j = 0; j < 3; j++) {
                    corner& K = S.corners[i + j];
                    K.vn = K.v;
                    Normals[3 * (size_t)K.vn + 0] += N.x; Normals[3 * (size_t)K.vn + 1] += N.y; Normals[3 * (size_t)K.vn + 2] += N.z;
                }
            }
        }
        for (size_t i = 0; i + 2 < Normals.size(); i += 3) {
            float L = length3({Normals[i], Normals[i + 1], Normals[i + 2]});
            if (L > 1e-9f) { Normals[i] /= L; Normals[i + 1] /= L; Normals[i + 2] /= L; }      // EPSILON, common.hpp
            else { Normals[i] = 0; Normals[i + 1] = 0; Normals[i + 2] = 1; }
        }
    }

    // scene.cpp:665-729: scene materials from the file's.
    std::unordered_map<std::string, texture*> TextureMap;
    std::vector<material*> Materials;
    for (obj_material const& FM : FileMaterials) {
        material* M = CreateMaterial(Scene, MATERIAL_TYPE_BASIC_DIFFUSE, FM.name.c_str());
        M->BaseColor = {FM.kd[0], FM.kd[1], FM.kd[2]};
        if (!FM.map_kd.empty()) {
            if (!TextureMap.count(FM.map_kd))
                TextureMap[FM.map_kd] = LoadTexture(Scene, (Directory + "/" + FM.map_kd).c_str(), TEXTURE_TYPE_REFLECTANCE_WITH_ALPHA, FM.map_kd.c_str());
            M->BaseTexture = TextureMap[FM.map_kd];
        }
        // The reference loads the emission map as well (scene.cpp:703-707), as a radiance texture of its OpenPBR material.  BasicDiffuse
        // has nowhere to attach it, but it is loaded all the same: the textures of a scene, their order and their atlas placements
        // stay the reference's.
        if (!FM.map_ke.empty() && !TextureMap.count(FM.map_ke))
            TextureMap[FM.map_ke] = LoadTexture(Scene, (Directory + "/" + FM.map_ke).c_str(), TEXTURE_TYPE_RADIANCE, FM.map_ke.c_str());
        Materials.push_back(M);
    }

    const std::string ModelName = Options->Name ? std::string(Options->Name) : stem_of(Path);

    // scene.cpp:743-781: shape origins and (shape, material) pairs.
    struct pair_t { size_t Shape; int Material; };
    std::vector<pair_t> Pairs;
    std::vector<vec3> Origins(Shapes.size());
    for (size_t si = 0; si < Shapes.size(); si++) {
        obj_shape const& S = Shapes[si];
        if (S.corners.empty()) continue;
        vec3 Lo{+INFINITY, +INFINITY, +INFINITY}, Hi{-INFINITY, -INFINITY, -INFINITY};
        for (corner const& K : S.corners) { const float* p = &Positions[3 * (size_t)K.v]; Lo = min3(Lo, {p[0], p[1], p[2]}); Hi = max3(Hi, {p[0], p[1], p[2]}); }
        Origins[si] = 0.5f * (Lo + Hi);
        std::vector<int> Seen;
        for (int id : S.material_ids) if (std::find(Seen.begin(), Seen.end(), id) == Seen.end()) { Seen.push_back(id); Pairs.push_back({si, id}); }
    }

    // scene.cpp:783-850: one welded mesh per pair.
    struct pending { std::string Name; std::vector<mesh_vertex> V; std::vector<mesh_face> F; material* Material; size_t Shape; };
    std::vector<pending> Pending;
    for (pair_t const& P : Pairs) {
        obj_shape const& S = Shapes[P.Shape];
        pending M;
        M.Shape = P.Shape;
        M.Material = P.Material >= 0 ? Materials[P.Material] : Options->DefaultMaterial;
        M.Name = !S.name.empty() ? S.name : ModelName + " " + std::to_string(P.Shape);
        const vec3 Origin = Origins[P.Shape];
        std::unordered_map<vertex_key, uint32_t, vertex_hash> Weld;
        for (size_t f = 0; f < S.material_ids.size(); f++) {
            if (S.material_ids[f] != P.Material) continue;
            mesh_face Face;
            for (int j = 0; j < 3; j++) {
                corner const& K = S.corners[3 * f + j];
                vertex_key Key;
                std::memset(&Key.V, 0, sizeof Key.V);
                const float* p = &Positions[3 * (size_t)K.v];
                vec4 Q = mul(Options->VertexTransform, vec4{p[0] - Origin.x, p[1] - Origin.y, p[2] - Origin.z, 1.0f});
                Key.V.Position = {Q.x, Q.y, Q.z};
                if (K.vn >= 0) {
                    const float* n = &Normals[3 * (size_t)K.vn];
                    vec4 N = mul(Options->NormalTransform, vec4{n[0], n[1], n[2], 0.0f});
                    Key.V.Normal = {N.x, N.y, N.z};
                }
                if (K.vt >= 0) {
                    const float* t = &TexCoords[2 * (size_t)K.vt];
                    const float* m = Options->TextureCoordinateTransform;           // column-major 3x3
                    Key.V.UV = {(m[0] * t[0] + m[3] * t[1]) + m[6], (m[1] * t[0] + m[4] * t[1]) + m[7]};
                }
                auto it = Weld.find(Key);
                if (it == Weld.end()) { it = Weld.emplace(Key, (uint32_t)M.V.size()).first; M.V.push_back(Key.V); }
                Face.VertexIndex[j] = it->second;
            }
            M.F.push_back(Face);
        }
        Pending.push_back(std::move(M));
    }

    // scene.cpp:852-866: BuildMeshNode per mesh — independent, so on all host threads; attached in order.
    std::vector<mesh*> Meshes(Pending.size(), nullptr);
    {
        std::atomic<size_t> next(0);
        unsigned T = std::max(1u, std::min((unsigned)Pending.size(), std::thread::hardware_concurrency()));
        std::vector<std::thread> pool;
        for (unsigned t = 0; t < T; t++) pool.emplace_back([&]() {
            for (;;) { size_t i = next.fetch_add(1); if (i >= Pending.size()) break;
                       Meshes[i] = BuildMesh(Pending[i].Name.c_str(), std::move(Pending[i].V), std::move(Pending[i].F)); }
        });
        for (auto& th : pool) th.join();
    }
    for (mesh* M : Meshes) AttachMesh(Scene, M);

    // scene.cpp:871-900
    prefab* Prefab = new prefab;
    auto instance = [&](size_t i) {
        entity* E = new entity;
        E->Type = ENTITY_TYPE_MESH_INSTANCE;
        E->Name = Meshes[i]->Name;
        E->Mesh = Meshes[i];
        E->Material = Pending[i].Material;
        return E;
    };
    if (Meshes.size() == 1) Prefab->Entity = instance(0);
    else {
        entity* Container = new entity;
        Container->Type = ENTITY_TYPE_CONTAINER;
        Container->Name = ModelName;
        for (size_t i = 0; i < Meshes.size(); i++) {
            entity* E = instance(i);
            const vec3 O = Origins[Pending[i].Shape];
            vec4 P = mul(Options->VertexTransform, vec4{O.x, O.y, O.z, 1.0f});
            E->Transform.Position = {P.x, P.y, P.z};
            E->Parent = Container;
            Container->Children.push_back(E);
        }
        Prefab->Entity = Container;
    }
    Scene->Prefabs.push_back(Prefab);
    return Prefab;
}

// scene.cpp:253-258: an instance of the prefab's entity tree under Parent (meshes and materials are shared).
entity* CreateEntity(scene* Scene, prefab* Prefab, entity* Parent) {
    if (!Parent) Parent = &Scene->Root;
    entity* E = clone_tree(Scene, Prefab->Entity, Parent);
    Scene->DirtyFlags |= SCENE_DIRTY_SHAPES;
    return E;
}

// scene.cpp:905-910
void DestroyPrefab(scene* Scene, prefab* Prefab) {
    delete_tree(Prefab->Entity);
    Scene->Prefabs.erase(std::remove(Scene->Prefabs.begin(), Scene->Prefabs.end(), Prefab), Scene->Prefabs.end());
    delete Prefab;
}

// ---- image files ------------------------------------------------------------------------
namespace {
uint32_t crc32_of(const unsigned char* p, size_t n, uint32_t crc) {
    static uint32_t table[256];
    static bool init = false;
    if (!init) {
        for (uint32_t i = 0; i < 256; i++) { uint32_t c = i; for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1; table[i] = c; }
        init = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xFF] ^ (crc >> 8);
    return ~crc;
}
void put_be32(std::vector<unsigned char>& v, uint32_t x) { for (int s = 24; s >= 0; s -= 8) v.push_back((unsigned char)(x >> s)); }
bool write_chunk(FILE* f, const char* type, const std::vector<unsigned char>& data) {
    std::vector<unsigned char> buf;
    put_be32(buf, (uint32_t)data.size());
    buf.insert(buf.end(), type, type + 4);
    buf.insert(buf.end(), data.begin(), data.end());
    put_be32(buf, crc32_of(buf.data() + 4, buf.size() - 4, 0));
    return std::fwrite(buf.data(), 1, buf.size(), f) == buf.size();
}
}  // namespace

// 8-bit RGBA, row 0 at the top (the layout ptb_film_resolve writes).  zlib stream of stored blocks: no compressor needed.
bool WriteImagePNG(char const* Path, uint32_t Width, uint32_t Height, uint8_t const* RGBA) {
    FILE* f = std::fopen(Path, "wb");
    if (!f) return false;
    static const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    bool ok = std::fwrite(sig, 1, 8, f) == 8;
    std::vector<unsigned char> ihdr;
    put_be32(ihdr, Width); put_be32(ihdr, Height);
    ihdr.insert(ihdr.end(), {8, 6, 0, 0, 0});          // 8 bits, RGBA, deflate, adaptive filter set, no interlace
    ok = ok && write_chunk(f, "IHDR", ihdr);
    ok = ok && write_chunk(f, "sRGB", {0});
    std::vector<unsigned char> raw;
    raw.reserve((size_t)Height * (1 + (size_t)Width * 4));
    for (uint32_t y = 0; y < Height; y++) { raw.push_back(0); raw.insert(raw.end(), RGBA + (size_t)y * Width * 4, RGBA + (size_t)(y + 1) * Width * 4); }
    std::vector<unsigned char> z = {0x78, 0x01};
    uint32_t a = 1, b = 0;
    for (size_t pos = 0; pos < raw.size() || pos == 0;) {
        const size_t n = std::min<size_t>(65535, raw.size() - pos);
        z.push_back(pos + n >= raw.size() ? 1 : 0);
        z.push_back((unsigned char)(n & 0xFF)); z.push_back((unsigned char)(n >> 8));
        z.push_back((unsigned char)(~n & 0xFF)); z.push_back((unsigned char)((~n >> 8) & 0xFF));
        z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
        for (size_t i = 0; i < n; i++) { a = (a + raw[pos + i]) % 65521u; b = (b + a) % 65521u; }
        pos += n;
        if (n == 0) break;
    }
    put_be32(z, (b << 16) | a);
    ok = ok && write_chunk(f, "IDAT", z);
    ok = ok && write_chunk(f, "IEND", {});
    return std::fclose(f) == 0 && ok;
}

// Linear RGB float, row 0 at the top in memory (PFM stores rows bottom to top).
bool WriteImagePFM(char const* Path, uint32_t Width, uint32_t Height, float const* RGB) {
    FILE* f = std::fopen(Path, "wb");
    if (!f) return false;
    std::fprintf(f, "PF\n%u %u\n-1.0\n", Width, Height);
    bool ok = true;
    for (uint32_t y = 0; y < Height && ok; y++)
        ok = std::fwrite(RGB + (size_t)(Height - 1 - y) * Width * 3, 4, (size_t)Width * 3, f) == (size_t)Width * 3;
    return std::fclose(f) == 0 && ok;
}

}}  // namespace ptb::host
