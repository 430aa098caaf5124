// Host scene layer, part 5: LoadScene / SaveScene.
//
// Same on-disk format as /root/reference/src/scene/serializer.cpp (cited per item): one JSON
// document {"Textures", "Materials", "Meshes", "Prefabs", "Root"} (:414-483) next to side files
// `<name>.texture` (:176-220), `<name>.mesh` (:271-313) and `spectrum.dat` (:485-516), whose bulk
// arrays are zlib streams prefixed with their compressed size (:16-46).  The reference uses
// nlohmann::json and miniz (both vendored there, neither copied here): the JSON subset the format
// needs is read and written by the small value class below, the streams by the system zlib.
//
// Where the reference's format has holes, the file stays readable by the reference and carries
// the missing data in places its reader ignores:
//   * mesh VERTICES are not saved by the reference (:294-295 writes faces and nodes only), so a
//     scene it saved cannot be rendered after loading.  Here a third block {u32 count, vertices}
//     follows the two the reference reads; a file without it fails to load with a clear error.
//   * the size prefix is `mz_ulong` = unsigned long (:18,27): 4 bytes where the reference builds
//     (MSVC), 8 on LP64.  Written as 4 bytes (8 with PTB_SCENE_FILE_SIZE_PREFIX=8), read as either (told apart by plausibility).
//   * Root.SkyboxSamplingProbability is not saved (:342-349); written here as an extra key.
#include "scene.hpp"

#include <zlib.h>

#include <algorithm>
#include <cctype>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <unordered_map>

namespace ptb { namespace host {

namespace {

// ---- minimal JSON ----------------------------------------------------------------------
struct json {
    enum kind { Null, Bool, Number, String, Array, Object } k = Null;
    bool b = false;
    double num = 0.0;
    bool integral = false;
    std::string str;
    std::vector<json> arr;
    std::map<std::string, json> obj;               // sorted keys, like nlohmann's default object

    json() {}
    json(bool v) : k(Bool), b(v) {}
    json(int v) : k(Number), num(v), integral(true) {}
    json(uint32_t v) : k(Number), num(v), integral(true) {}
    json(float v) : k(Number), num(v) {}
    json(const std::string& v) : k(String), str(v) {}
    json(const vec2& v) : k(Array) { arr = {json(v.x), json(v.y)}; }
    json(const vec3& v) : k(Array) { arr = {json(v.x), json(v.y), json(v.z)}; }

    json& operator[](const char* key) { if (k == Null) k = Object; return obj[key]; }
    const json& at(const char* key) const {
        static const json none;
        auto it = obj.find(key);
        return it == obj.end() ? none : it->second;
    }
    const json& at(size_t i) const { static const json none; return i < arr.size() ? arr[i] : none; }
    size_t size() const { return k == Array ? arr.size() : k == Object ? obj.size() : 0; }
    void push(json v) { k = Array; arr.push_back(std::move(v)); }
    float f(float dflt = 0.0f) const { return k == Number ? (float)num : dflt; }
    int i(int dflt = 0) const { return k == Number ? (int)num : k == Bool ? (int)b : dflt; }
    bool boolean(bool dflt = false) const { return k == Bool ? b : k == Number ? num != 0 : dflt; }
    vec3 v3(vec3 d = {0, 0, 0}) const { return size() >= 3 ? vec3{arr[0].f(), arr[1].f(), arr[2].f()} : d; }
    vec2 v2(vec2 d = {0, 0}) const { return size() >= 2 ? vec2{arr[0].f(), arr[1].f()} : d; }

    void dump(std::string& out, int indent, int depth = 0) const {
        auto nl = [&](int d) { out += '\n'; out.append((size_t)indent * d, ' '); };
        switch (k) {
        case Null: out += "null"; break;
        case Bool: out += b ? "true" : "false"; break;
        case Number: {
            char buf[40];
            if (integral) { auto r = std::to_chars(buf, buf + sizeof buf, (long long)num); out.append(buf, r.ptr); }
            else if (!std::isfinite(num)) out += "null";
            else {
                auto r = std::to_chars(buf, buf + sizeof buf, num);       // shortest round-trip form
                std::string s(buf, r.ptr);
                if (s.find_first_of(".eE") == std::string::npos) s += ".0";
                out += s;
            }
            break;
        }
        case String:
            out += '"';
            for (char c : str) {
                if (c == '"' || c == '\\') { out += '\\'; out += c; }
                else if (c == '\n') out += "\\n";
                else if (c == '\t') out += "\\t";
                else if ((unsigned char)c < 0x20) { char b4[8]; std::snprintf(b4, sizeof b4, "\\u%04x", c); out += b4; }
                else out += c;
            }
            out += '"';
            break;
        case Array:
            if (arr.empty()) { out += "[]"; break; }
            out += '[';
            for (size_t n = 0; n < arr.size(); n++) { nl(depth + 1); arr[n].dump(out, indent, depth + 1); if (n + 1 < arr.size()) out += ','; }
            nl(depth); out += ']';
            break;
        case Object: {
            if (obj.empty()) { out += "{}"; break; }
            out += '{';
            size_t n = 0;
            for (auto& kv : obj) { nl(depth + 1); json(kv.first).dump(out, indent); out += ": "; kv.second.dump(out, indent, depth + 1); if (++n < obj.size()) out += ','; }
            nl(depth); out += '}';
            break;
        }
        }
    }
};

struct parser {
    const char* p; const char* end;
    void ws() { while (p < end && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) p++; }
    [[noreturn]] void fail(const char* what) { throw std::runtime_error(std::string("scene file: ") + what); }
    json value() {
        ws();
        if (p >= end) fail("unexpected end");
        if (*p == '{') {
            json j; j.k = json::Object; p++; ws();
            if (p < end && *p == '}') { p++; return j; }
            for (;;) {
                ws(); if (p >= end || *p != '"') fail("expected a key");
                std::string key = string();
                ws(); if (p >= end || *p != ':') fail("expected ':'");
                p++;
                j.obj[key] = value();
                ws(); if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == '}') { p++; return j; }
                fail("expected ',' or '}'");
            }
        }
        if (*p == '[') {
            json j; j.k = json::Array; p++; ws();
            if (p < end && *p == ']') { p++; return j; }
            for (;;) {
                j.arr.push_back(value());
                ws(); if (p < end && *p == ',') { p++; continue; }
                if (p < end && *p == ']') { p++; return j; }
                fail("expected ',' or ']'");
            }
        }
        if (*p == '"') return json(string());
        if (!std::strncmp(p, "true", 4)) { p += 4; return json(true); }
        if (!std::strncmp(p, "false", 5)) { p += 5; return json(false); }
        if (!std::strncmp(p, "null", 4)) { p += 4; return json(); }
        const char* s = p;
        while (p < end && (std::isdigit((unsigned char)*p) || *p == '-' || *p == '+' || *p == '.' || *p == 'e' || *p == 'E')) p++;
        if (s == p) fail("unexpected character");
        json j; j.k = json::Number;
        std::string t(s, p);
        j.num = std::strtod(t.c_str(), nullptr);
        j.integral = t.find_first_of(".eE") == std::string::npos;
        return j;
    }
    std::string string() {
        std::string out;
        p++;
        while (p < end && *p != '"') {
            if (*p == '\\' && p + 1 < end) {
                p++;
                switch (*p) {
                case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
                case 'b': out += '\b'; break; case 'f': out += '\f'; break;
                case 'u': { if (p + 4 >= end) fail("bad escape"); unsigned c = (unsigned)std::strtoul(std::string(p + 1, p + 5).c_str(), nullptr, 16);
                            if (c < 0x80) out += (char)c; else if (c < 0x800) { out += (char)(0xC0 | (c >> 6)); out += (char)(0x80 | (c & 0x3F)); }
                            else { out += (char)(0xE0 | (c >> 12)); out += (char)(0x80 | ((c >> 6) & 0x3F)); out += (char)(0x80 | (c & 0x3F)); }
                            p += 4; break; }
                default: out += *p;
                }
                p++;
            } else out += *p++;
        }
        if (p >= end) fail("unterminated string");
        p++;
        return out;
    }
};

// ---- compressed blocks (serializer.cpp:16-46) ----------------------------------------------
void write_compressed(std::ostream& out, const void* data, size_t size) {
    uLongf n = compressBound((uLong)size);
    std::vector<unsigned char> buf(n);
    if (compress(buf.data(), &n, (const Bytef*)data, (uLong)size) != Z_OK) throw std::runtime_error("scene file: compression failed");
    // PTB_SCENE_FILE_SIZE_PREFIX=8: the prefix an LP64 build of the reference writes and expects (tests exchange files with one).
    const char* width = std::getenv("PTB_SCENE_FILE_SIZE_PREFIX");
    if (width && std::atoi(width) == 8) { const uint64_t n64 = (uint64_t)n; out.write((const char*)&n64, 8); }
    else { const uint32_t n32 = (uint32_t)n; out.write((const char*)&n32, 4); }
    out.write((const char*)buf.data(), n);
}
void read_compressed(std::istream& in, void* data, size_t size, size_t file_size) {
    const std::streampos at = in.tellg();
    uint64_t n64 = 0;
    in.read((char*)&n64, 8);
    uint64_t n = n64;
    const uint64_t left = file_size - (uint64_t)at;
    if (!in || n64 + 8 > left) {                       // 4-byte prefix (MSVC unsigned long)
        in.clear();
        in.seekg(at);
        uint32_t n32 = 0;
        in.read((char*)&n32, 4);
        n = n32;
        if (!in || n + 4 > left) throw std::runtime_error("scene file: truncated compressed block");
    }
    std::vector<unsigned char> buf(n);
    in.read((char*)buf.data(), (std::streamsize)n);
    uLongf got = (uLongf)size;
    if (!in || uncompress((Bytef*)data, &got, buf.data(), (uLong)n) != Z_OK || got != size) throw std::runtime_error("scene file: corrupt compressed block");
}
size_t file_size_of(std::ifstream& f) { f.seekg(0, std::ios::end); size_t n = (size_t)f.tellg(); f.seekg(0); return n; }

std::string make_file_name(const std::string& name, const char* ext) {      // serializer.cpp:48-61
    std::string s = name;
    for (char& c : s) if (!std::isalnum((unsigned char)c)) c = '_';
    return s + "." + ext;
}
std::string dir_of(const std::string& p) {
    size_t s = p.find_last_of("/\\");
    return s == std::string::npos ? std::string(".") : p.substr(0, s);
}

struct header { uint32_t Magic, Version, A, B; };
constexpr uint32_t MAGIC_TEX = 0x54455820u, MAGIC_MESH = 0x4D455348u, MAGIC_SPEC = 0x53504543u;   // 'TEX ', 'MESH', 'SPEC'

struct context {
    scene* Scene; std::string Dir; bool Writing;
    std::unordered_map<const texture*, int> TextureIndex;
    std::unordered_map<const material*, int> MaterialIndex;
    std::unordered_map<const mesh*, int> MeshIndex;
    int index(const texture* t) { auto it = TextureIndex.find(t); return it == TextureIndex.end() ? -1 : it->second; }
    int index(const material* m) { auto it = MaterialIndex.find(m); return it == MaterialIndex.end() ? -1 : it->second; }
    int index(const mesh* m) { auto it = MeshIndex.find(m); return it == MeshIndex.end() ? -1 : it->second; }
    texture* tex(const json& j) { int i = j.i(-1); return i >= 0 && i < (int)Scene->Textures.size() ? Scene->Textures[i] : nullptr; }
    material* mat(const json& j) { int i = j.i(-1); return i >= 0 && i < (int)Scene->Materials.size() ? Scene->Materials[i] : nullptr; }
    mesh* msh(const json& j) { int i = j.i(-1); return i >= 0 && i < (int)Scene->Meshes.size() ? Scene->Meshes[i] : nullptr; }
};

// serializer.cpp:176-220
void save_texture(context& C, json& J, texture& T) {
    J["Type"] = json((int)T.Type); J["Name"] = json(T.Name); J["EnableNearestFiltering"] = json(T.EnableNearestFiltering);
    std::ofstream f(C.Dir + "/" + make_file_name(T.Name, "texture"), std::ios::binary);
    header H{MAGIC_TEX, 0, T.Width, T.Height};
    f.write((const char*)&H, sizeof H);
    write_compressed(f, T.Pixels.data(), sizeof(vec4) * T.Pixels.size());
    if (!f) throw std::runtime_error("scene file: cannot write texture " + T.Name);
}
// An enumerator read from a file must be one the packer knows (a corrupt file otherwise reaches PackSceneData with it).
// Counts: scene.hpp:10-46,230-240 (TEXTURE_TYPE__COUNT 3, MATERIAL_TYPE__COUNT 4, CAMERA_MODEL__COUNT 3, ENTITY_TYPE__COUNT 7).
int checked_enum(const json& J, int count, const char* what) {
    const int v = (int)J.i();
    if (v < 0 || v >= count) throw std::runtime_error(std::string("scene file: unknown ") + what);
    return v;
}
void load_texture(context& C, const json& J, texture& T) {
    T.Type = (texture_type)checked_enum(J.at("Type"), 3, "texture type"); T.Name = J.at("Name").str; T.EnableNearestFiltering = J.at("EnableNearestFiltering").boolean();
    std::ifstream f(C.Dir + "/" + make_file_name(T.Name, "texture"), std::ios::binary);
    if (!f) throw std::runtime_error("scene file: missing texture file for " + T.Name);
    const size_t fs = file_size_of(f);
    header H{};
    f.read((char*)&H, sizeof H);
    if (!f || H.Magic != MAGIC_TEX) throw std::runtime_error("scene file: bad texture file for " + T.Name);
    T.Width = H.A; T.Height = H.B;
    T.Pixels.resize((size_t)T.Width * T.Height);
    read_compressed(f, T.Pixels.data(), sizeof(vec4) * T.Pixels.size(), fs);
}

// serializer.cpp:238-246 + the per-type field lists (basic_diffuse.hpp:44-56, basic_metal.hpp:74-92, basic_translucent.hpp:87-105)
void save_material(context& C, json& J, material& M) {
    J["Type"] = json((int)M.Type); J["Name"] = json(M.Name); J["Flags"] = json(0); J["Opacity"] = json(1.0f);
    if (M.Type == MATERIAL_TYPE_BASIC_DIFFUSE || M.Type == MATERIAL_TYPE_BASIC_METAL) { J["BaseColor"] = json(M.BaseColor); J["BaseTexture"] = json(C.index(M.BaseTexture)); }
    if (M.Type == MATERIAL_TYPE_BASIC_METAL) { J["SpecularColor"] = json(M.SpecularColor); J["SpecularTexture"] = json(C.index(M.SpecularTexture)); }
    if (M.Type == MATERIAL_TYPE_BASIC_METAL || M.Type == MATERIAL_TYPE_BASIC_TRANSLUCENT) {
        J["Roughness"] = json(M.Roughness); J["RoughnessTexture"] = json(C.index(M.RoughnessTexture));
        J["RoughnessAnisotropy"] = json(M.RoughnessAnisotropy); J["RoughnessAnisotropyTexture"] = json(C.index(M.RoughnessAnisotropyTexture));
    }
    if (M.Type == MATERIAL_TYPE_BASIC_TRANSLUCENT) {
        J["IOR"] = json(M.IOR); J["AbbeNumber"] = json(M.AbbeNumber); J["TransmissionColor"] = json(M.TransmissionColor);
        J["TransmissionDepth"] = json(M.TransmissionDepth); J["ScatteringColor"] = json(M.ScatteringColor); J["ScatteringAnisotropy"] = json(M.ScatteringAnisotropy);
    }
}
void load_material(context& C, const json& J, material& M) {
    material D;
    M.Name = J.at("Name").str;
    M.BaseColor = J.at("BaseColor").v3(D.BaseColor); M.BaseTexture = C.tex(J.at("BaseTexture"));
    M.SpecularColor = J.at("SpecularColor").v3(D.SpecularColor); M.SpecularTexture = C.tex(J.at("SpecularTexture"));
    M.Roughness = J.at("Roughness").f(D.Roughness); M.RoughnessTexture = C.tex(J.at("RoughnessTexture"));
    M.RoughnessAnisotropy = J.at("RoughnessAnisotropy").f(D.RoughnessAnisotropy); M.RoughnessAnisotropyTexture = C.tex(J.at("RoughnessAnisotropyTexture"));
    M.IOR = J.at("IOR").f(D.IOR); M.AbbeNumber = J.at("AbbeNumber").f(D.AbbeNumber);
    M.TransmissionColor = J.at("TransmissionColor").v3(D.TransmissionColor); M.TransmissionDepth = J.at("TransmissionDepth").f(D.TransmissionDepth);
    M.ScatteringColor = J.at("ScatteringColor").v3(D.ScatteringColor); M.ScatteringAnisotropy = J.at("ScatteringAnisotropy").f(D.ScatteringAnisotropy);
}

// serializer.cpp:271-313; the third block (vertices) is this implementation's addition.
void save_mesh(context& C, json& J, mesh& M) {
    J["Name"] = json(M.Name);
    std::ofstream f(C.Dir + "/" + make_file_name(M.Name, "mesh"), std::ios::binary);
    header H{MAGIC_MESH, 0, (uint32_t)M.Faces.size(), (uint32_t)M.Nodes.size()};
    f.write((const char*)&H, sizeof H);
    static_assert(sizeof(mesh_face) == 12 && sizeof(mesh_node) == 36 && sizeof(mesh_vertex) == 32, "reference record sizes (scene.hpp:199-217)");
    write_compressed(f, M.Faces.data(), sizeof(mesh_face) * M.Faces.size());
    write_compressed(f, M.Nodes.data(), sizeof(mesh_node) * M.Nodes.size());
    const uint32_t nv = (uint32_t)M.Vertices.size();
    f.write((const char*)&nv, 4);
    write_compressed(f, M.Vertices.data(), sizeof(mesh_vertex) * M.Vertices.size());
    if (!f) throw std::runtime_error("scene file: cannot write mesh " + M.Name);
}
void load_mesh(context& C, const json& J, mesh& M) {
    M.Name = J.at("Name").str;
    std::ifstream f(C.Dir + "/" + make_file_name(M.Name, "mesh"), std::ios::binary);
    if (!f) throw std::runtime_error("scene file: missing mesh file for " + M.Name);
    const size_t fs = file_size_of(f);
    header H{};
    f.read((char*)&H, sizeof H);
    if (!f || H.Magic != MAGIC_MESH) throw std::runtime_error("scene file: bad mesh file for " + M.Name);
    M.Faces.resize(H.A); M.Nodes.resize(H.B);
    read_compressed(f, M.Faces.data(), sizeof(mesh_face) * M.Faces.size(), fs);
    read_compressed(f, M.Nodes.data(), sizeof(mesh_node) * M.Nodes.size(), fs);
    uint32_t nv = 0;
    f.read((char*)&nv, 4);
    if (!f) throw std::runtime_error("scene file: mesh '" + M.Name + "' has no vertex block (files saved by the reference do not store vertices, serializer.cpp:294-295)");
    M.Vertices.resize(nv);
    read_compressed(f, M.Vertices.data(), sizeof(mesh_vertex) * M.Vertices.size(), fs);
    for (auto const& F : M.Faces) for (uint32_t v : F.VertexIndex) if (v >= nv) throw std::runtime_error("scene file: face index out of range in " + M.Name);
    M.Depth = 0;
    std::vector<std::pair<uint32_t, uint32_t>> stack{{0u, 0u}};            // depth is not stored: recompute
    while (!stack.empty() && !M.Nodes.empty()) {
        auto [n, d] = stack.back(); stack.pop_back();
        M.Depth = std::max(M.Depth, d);
        const uint32_t c = M.Nodes[n].ChildNodeIndex;
        if (c > 0 && c + 1 < M.Nodes.size()) { stack.push_back({c, d + 1}); stack.push_back({c + 1, d + 1}); }
    }
}

// serializer.cpp:332-406
void save_entity(context& C, json& J, entity& E) {
    J["Type"] = json((int)E.Type); J["Position"] = json(E.Transform.Position); J["Rotation"] = json(E.Transform.Rotation); J["Scale"] = json(E.Transform.Scale);
    J["Name"] = json(E.Name); J["Active"] = json(E.Active); J["Material"] = json(C.index(E.Material));
    if (E.Type == ENTITY_TYPE_ROOT) {
        J["ScatterRate"] = json(E.ScatterRate); J["SkyboxBrightness"] = json(E.SkyboxBrightness); J["SkyboxTexture"] = json(C.index(E.SkyboxTexture));
        J["SkyboxSamplingProbability"] = json(E.SkyboxSamplingProbability);
    } else if (E.Type == ENTITY_TYPE_CAMERA) {
        J["CameraModel"] = json((int)E.CameraModel);
        J["Pinhole"]["FieldOfViewInDegrees"] = json(E.Pinhole.FieldOfViewInDegrees); J["Pinhole"]["ApertureDiameterInMM"] = json(E.Pinhole.ApertureDiameterInMM);
        J["ThinLens"]["SensorSizeInMM"] = json(E.ThinLens.SensorSizeInMM); J["ThinLens"]["FocalLengthInMM"] = json(E.ThinLens.FocalLengthInMM);
        J["ThinLens"]["ApertureDiameterInMM"] = json(E.ThinLens.ApertureDiameterInMM); J["ThinLens"]["FocusDistance"] = json(E.ThinLens.FocusDistance);
    } else if (E.Type == ENTITY_TYPE_MESH_INSTANCE) {
        J["Mesh"] = json(C.index(E.Mesh));
    }
    json& K = J["Children"];                           // stays null when there are none, like the reference's dump
    for (entity* Child : E.Children) { json V; save_entity(C, V, *Child); K.push(std::move(V)); }
}
void load_entity(context& C, const json& J, entity& E) {
    entity D;
    E.Type = (entity_type)checked_enum(J.at("Type"), 7, "entity type");
    E.Transform.Position = J.at("Position").v3(); E.Transform.Rotation = J.at("Rotation").v3(); E.Transform.Scale = J.at("Scale").v3({1, 1, 1});
    E.Name = J.at("Name").str; E.Active = J.at("Active").boolean(true); E.Material = C.mat(J.at("Material"));
    if (E.Type == ENTITY_TYPE_ROOT) {
        E.ScatterRate = J.at("ScatterRate").f(); E.SkyboxBrightness = J.at("SkyboxBrightness").f(1.0f); E.SkyboxTexture = C.tex(J.at("SkyboxTexture"));
        E.SkyboxSamplingProbability = J.at("SkyboxSamplingProbability").f(0.0f);
    } else if (E.Type == ENTITY_TYPE_CAMERA) {
        E.CameraModel = (camera_model)checked_enum(J.at("CameraModel"), 3, "camera model");
        E.Pinhole.FieldOfViewInDegrees = J.at("Pinhole").at("FieldOfViewInDegrees").f(D.Pinhole.FieldOfViewInDegrees);
        E.Pinhole.ApertureDiameterInMM = J.at("Pinhole").at("ApertureDiameterInMM").f(D.Pinhole.ApertureDiameterInMM);
        E.ThinLens.SensorSizeInMM = J.at("ThinLens").at("SensorSizeInMM").v2(D.ThinLens.SensorSizeInMM);
        E.ThinLens.FocalLengthInMM = J.at("ThinLens").at("FocalLengthInMM").f(D.ThinLens.FocalLengthInMM);
        E.ThinLens.ApertureDiameterInMM = J.at("ThinLens").at("ApertureDiameterInMM").f(D.ThinLens.ApertureDiameterInMM);
        E.ThinLens.FocusDistance = J.at("ThinLens").at("FocusDistance").f(D.ThinLens.FocusDistance);
    } else if (E.Type == ENTITY_TYPE_MESH_INSTANCE) {
        E.Mesh = C.msh(J.at("Mesh"));
    }
    const json& K = J.at("Children");
    for (size_t n = 0; n < K.size(); n++) {
        entity* Child = new entity;
        Child->Parent = &E;
        E.Children.push_back(Child);
        load_entity(C, K.at(n), *Child);
    }
}

}  // namespace

// serializer.cpp:518-529.  Returns false (and a message through *Error when given) on I/O failure.
bool SaveScene(char const* Path, scene* Scene, std::string* Error) {
    try {
        context C{Scene, dir_of(Path), true, {}, {}, {}};
        // Side files are named after the asset (serializer.cpp:48-61,195,285): two meshes called "Box" (one shape,
        // two materials) would share one file and the second would overwrite the first.  Names are made unique first.
        auto make_unique = [](auto& assets, const char* ext) {
            std::map<std::string, int> used;
            for (auto* a : assets) {
                std::string base = a->Name;
                for (int n = 2; used.count(make_file_name(a->Name, ext)); n++) a->Name = base + " " + std::to_string(n);
                used[make_file_name(a->Name, ext)] = 1;
            }
        };
        make_unique(Scene->Textures, "texture");
        make_unique(Scene->Meshes, "mesh");
        json J;
        J["Textures"].k = json::Array; J["Materials"].k = json::Array; J["Meshes"].k = json::Array; J["Prefabs"].k = json::Array;
        for (size_t n = 0; n < Scene->Textures.size(); n++) { C.TextureIndex[Scene->Textures[n]] = (int)n; json V; save_texture(C, V, *Scene->Textures[n]); J["Textures"].push(std::move(V)); }
        for (size_t n = 0; n < Scene->Materials.size(); n++) { C.MaterialIndex[Scene->Materials[n]] = (int)n; json V; save_material(C, V, *Scene->Materials[n]); J["Materials"].push(std::move(V)); }
        for (size_t n = 0; n < Scene->Meshes.size(); n++) { C.MeshIndex[Scene->Meshes[n]] = (int)n; json V; save_mesh(C, V, *Scene->Meshes[n]); J["Meshes"].push(std::move(V)); }
        for (prefab* P : Scene->Prefabs) { json V; save_entity(C, V, *P->Entity); J["Prefabs"].push(std::move(V)); }
        save_entity(C, J["Root"], Scene->Root);
        std::string text;
        J.dump(text, 4);
        std::ofstream f(Path);
        f << text;
        if (!f) throw std::runtime_error(std::string("scene file: cannot write ") + Path);
        if (Scene->RGBSpectrumTable) {                                      // serializer.cpp:485-516
            std::ofstream s(C.Dir + "/spectrum.dat", std::ios::binary);
            const uint32_t H[2] = {MAGIC_SPEC, 0};
            s.write((const char*)H, 8);
            write_compressed(s, Scene->RGBSpectrumTable->Coefficients, sizeof(Scene->RGBSpectrumTable->Coefficients));
        }
        return true;
    } catch (std::exception const& e) { if (Error) *Error = e.what(); return false; }
}

// serializer.cpp:503-516.  The spectrum table is the process-wide one (the file's copy is checked for presence only:
// it is the same table by construction).
scene* LoadScene(char const* Path, std::string* Error) {
    scene* S = nullptr;
    try {
        std::ifstream f(Path);
        if (!f) throw std::runtime_error(std::string("scene file: cannot open ") + Path);
        std::stringstream ss; ss << f.rdbuf();
        const std::string text = ss.str();
        parser P{text.data(), text.data() + text.size()};
        const json J = P.value();
        S = CreateScene(false);
        context C{S, dir_of(Path), false, {}, {}, {}};
        const json &JT = J.at("Textures"), &JM = J.at("Materials"), &JG = J.at("Meshes"), &JP = J.at("Prefabs");
        for (size_t n = 0; n < JT.size(); n++) { texture* T = new texture; S->Textures.push_back(T); load_texture(C, JT.at(n), *T); }
        for (size_t n = 0; n < JM.size(); n++) CreateMaterial(S, (material_type)checked_enum(JM.at(n).at("Type"), 4, "material type"), "");
        for (size_t n = 0; n < JM.size(); n++) load_material(C, JM.at(n), *S->Materials[n]);
        for (size_t n = 0; n < JG.size(); n++) { mesh* M = new mesh; S->Meshes.push_back(M); load_mesh(C, JG.at(n), *M); }
        for (size_t n = 0; n < JP.size(); n++) { prefab* R = new prefab; R->Entity = new entity; S->Prefabs.push_back(R); load_entity(C, JP.at(n), *R->Entity); }
        load_entity(C, J.at("Root"), S->Root);
        S->Root.Type = ENTITY_TYPE_ROOT;
        S->Root.Parent = nullptr;
        S->DirtyFlags = SCENE_DIRTY_ALL;
        return S;
    } catch (std::exception const& e) {
        if (Error) *Error = e.what();
        if (S) DestroyScene(S);
        return nullptr;
    }
}

}}  // namespace ptb::host
