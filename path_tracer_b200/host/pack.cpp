// Host scene layer, part 2: PackSceneData — flattening of the source
// description into the packed arrays the integrator consumes.
// Follows /root/reference/src/scene/scene.cpp:1031-1093 (ShapeBounds) and
// :1115-1621 (PackSceneData); material packers follow basic_diffuse.hpp:17-28,
// basic_metal.hpp:26-52, basic_translucent.hpp:26-61.
#include "scene.hpp"
#include <algorithm>
#include <atomic>
#include <cassert>
#include <thread>
#include <unordered_map>
#include <functional>

namespace ptb { namespace host {

namespace {

inline uint32_t f2u(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
inline uint32_t tex_index(texture const* T) { return T ? T->PackedTextureIndex : TEXTURE_INDEX_NONE; }   // scene.hpp:391-395
inline uint32_t mat_index(material const* M) { return M ? M->PackedMaterialIndex : 0u; }                 // scene.hpp:397-401

void store_mat4(float* dst, mat4 const& m) { std::memcpy(dst, &m, 64); }
ptb_packed_transform pack_transform(mat4 const& m) {     // scene.hpp:403-406
    ptb_packed_transform t;
    store_mat4(t.To, m);
    store_mat4(t.From, inverse(m));
    return t;
}

// Post-order walk with accumulated transforms; inactive entities prune their
// subtree (scene.cpp:127-149).
void walk(entity* E, mat4 const& Outer, std::function<void(entity*, mat4 const&)> const& Fn) {
    if (!E->Active) return;
    mat4 M = mul(Outer, make_transform(E->Transform.Position, E->Transform.Rotation, E->Transform.Scale));
    for (entity* C : E->Children) walk(C, M, Fn);
    Fn(E, M);
}

void put_spectrum(scene* S, uint32_t* A, uint32_t at, vec3 color) {
    vec3 b = GetParametricSpectrumCoefficients(S->RGBSpectrumTable, color);
    A[at + 0] = f2u(b.x); A[at + 1] = f2u(b.y); A[at + 2] = f2u(b.z);
}

// Word offsets are the `const uint` tables at the top of each material's
// .glsl.inc / .hpp pair.
void pack_material(scene* S, material const* M, uint32_t* A) {
    switch (M->Type) {
    case MATERIAL_TYPE_BASIC_DIFFUSE:
        put_spectrum(S, A, 1, M->BaseColor);
        A[4] = tex_index(M->BaseTexture);
        break;
    case MATERIAL_TYPE_BASIC_METAL:
        put_spectrum(S, A, 1, M->BaseColor);       A[4] = tex_index(M->BaseTexture);
        put_spectrum(S, A, 5, M->SpecularColor);   A[8] = tex_index(M->SpecularTexture);
        A[9] = f2u(M->Roughness);                  A[10] = tex_index(M->RoughnessTexture);
        A[11] = f2u(M->RoughnessAnisotropy);       A[12] = tex_index(M->RoughnessAnisotropyTexture);
        break;
    case MATERIAL_TYPE_BASIC_TRANSLUCENT:
        A[1] = f2u(M->IOR);
        A[2] = f2u(M->AbbeNumber);
        A[3] = f2u(M->Roughness);                  A[4] = tex_index(M->RoughnessTexture);
        A[5] = f2u(M->RoughnessAnisotropy);        A[6] = tex_index(M->RoughnessAnisotropyTexture);
        put_spectrum(S, A, 7, M->TransmissionColor);
        A[10] = f2u(M->TransmissionDepth);
        put_spectrum(S, A, 11, M->ScatteringColor);
        A[14] = f2u(M->ScatteringAnisotropy);
        break;
    case MATERIAL_TYPE_OPENPBR:
        // Shader include disabled in the reference (scene.glsl.inc:685): the record only
        // has to occupy its two slots so that slot numbering is unchanged.
        break;
    }
}
inline uint32_t material_words(material_type T) { return T == MATERIAL_TYPE_OPENPBR ? 64u : 32u; }   // scene.hpp:468-483

// ---- texture atlas (scene.cpp:1120-1233) ----
// The reference places the rectangles of one atlas layer with stb_rect_pack (stbrp_init_target with one skyline node
// per atlas column, stbrp_pack_rects with the default heuristic, scene.cpp:1144-1147).  That algorithm, restated:
// rectangles are taken by height descending, then width descending (ties: input order); the skyline is a list of
// (x, height) steps; a rectangle of width w is tried with its left edge on every step whose x + w fits, resting at
// the highest skyline point under it, and goes where it rests LOWEST, leftmost among equals ("bottom-left");
// a rectangle that would rise above the atlas stays unplaced for the next layer.  Placements are equal to the
// reference's (tests/test_reference_host.py compares them with the reference's own PackSceneData).
struct skyline_packer {
    struct step { uint32_t x, y; };
    uint32_t W, H;
    std::vector<step> line;                        // ascending x; the last entry is the sentinel {W, 65535}
    skyline_packer(uint32_t w, uint32_t h) : W(w), H(h) { line.push_back({0u, 0u}); line.push_back({w, 65535u}); }
    bool place(uint32_t w, uint32_t h, uint32_t& ox, uint32_t& oy) {
        if (w > W || h > H) return false;
        size_t best = line.size();
        uint32_t best_y = 1u << 30;
        for (size_t i = 0; i + 1 < line.size() && line[i].x + w <= W; i++) {
            const uint32_t x1 = line[i].x + w;
            uint32_t y = 0;
            for (size_t j = i; line[j].x < x1; j++) y = std::max(y, line[j].y);
            if (y < best_y) { best_y = y; best = i; }
        }
        if (best == line.size() || best_y + h > H) return false;
        ox = line[best].x; oy = best_y;
        // the new step replaces every step it covers; the first step reaching beyond it is cut back to its right edge
        const uint32_t x1 = ox + w;
        size_t k = best;
        while (k + 1 < line.size() && line[k + 1].x <= x1) k++;      // line[k]: the last step starting at or before the right edge
        std::vector<step> next(line.begin(), line.begin() + best);
        next.push_back({ox, best_y + h});
        next.push_back({std::max(line[k].x, x1), line[k].y});       // it survives, cut back to the right edge if it started inside
        next.insert(next.end(), line.begin() + k + 1, line.end());
        line.swap(next);
        return true;
    }
};

void pack_textures(scene* S) {
    S->Images.clear();
    S->TexturePack.clear();
    std::vector<texture*> Pending(S->Textures.begin(), S->Textures.end());
    const uint32_t AW = S->AtlasWidth, AH = S->AtlasHeight;
    while (!Pending.empty()) {
        skyline_packer P(AW, AH);
        // stbrp_pack_rects: by height, then width, both descending; unplaced rectangles wait for the next layer
        std::vector<texture*> Order(Pending);
        std::stable_sort(Order.begin(), Order.end(), [](texture const* A, texture const* B) {
            return A->Height != B->Height ? A->Height > B->Height : A->Width > B->Width; });
        std::unordered_map<texture const*, std::pair<uint32_t, uint32_t>> Spot;
        for (texture* T : Order) {
            uint32_t rx = 0, ry = 0;
            if (T->Width == 0 || T->Height == 0 || P.place(T->Width, T->Height, rx, ry)) Spot[T] = {rx, ry};
        }
        std::vector<float> Layer((size_t)AW * AH * 4, 0.0f);
        const uint32_t ImageIndex = (uint32_t)S->Images.size();
        std::vector<texture*> Left;
        bool any = false;
        struct job { texture const* T; uint32_t rx, ry; };
        std::vector<job> Jobs;
        for (texture* T : Pending) {
            auto Found = Spot.find(T);
            if (Found == Spot.end()) { Left.push_back(T); continue; }
            const uint32_t rx = Found->second.first, ry = Found->second.second;
            any = true;
            T->PackedTextureIndex = (uint32_t)S->TexturePack.size();
            ptb_packed_texture K{};
            K.AtlasImageIndex = ImageIndex;
            // half-texel inset, V flipped: scene.cpp:1168-1177
            K.AtlasPlacementMinimum[0] = (rx + 0.5f) / float(AW);
            K.AtlasPlacementMinimum[1] = (ry + T->Height - 0.5f) / float(AH);
            K.AtlasPlacementMaximum[0] = (rx + T->Width - 0.5f) / float(AW);
            K.AtlasPlacementMaximum[1] = (ry + 0.5f) / float(AH);
            K.Type = (uint32_t)T->Type;       // (left uninitialised by the reference)
            K.Flags = T->EnableNearestFiltering ? (uint32_t)TEXTURE_FLAG_FILTER_NEAREST : 0u;
            Jobs.push_back({T, rx, ry});
            S->TexturePack.push_back(K);
        }
        if (!any) break;   // a texture larger than the atlas: dropped (the reference would loop forever)
        // Texel conversion (scene.cpp:1179-1212): every texel is a pure function of its source texel, rows are dealt to all host threads.
        {
            struct rows { job const* J; uint32_t Y0, Y1; };
            std::vector<rows> Work;
            for (job const& J : Jobs) {
                const uint32_t step = std::max<uint32_t>(1, 65536 / std::max<uint32_t>(1, J.T->Width));
                for (uint32_t Y = 0; Y < J.T->Height; Y += step) Work.push_back({&J, Y, std::min(J.T->Height, Y + step)});
            }
            std::atomic<size_t> Next(0);
            auto Convert = [&]() {
                for (size_t w = Next.fetch_add(1); w < Work.size(); w = Next.fetch_add(1)) {
                    texture const* T = Work[w].J->T;
                    const uint32_t rx = Work[w].J->rx, ry = Work[w].J->ry;
                    for (uint32_t Y = Work[w].Y0; Y < Work[w].Y1; Y++) {
                        vec4 const* Src = T->Pixels.data() + (size_t)Y * T->Width;
                        float* Dst = Layer.data() + ((size_t)(ry + Y) * AW + rx) * 4;
                        for (uint32_t X = 0; X < T->Width; X++, Dst += 4) {
                            vec4 V = Src[X];
                            if (T->Type == TEXTURE_TYPE_RAW) {
                                Dst[0] = V.x; Dst[1] = V.y; Dst[2] = V.z; Dst[3] = V.w;
                            } else if (T->Type == TEXTURE_TYPE_REFLECTANCE_WITH_ALPHA) {       // scene.cpp:1187-1195
                                vec3 B = GetParametricSpectrumCoefficients(S->RGBSpectrumTable, {V.x, V.y, V.z});
                                Dst[0] = B.x; Dst[1] = B.y; Dst[2] = B.z; Dst[3] = V.w;
                            } else {                                                            // RADIANCE, scene.cpp:1196-1211
                                float I = 2 * std::max(std::max(V.x, V.y), V.z);
                                if (I > 1e-6f) {
                                    vec3 B = GetParametricSpectrumCoefficients(S->RGBSpectrumTable, {V.x / I, V.y / I, V.z / I});
                                    Dst[0] = B.x; Dst[1] = B.y; Dst[2] = B.z; Dst[3] = I;
                                } else { Dst[0] = Dst[1] = Dst[2] = Dst[3] = 0.0f; }
                            }
                        }
                    }
                }
            };
            const unsigned Threads = (unsigned)std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), std::max<size_t>(1, Work.size()));
            std::vector<std::thread> Pool;
            for (unsigned t = 1; t < Threads; t++) Pool.emplace_back(Convert);
            Convert();
            for (auto& Th : Pool) Th.join();
        }
        S->Images.push_back(std::move(Layer));
        Pending.swap(Left);
    }
}

// scene.cpp:1031-1093
bounds shape_bounds(scene const* S, ptb_packed_shape const& O) {
    vec3 lo, hi;
    switch (O.Type) {
    case SHAPE_TYPE_MESH_INSTANCE: {
        auto const& N = S->MeshNodePack[O.MeshRootNodeIndex];
        lo = {N.Minimum[0], N.Minimum[1], N.Minimum[2]}; hi = {N.Maximum[0], N.Maximum[1], N.Maximum[2]};
        break; }
    case SHAPE_TYPE_PLANE: lo = {-1e9f, -1e9f, -1e-9f}; hi = {+1e9f, +1e9f, +1e-9f}; break;    // Q11
    default: lo = {-1, -1, -1}; hi = {+1, +1, +1}; break;
    }
    mat4 To; std::memcpy(&To, O.Transform.To, 64);
    bounds B;
    for (int k = 0; k < 8; k++) {
        vec4 C{(k & 1) ? hi.x : lo.x, (k & 2) ? hi.y : lo.y, (k & 4) ? hi.z : lo.z, 1.0f};
        vec4 Wc = mul(To, C);
        B.Minimum = min3(B.Minimum, {Wc.x, Wc.y, Wc.z});
        B.Maximum = max3(B.Maximum, {Wc.x, Wc.y, Wc.z});
    }
    return B;
}

// Top-level BVH over shapes: greedy nearest-neighbour agglomeration with 16-bit
// child indices, root swapped into slot 0 (scene.cpp:1402-1492, quirk Q9 incl. the
// `Size.z*Size.z` area term at :1437).
void build_shape_nodes(scene* S) {
    auto& Nodes = S->ShapeNodePack;
    Nodes.resize(1);
    // Node 0 of a scene without shapes: the reference's resize(1) keeps whatever its previous pack left there (scene.cpp:1349), a
    // root whose children no longer exist; here it is an empty leaf (found by tools/fuzz_pack.py's edit rounds).
    std::memset(&Nodes[0], 0, sizeof(Nodes[0]));
    std::vector<uint16_t> Map;
    for (uint32_t s = 0; s < S->ShapePack.size(); s++) {
        bounds B = shape_bounds(S, S->ShapePack[s]);
        Map.push_back((uint16_t)Nodes.size());
        ptb_packed_shape_node N{};
        N.Minimum[0] = B.Minimum.x; N.Minimum[1] = B.Minimum.y; N.Minimum[2] = B.Minimum.z;
        N.Maximum[0] = B.Maximum.x; N.Maximum[1] = B.Maximum.y; N.Maximum[2] = B.Maximum.z;
        N.ChildNodeIndices = 0; N.ShapeIndex = s;
        Nodes.push_back(N);
    }
    if (S->ShapePack.empty()) return;

    auto best_match = [&](uint16_t A) -> uint16_t {
        auto const& NA = Nodes[Map[A]];
        float BestArea = INFINITY;
        uint16_t Best = 0xFFFF;
        for (uint16_t B = 0; B < Map.size(); B++) {
            if (A == B) continue;
            auto const& NB = Nodes[Map[B]];
            float sx = std::max(NA.Maximum[0], NB.Maximum[0]) - std::min(NA.Minimum[0], NB.Minimum[0]);
            float sy = std::max(NA.Maximum[1], NB.Maximum[1]) - std::min(NA.Minimum[1], NB.Minimum[1]);
            float sz = std::max(NA.Maximum[2], NB.Maximum[2]) - std::min(NA.Minimum[2], NB.Minimum[2]);
            float Area = (sx * sy + sy * sz) + sz * sz;
            if (Area <= BestArea) { BestArea = Area; Best = B; }
        }
        return Best;
    };

    uint16_t A = 0, B = best_match(A);
    while (Map.size() > 1) {
        uint16_t C = best_match(B);
        if (A == C) {
            uint16_t IA = Map[A], IB = Map[B];
            ptb_packed_shape_node N{};
            for (int k = 0; k < 3; k++) {
                N.Minimum[k] = std::min(Nodes[IA].Minimum[k], Nodes[IB].Minimum[k]);
                N.Maximum[k] = std::max(Nodes[IA].Maximum[k], Nodes[IB].Maximum[k]);
            }
            N.ChildNodeIndices = uint32_t(IA) | (uint32_t(IB) << 16);
            N.ShapeIndex = SHAPE_INDEX_NONE;
            Map[A] = (uint16_t)Nodes.size();
            Map[B] = Map.back();
            Map.pop_back();
            if (A == Map.size()) A = B;
            Nodes.push_back(N);
            B = best_match(A);
        } else {
            A = B; B = C;
        }
    }
    Nodes[0] = Nodes[Map[A]];
    Nodes[Map[A]] = Nodes.back();
    Nodes.pop_back();
}

}  // namespace

uint32_t PackSceneData(scene* Scene) {
    uint32_t Dirty = Scene->DirtyFlags;

    if (Dirty & SCENE_DIRTY_TEXTURES) {
        pack_textures(Scene);
        Dirty |= SCENE_DIRTY_MATERIALS;
    }

    // Materials (scene.cpp:1236-1263): slots 0-1 = fallback OpenPBR; slot = word offset / 32.
    if (Dirty & SCENE_DIRTY_MATERIALS) {
        auto& P = Scene->MaterialAttributePack;
        P.clear();
        P.resize(64, 0u);
        P[0] = MATERIAL_TYPE_OPENPBR;
        for (material* M : Scene->Materials) {
            size_t Off = P.size();
            P.resize(Off + material_words(M->Type), 0u);
            P[Off] = (uint32_t)M->Type;
            pack_material(Scene, M, &P[Off]);
            M->PackedMaterialIndex = (uint32_t)Off / 32;
        }
        Dirty |= SCENE_DIRTY_SHAPES;
    }

    // Meshes (scene.cpp:1266-1343): global vertex / face / node index bases.
    if (Dirty & SCENE_DIRTY_MESHES) {
        auto& VP = Scene->MeshVertexPack; auto& FP = Scene->MeshFacePack; auto& NP = Scene->MeshNodePack;
        // Index bases first (a prefix sum over the meshes, in order), then every record is a pure function of its
        // source and its mesh's bases: the arrays are filled in pieces on all host threads and come out as the
        // reference's sequential push_backs do.
        struct base { uint32_t V, F, N; };
        std::vector<base> Bases(Scene->Meshes.size());
        size_t nv = 0, nf = 0, nn = 0;
        for (size_t m = 0; m < Scene->Meshes.size(); m++) {
            mesh* M = Scene->Meshes[m];
            Bases[m] = {(uint32_t)nv, (uint32_t)nf, (uint32_t)nn};
            M->PackedRootNodeIndex = (uint32_t)nn;
            nv += M->Vertices.size(); nf += M->Faces.size(); nn += M->Nodes.size();
        }
        VP.resize(nv); FP.resize(nf); NP.resize(nn);
        struct piece { uint32_t Mesh, Kind, Begin, End; };
        std::vector<piece> Pieces;
        constexpr uint32_t PIECE = 32768;
        for (size_t m = 0; m < Scene->Meshes.size(); m++) {
            mesh* M = Scene->Meshes[m];
            const size_t Counts[3] = {M->Vertices.size(), M->Faces.size(), M->Nodes.size()};
            for (uint32_t k = 0; k < 3; k++)
                for (size_t b = 0; b < Counts[k]; b += PIECE)
                    Pieces.push_back({(uint32_t)m, k, (uint32_t)b, (uint32_t)std::min(Counts[k], b + PIECE)});
        }
        std::atomic<size_t> Next(0);
        auto Work = [&]() {
            for (size_t i = Next.fetch_add(1); i < Pieces.size(); i = Next.fetch_add(1)) {
                piece const& W = Pieces[i];
                mesh const* M = Scene->Meshes[W.Mesh];
                const uint32_t VB = Bases[W.Mesh].V, FB = Bases[W.Mesh].F, NB = Bases[W.Mesh].N;
                if (W.Kind == 0) {
                    for (uint32_t v = W.Begin; v < W.End; v++) {
                        mesh_vertex const& V = M->Vertices[v];
                        VP[VB + v] = {pack_unit_vector(V.Normal), pack_half2x16(V.UV.x, V.UV.y)};
                    }
                } else if (W.Kind == 1) {
                    for (uint32_t f = W.Begin; f < W.End; f++) {
                        mesh_face const& F = M->Faces[f];
                        ptb_packed_mesh_face K;
                        vec3 const& p0 = M->Vertices[F.VertexIndex[0]].Position;
                        vec3 const& p1 = M->Vertices[F.VertexIndex[1]].Position;
                        vec3 const& p2 = M->Vertices[F.VertexIndex[2]].Position;
                        K.Position0[0] = p0.x; K.Position0[1] = p0.y; K.Position0[2] = p0.z; K.VertexIndex0 = VB + F.VertexIndex[0];
                        K.Position1[0] = p1.x; K.Position1[1] = p1.y; K.Position1[2] = p1.z; K.VertexIndex1 = VB + F.VertexIndex[1];
                        K.Position2[0] = p2.x; K.Position2[1] = p2.y; K.Position2[2] = p2.z; K.VertexIndex2 = VB + F.VertexIndex[2];
                        FP[FB + f] = K;
                    }
                } else {
                    for (uint32_t n = W.Begin; n < W.End; n++) {
                        mesh_node const& N = M->Nodes[n];
                        ptb_packed_mesh_node K;
                        K.Minimum[0] = N.Bounds.Minimum.x; K.Minimum[1] = N.Bounds.Minimum.y; K.Minimum[2] = N.Bounds.Minimum.z;
                        K.Maximum[0] = N.Bounds.Maximum.x; K.Maximum[1] = N.Bounds.Maximum.y; K.Maximum[2] = N.Bounds.Maximum.z;
                        if (N.ChildNodeIndex > 0) { K.FaceBeginOrNodeIndex = NB + N.ChildNodeIndex; K.FaceEndIndex = 0; }
                        else { K.FaceBeginOrNodeIndex = FB + N.FaceBeginIndex; K.FaceEndIndex = FB + N.FaceEndIndex; }
                        NP[NB + n] = K;
                    }
                }
            }
        };
        const unsigned T = (unsigned)std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), std::max<size_t>(1, Pieces.size() / 4));
        std::vector<std::thread> Pool;
        for (unsigned t = 1; t < T; t++) Pool.emplace_back(Work);
        Work();
        for (auto& Th : Pool) Th.join();
        Dirty |= SCENE_DIRTY_SHAPES;
    }

    // Shapes + top-level BVH (scene.cpp:1346-1498).
    if (Dirty & SCENE_DIRTY_SHAPES) {
        Scene->ShapePack.clear();
        walk(&Scene->Root, mat4::identity(), [Scene](entity* E, mat4 const& M) {
            ptb_packed_shape K{};
            switch (E->Type) {
            case ENTITY_TYPE_MESH_INSTANCE:
                if (!E->Mesh) return;
                K.Type = SHAPE_TYPE_MESH_INSTANCE; K.MeshRootNodeIndex = E->Mesh->PackedRootNodeIndex; break;
            case ENTITY_TYPE_PLANE: K.Type = SHAPE_TYPE_PLANE; break;
            case ENTITY_TYPE_SPHERE: K.Type = SHAPE_TYPE_SPHERE; break;
            case ENTITY_TYPE_CUBE: K.Type = SHAPE_TYPE_CUBE; break;
            default: return;
            }
            K.MaterialIndex = mat_index(E->Material);
            K.Transform = pack_transform(M);
            E->PackedShapeIndex = (uint32_t)Scene->ShapePack.size();
            Scene->ShapePack.push_back(K);
        });
        build_shape_nodes(Scene);
        Dirty |= SCENE_DIRTY_GLOBALS;
    }

    // Cameras (scene.cpp:1501-1539).  Pinhole sensor aspect is fixed at 2.0 (Q12).
    if (Dirty & SCENE_DIRTY_CAMERAS) {
        Scene->CameraPack.clear();
        walk(&Scene->Root, mat4::identity(), [Scene](entity* E, mat4 const& M) {
            if (E->Type != ENTITY_TYPE_CAMERA) return;
            ptb_packed_camera K{};
            K.Model = (uint32_t)E->CameraModel;
            if (E->CameraModel == CAMERA_MODEL_PINHOLE) {
                const float AspectRatio = 2.0f;
                K.ApertureRadius = E->Pinhole.ApertureDiameterInMM / 2000.0f;
                K.SensorSize[0] = 2 * std::tan((E->Pinhole.FieldOfViewInDegrees / 2) * 0.01745329251994329576923690768489f);
                K.SensorSize[1] = K.SensorSize[0] / AspectRatio;
                K.SensorDistance = 1.0f;
            }
            if (E->CameraModel == CAMERA_MODEL_THIN_LENS) {
                K.FocalLength = E->ThinLens.FocalLengthInMM / 1000.0f;
                K.ApertureRadius = E->ThinLens.ApertureDiameterInMM / 2000.0f;
                K.SensorDistance = 1.0f / (1000.0f / E->ThinLens.FocalLengthInMM - 1.0f / E->ThinLens.FocusDistance);
                K.SensorSize[0] = E->ThinLens.SensorSizeInMM.x / 1000.0f;
                K.SensorSize[1] = E->ThinLens.SensorSizeInMM.y / 1000.0f;
            }
            K.Transform = pack_transform(M);
            E->PackedCameraIndex = (uint32_t)Scene->CameraPack.size();
            Scene->CameraPack.push_back(K);
        });
    }

    // Skybox vMF lobe (scene.cpp:1542-1604, quirk Q23): luminance^2 * cos(theta) weights on raw RGB.
    if (Dirty & SCENE_DIRTY_SKYBOX_TEXTURE) {
        ptb_packed_scene_globals* G = &Scene->Globals;
        texture* Sky = Scene->Root.SkyboxTexture;
        G->SkyboxTextureIndex = tex_index(Sky);
        if (Sky) {
            const float PI_ = 3.141592653f, TAU_ = 6.283185306f;
            vec3 Mean{0, 0, 0};
            float WeightSum = 0.0f;
            for (uint32_t Y = 0; Y < Sky->Height; Y++) {
                float Theta = (0.5f - (Y + 0.5f) / Sky->Height) * PI_;
                for (uint32_t X = 0; X < Sky->Width; X++) {
                    float Phi = ((X + 0.5f) / Sky->Width - 0.5f) * TAU_;
                    vec4 C = Sky->Pixels[(size_t)Y * Sky->Width + X];
                    float Lum = dot3({0.2126f, 0.7152f, 0.0722f}, {C.x, C.y, C.z});
                    float Weight = std::cos(Theta) * Lum * Lum;
                    vec3 Dir{std::cos(Theta) * std::cos(Phi), std::cos(Theta) * std::sin(Phi), std::sin(Theta)};
                    Mean = Mean + Weight * Dir;
                    WeightSum += Weight;
                }
            }
            Mean = Mean / WeightSum;
            float R = length3(Mean);
            vec3 D = Mean / R;
            G->SkyboxMeanDirection[0] = D.x; G->SkyboxMeanDirection[1] = D.y; G->SkyboxMeanDirection[2] = D.z;
            G->SkyboxConcentration = R * (3.0f - R * R) / (1 - R * R);
        }
        Dirty |= SCENE_DIRTY_GLOBALS;
    }

    if (Dirty & SCENE_DIRTY_GLOBALS) {           // scene.cpp:1607-1616
        ptb_packed_scene_globals* G = &Scene->Globals;
        G->SkyboxSamplingProbability = Scene->Root.SkyboxSamplingProbability;
        G->SkyboxBrightness = Scene->Root.SkyboxBrightness;
        G->SceneScatterRate = Scene->Root.ScatterRate;
        G->ShapeCount = (uint32_t)Scene->ShapePack.size();
    }

    Scene->DirtyFlags = 0;
    return Dirty;
}

ptb_scene_view scene::View() {
    ImagePointers.clear();
    for (auto const& L : Images) ImagePointers.push_back(L.data());
    ptb_scene_view V{};
    V.Textures = TexturePack.data();           V.TextureCount = (uint32_t)TexturePack.size();
    V.MaterialData = MaterialAttributePack.data(); V.MaterialWordCount = (uint32_t)MaterialAttributePack.size();
    V.Shapes = ShapePack.data();               V.ShapeCount = (uint32_t)ShapePack.size();
    V.ShapeNodes = ShapeNodePack.data();       V.ShapeNodeCount = (uint32_t)ShapeNodePack.size();
    V.MeshFaces = MeshFacePack.data();         V.MeshFaceCount = (uint32_t)MeshFacePack.size();
    V.MeshVertices = MeshVertexPack.data();    V.MeshVertexCount = (uint32_t)MeshVertexPack.size();
    V.MeshNodes = MeshNodePack.data();         V.MeshNodeCount = (uint32_t)MeshNodePack.size();
    V.Cameras = CameraPack.data();             V.CameraCount = (uint32_t)CameraPack.size();
    V.AtlasLayers = ImagePointers.data();      V.AtlasLayerCount = (uint32_t)ImagePointers.size();
    V.AtlasWidth = AtlasWidth;                 V.AtlasHeight = AtlasHeight;
    V.Globals = &Globals;
    return V;
}

}}  // namespace ptb::host
