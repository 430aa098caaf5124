// Host scene layer, part 1: source-description API and the mesh BVH2 builder.
// See scene.hpp.  Citations are to /root/reference/src/scene/scene.cpp.
#include "scene.hpp"
#include <cassert>
#include <cstdio>
#include <atomic>
#include <mutex>
#include <thread>

namespace ptb { namespace host {

// ---- process-wide RGB->spectrum table (scene.cpp:918-927) ------------------
namespace {
std::mutex g_table_mutex;
parametric_spectrum_table* g_table = nullptr;
std::string g_table_path = "sRGBSpectrumTable.dat";
}

void SetSpectrumTablePath(char const* Path) { std::lock_guard<std::mutex> l(g_table_mutex); g_table_path = Path; }

parametric_spectrum_table const* SharedSpectrumTable(char const* Path) {
    std::lock_guard<std::mutex> l(g_table_mutex);
    if (g_table) return g_table;
    if (Path) g_table_path = Path;
    g_table = new parametric_spectrum_table;
    if (!LoadParametricSpectrumTable(g_table, g_table_path.c_str())) {
        std::fprintf(stderr, "[ptb-host] %s not found, generating it (all host threads)...\n", g_table_path.c_str());
        BuildParametricSpectrumTableForSRGB(g_table, 0);
        if (!SaveParametricSpectrumTable(g_table, g_table_path.c_str()))
            std::fprintf(stderr, "[ptb-host] warning: could not save %s\n", g_table_path.c_str());
    }
    return g_table;
}

// ---- entities (scene.cpp:161-268) -------------------------------------------
entity* CreateEntity(scene* Scene, entity_type Type, entity* Parent) {
    entity* E = new entity;
    E->Type = Type;
    if (!Parent) Parent = &Scene->Root;
    E->Parent = Parent;
    Parent->Children.push_back(E);
    if (Type == ENTITY_TYPE_CAMERA) Scene->DirtyFlags |= SCENE_DIRTY_CAMERAS;
    else Scene->DirtyFlags |= SCENE_DIRTY_SHAPES;
    return E;
}

void DestroyEntity(scene* Scene, entity* Entity) {
    if (entity* P = Entity->Parent) {
        auto& C = P->Children;
        C.erase(std::remove(C.begin(), C.end(), Entity), C.end());
    }
    std::vector<entity*> Kids = Entity->Children;
    for (entity* K : Kids) { K->Parent = nullptr; DestroyEntity(Scene, K); }
    Scene->DirtyFlags |= SCENE_DIRTY_SHAPES | SCENE_DIRTY_CAMERAS;
    delete Entity;
}

// scene.cpp:380-410
material* CreateMaterial(scene* Scene, material_type Type, char const* Name) {
    material* M = new material;
    M->Type = Type;
    M->Name = Name ? Name : "New Material";
    Scene->Materials.push_back(M);
    Scene->DirtyFlags |= SCENE_DIRTY_MATERIALS;
    return M;
}

// scene.cpp:270-292: 2x2 checker A B / B A.
texture* CreateCheckerTexture(scene* Scene, char const* Name, texture_type Type, vec4 const& A, vec4 const& B) {
    vec4 Px[4] = {A, B, B, A};
    return CreateTexture(Scene, Name, Type, 2, 2, Px);
}

texture* CreateTexture(scene* Scene, char const* Name, texture_type Type, uint32_t Width, uint32_t Height, vec4 const* Pixels) {
    texture* T = new texture;
    T->Name = Name ? Name : "New Texture";
    T->Type = Type;
    T->Width = Width; T->Height = Height;
    T->Pixels.assign(Pixels, Pixels + (size_t)Width * Height);
    Scene->Textures.push_back(T);
    Scene->DirtyFlags |= SCENE_DIRTY_TEXTURES;
    return T;
}

// ---- mesh BVH2 (scene.cpp:424-599) -------------------------------------------
namespace {

inline void grow(bounds& B, vec3 P) { B.Minimum = min3(B.Minimum, P); B.Maximum = max3(B.Maximum, P); }
inline void grow(bounds& B, bounds const& O) { B.Minimum = min3(B.Minimum, O.Minimum); B.Maximum = max3(B.Maximum, O.Maximum); }
// scene.cpp:53-57
inline float half_area(bounds const& B) {
    vec3 E = B.Maximum - B.Minimum;
    return (E.x * E.y + E.y * E.z) + E.z * E.x;
}

// Per-face cache kept in lock-step with Mesh->Faces while partitioning; holds
// exactly the values the reference recomputes on the fly (scene.cpp:424-433,443-450).
struct face_cache { vec3 Centroid; bounds Box; };

// Threads the mesh builders may still start, process-wide (the caller's own thread is not counted).
std::atomic<int>& free_host_threads() {
    static std::atomic<int> Free{(int)std::max(1u, std::thread::hardware_concurrency()) - 1};
    return Free;
}

// Runs Body(piece, first, last) over [Begin, End) cut into T contiguous pieces, one thread each (T = 1: inline).
template <typename F> void in_pieces(uint32_t Begin, uint32_t End, unsigned T, F const& Body) {
    if (T <= 1) { Body(0u, Begin, End); return; }
    std::vector<std::thread> Pool;
    const uint64_t N = End - Begin;
    for (unsigned t = 0; t < T; t++)
        Pool.emplace_back([&, t]() { Body(t, Begin + (uint32_t)(N * t / T), Begin + (uint32_t)(N * (t + 1) / T)); });
    for (auto& Th : Pool) Th.join();
}

// The reference's builder is one sequential recursion.  This one produces the SAME nodes in the SAME order with
// the SAME face permutation on several threads:
//  * a node pair is allocated when its parent splits, then the whole left subtree, then the whole right one
//    (scene.cpp:575-598), so the descendants of a node are one contiguous run whose length is the only thing the
//    right sibling needs from the left: the two subtrees are built apart (the left one where it will stay, the
//    right one with indices relative to itself) and the right one is appended with its child indices shifted;
//  * inside a large node the passes over its faces (bounds, centroid range, binning) are cut into pieces whose
//    partial results merge exactly - min / max of floats and integer counts do not depend on the order;
//  * the in-place partition (scene.cpp:556-573) decides the face order and stays sequential.
struct builder {
    static constexpr uint32_t BINS = 32;
    static constexpr uint32_t SUBTREE_TASK_FACES = 2048;        // smaller subtrees stay on the thread that reached them
    static constexpr uint32_t PIECEWISE_FACES = 131072;         // larger nodes cut their passes into pieces
    static constexpr unsigned MAX_PIECES = 32;
    mesh* Mesh;
    std::vector<face_cache> Cache;
    std::atomic<int>& ThreadsFree = free_host_threads();        // one budget for all builds running at once
    std::atomic<uint32_t> MaxDepth{0};
    unsigned Hardware = 1;

    explicit builder(mesh* M) : Mesh(M), Cache(M->Faces.size()) {
        Hardware = std::max(1u, std::thread::hardware_concurrency());
        const uint32_t N = (uint32_t)M->Faces.size();
        in_pieces(0, N, N >= PIECEWISE_FACES ? std::min(Hardware, MAX_PIECES) : 1, [&](unsigned, uint32_t b, uint32_t e) {
            for (uint32_t f = b; f < e; f++) {
                face_cache& C = Cache[f];
                for (int a = 0; a < 3; a++) {
                    float Sum = 0.0f;
                    for (int k = 0; k < 3; k++) Sum += M->Vertices[M->Faces[f].VertexIndex[k]].Position[a];
                    C.Centroid[a] = Sum / 3.0f;
                }
                for (int k = 0; k < 3; k++) grow(C.Box, M->Vertices[M->Faces[f].VertexIndex[k]].Position);
            }
        });
    }

    int take_threads(int Wanted) {
        int Free = ThreadsFree.load();
        while (Free > 0) {
            int Take = std::min(Free, Wanted);
            if (ThreadsFree.compare_exchange_weak(Free, Free - Take)) return Take;
        }
        return 0;
    }

    struct axis_bins { float Lo, Hi; bounds Box[BINS]; uint32_t Count[BINS]; };

    // BuildMeshNode (scene.cpp:435-599): top-down binned SAH, 32 bins per axis, cost N_L*A_L + N_R*A_R against
    // N*A, children allocated adjacently.  Out[Self] is the node; a node at Out[k] has index Bias + k.
    void build(std::vector<mesh_node>& Out, uint32_t Self, uint32_t Bias, uint32_t Depth) {
        const uint32_t Begin = Out[Self].FaceBeginIndex;
        const uint32_t End = Out[Self].FaceEndIndex;
        const uint32_t FaceCount = End - Begin;
        const int Extra = FaceCount >= PIECEWISE_FACES ? take_threads((int)std::min<uint32_t>(std::min(Hardware, MAX_PIECES) - 1, FaceCount / (PIECEWISE_FACES / 4))) : 0;
        const unsigned Pieces = 1 + Extra;                          // <= MAX_PIECES

        bounds NodeBounds, CentroidBounds;
        if (Pieces == 1) {
            for (uint32_t f = Begin; f < End; f++) { grow(NodeBounds, Cache[f].Box); grow(CentroidBounds, Cache[f].Centroid); }
        } else {
            std::vector<bounds> Part(2 * (size_t)Pieces);
            in_pieces(Begin, End, Pieces, [&](unsigned t, uint32_t b, uint32_t e) {
                bounds B, C;
                for (uint32_t f = b; f < e; f++) { grow(B, Cache[f].Box); grow(C, Cache[f].Centroid); }
                Part[2 * t] = B; Part[2 * t + 1] = C;
            });
            for (unsigned t = 0; t < Pieces; t++) { grow(NodeBounds, Part[2 * t]); grow(CentroidBounds, Part[2 * t + 1]); }
        }
        float Lo[3], Hi[3];
        for (int a = 0; a < 3; a++) { Lo[a] = CentroidBounds.Minimum[a]; Hi[a] = CentroidBounds.Maximum[a]; }
        Out[Self].Bounds = NodeBounds;

        // Binning of the three axes in one pass over the faces (the reference makes one pass per axis; the bins
        // of an axis depend on that axis alone).
        axis_bins Own[3];
        std::vector<axis_bins> Shared;
        axis_bins* Bins = Own;
        if (Pieces > 1) { Shared.resize((size_t)Pieces * 3); Bins = Shared.data(); }
        in_pieces(Begin, End, Pieces, [&](unsigned t, uint32_t b, uint32_t e) {
            axis_bins* A = &Bins[(size_t)t * 3];
            float BinsPerUnit[3];
            for (int a = 0; a < 3; a++) {
                for (uint32_t i = 0; i < BINS; i++) { A[a].Box[i] = bounds(); A[a].Count[i] = 0; }
                BinsPerUnit[a] = float(BINS) / (Hi[a] - Lo[a]);
            }
            for (uint32_t f = b; f < e; f++) {
                face_cache const& C = Cache[f];
                for (int a = 0; a < 3; a++) {
                    if (Lo[a] == Hi[a]) continue;
                    uint32_t i = std::min(static_cast<uint32_t>(BinsPerUnit[a] * (C.Centroid[a] - Lo[a])), BINS - 1);
                    grow(A[a].Box[i], C.Box);
                    A[a].Count[i]++;
                }
            }
        });
        if (Extra) ThreadsFree.fetch_add(Extra);

        for (unsigned t = 1; t < Pieces; t++)                        // exact: min / max and integer sums
            for (int a = 0; a < 3; a++)
                for (uint32_t i = 0; i < BINS; i++) {
                    axis_bins const& A = Bins[(size_t)t * 3 + a];
                    if (A.Count[i]) { grow(Bins[a].Box[i], A.Box[i]); Bins[a].Count[i] += A.Count[i]; }
                }

        int SplitAxis = 0;
        float SplitPosition = 0.0f;
        float SplitCost = +INFINITY;
        for (int Axis = 0; Axis < 3; Axis++) {
            if (Lo[Axis] == Hi[Axis]) continue;
            bounds const* BinBox = Bins[Axis].Box;
            uint32_t const* BinCount = Bins[Axis].Count;

            // Prefix (left) and suffix (right) sweeps: scene.cpp:506-532.
            float LeftArea[BINS - 1], RightArea[BINS - 1];
            uint32_t LeftCount[BINS - 1], RightCount[BINS - 1];
            bounds LB, RB;
            uint32_t LN = 0, RN = 0;
            for (uint32_t i = 0; i < BINS - 1; i++) {
                uint32_t j = BINS - 2 - i;
                if (BinCount[i] > 0) { LN += BinCount[i]; grow(LB, BinBox[i]); }
                LeftCount[i] = LN; LeftArea[i] = half_area(LB);
                if (BinCount[j + 1] > 0) { RN += BinCount[j + 1]; grow(RB, BinBox[j + 1]); }
                RightCount[j] = RN; RightArea[j] = half_area(RB);
            }

            const float Interval = (Hi[Axis] - Lo[Axis]) / float(BINS);
            float Position = Lo[Axis] + Interval;
            for (uint32_t i = 0; i < BINS - 1; i++) {
                float Cost = LeftCount[i] * LeftArea[i] + RightCount[i] * RightArea[i];
                if (Cost < SplitCost) { SplitCost = Cost; SplitAxis = Axis; SplitPosition = Position; }
                Position += Interval;
            }
        }

        // Leaf if splitting does not pay (scene.cpp:552-554); no leaf-size clamp (Q20).
        if (SplitCost >= FaceCount * half_area(NodeBounds)) return;

        // In-place partition, scene.cpp:556-573 (the last candidate is never tested).
        uint32_t Split = Begin, Swap = End - 1;
        while (Split < Swap) {
            if (Cache[Split].Centroid[SplitAxis] < SplitPosition) {
                Split++;
            } else {
                std::swap(Mesh->Faces[Split], Mesh->Faces[Swap]);
                std::swap(Cache[Split], Cache[Swap]);
                Swap--;
            }
        }
        if (Split == Begin || Split == End) return;

        const uint32_t Left = static_cast<uint32_t>(Out.size());
        Out[Self].ChildNodeIndex = Bias + Left;
        mesh_node L; L.FaceBeginIndex = Begin; L.FaceEndIndex = Split;
        mesh_node R; R.FaceBeginIndex = Split; R.FaceEndIndex = End;
        Out.push_back(L);
        Out.push_back(R);
        uint32_t Seen = MaxDepth.load();
        while (Seen < Depth + 1 && !MaxDepth.compare_exchange_weak(Seen, Depth + 1)) {}

        if (std::min(Split - Begin, End - Split) >= SUBTREE_TASK_FACES && take_threads(1) == 1) {
            // Right subtree on another thread, numbered from itself (Side[0] = R, Side[k] = its k-th descendant);
            // the left one goes where it belongs, and tells how far the right one has to move.
            std::vector<mesh_node> Side;
            Side.reserve(2 * (size_t)(End - Split));
            Side.push_back(R);
            std::thread Th([&]() { build(Side, 0, 0, Depth + 1); ThreadsFree.fetch_add(1); });
            build(Out, Left, Bias, Depth + 1);
            Th.join();
            const uint32_t Shift = Bias + static_cast<uint32_t>(Out.size()) - 1;      // Side[1] lands at Out.size()
            for (mesh_node& N : Side) if (N.ChildNodeIndex) N.ChildNodeIndex += Shift;
            Out[Left + 1] = Side[0];
            Out.insert(Out.end(), Side.begin() + 1, Side.end());
        } else {
            build(Out, Left, Bias, Depth + 1);
            build(Out, Left + 1, Bias, Depth + 1);
        }
    }
};

}  // namespace

// The mesh and its BVH without touching any scene: safe to run for several meshes at once (each build is the
// reference's sequential algorithm, so node and face order per mesh are unchanged).
mesh* BuildMesh(char const* Name, std::vector<mesh_vertex>&& Vertices, std::vector<mesh_face>&& Faces) {
    mesh* M = new mesh;
    M->Name = Name ? Name : "Mesh";
    M->Vertices = std::move(Vertices);
    M->Faces = std::move(Faces);
    M->Nodes.reserve(2 * M->Faces.size());          // scene.cpp:853
    mesh_node Root; Root.FaceBeginIndex = 0; Root.FaceEndIndex = static_cast<uint32_t>(M->Faces.size());
    M->Nodes.push_back(Root);
    builder B(M);
    B.build(M->Nodes, 0, 0, 0);
    M->Depth = B.MaxDepth.load();
    return M;
}
mesh* AttachMesh(scene* Scene, mesh* M) {
    Scene->Meshes.push_back(M);
    Scene->DirtyFlags |= SCENE_DIRTY_MATERIALS | SCENE_DIRTY_MESHES;   // scene.cpp:868-869
    return M;
}
mesh* CreateMesh(scene* Scene, char const* Name, std::vector<mesh_vertex>&& Vertices, std::vector<mesh_face>&& Faces) {
    return AttachMesh(Scene, BuildMesh(Name, std::move(Vertices), std::move(Faces)));
}

// scene.cpp:912-943
scene* CreateScene(bool WithDefaults) {
    scene* S = new scene;
    S->Root.Name = "Scene";
    S->Root.Type = ENTITY_TYPE_ROOT;
    S->RGBSpectrumTable = SharedSpectrumTable();
    std::memset(&S->Globals, 0, sizeof(S->Globals));
    S->Globals.SkyboxConcentration = 1.0f;           // scene.hpp:152-161 defaults
    S->Globals.SkyboxBrightness = 1.0f;
    S->Globals.SkyboxTextureIndex = TEXTURE_INDEX_NONE;
    if (WithDefaults) {
        material* PlaneMaterial = CreateMaterial(S, MATERIAL_TYPE_BASIC_DIFFUSE, "Plane Material");
        PlaneMaterial->BaseTexture = CreateCheckerTexture(S, "Plane Texture", TEXTURE_TYPE_REFLECTANCE_WITH_ALPHA, {1, 1, 1, 1}, {0.5f, 0.5f, 0.5f, 1});
        PlaneMaterial->BaseTexture->EnableNearestFiltering = true;
        entity* Plane = CreateEntity(S, ENTITY_TYPE_PLANE);
        Plane->Name = "Plane";
        Plane->Material = PlaneMaterial;
        entity* Camera = CreateEntity(S, ENTITY_TYPE_CAMERA);
        Camera->Name = "Camera";
        Camera->Transform.Position = {0, 0, 1};
    }
    S->DirtyFlags = SCENE_DIRTY_ALL;
    return S;
}

// scene.cpp:945-959
void DestroyScene(scene* Scene) {
    while (!Scene->Root.Children.empty()) DestroyEntity(Scene, Scene->Root.Children[0]);
    while (!Scene->Prefabs.empty()) DestroyPrefab(Scene, Scene->Prefabs.back());
    for (mesh* M : Scene->Meshes) delete M;
    for (material* M : Scene->Materials) delete M;
    for (texture* T : Scene->Textures) delete T;
    delete Scene;
}

static entity* find_shape(entity* E, uint32_t Index) {   // scene.cpp:1623-1640
    if (!E->Active) return nullptr;
    if (E->PackedShapeIndex == Index) return E;
    for (entity* C : E->Children) if (entity* R = find_shape(C, Index)) return R;
    return nullptr;
}
entity* FindEntityByPackedShapeIndex(scene* Scene, uint32_t PackedShapeIndex) { return find_shape(&Scene->Root, PackedShapeIndex); }

}}  // namespace ptb::host
