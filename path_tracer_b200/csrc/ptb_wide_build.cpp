// See ptb_wide_build.h.  Host-only C++ (no CUDA): binned-SAH BVH2 over world-space
// primitive bounds -> SAH-optimal collapse to 8-wide by dynamic programming (the
// forest-cost recurrence of Ylitie, Karras, Laine, "Efficient Incoherent Ray
// Traversal on GPUs Through Compressed Wide BVHs", HPG 2017, section 3.1) -> octant
// slot assignment -> 8-bit quantisation (outward rounded).
#include "ptb_wide_build.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace ptb {
namespace {

struct Box {
    float lo[3], hi[3];
    void reset() { for (int a = 0; a < 3; a++) { lo[a] = INFINITY; hi[a] = -INFINITY; } }
    void grow(const Box& o) { for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], o.lo[a]); hi[a] = std::max(hi[a], o.hi[a]); } }
    void grow(const float* p) { for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], p[a]); hi[a] = std::max(hi[a], p[a]); } }
    double area() const {
        double x = (double)hi[0] - lo[0], y = (double)hi[1] - lo[1], z = (double)hi[2] - lo[2];
        if (x < 0 || y < 0 || z < 0) return 0.0;
        return 2.0 * (x * y + y * z + z * x);
    }
};

struct Prim { uint32_t shape, face, kind, tag; };   // kind: 0 triangle, 2 sphere, 3 cube (shape types); tag: see ptb_wide.cuh

struct B2 { Box b; uint32_t left, right, first, count; };   // leaf when count > 0

constexpr int BINS = 16;
static uint32_t MAX_LEAF = 2;           // a wide-node child slot holds 1..3 primitives; 2 measured best (profiles/r01_v6_build_sweep.txt)

inline void xform_point(const float* m, const float* p, double* out) {   // column-major To * (p,1), in double
    for (int r = 0; r < 3; r++) out[r] = (double)m[r] * p[0] + (double)m[4 + r] * p[1] + (double)m[8 + r] * p[2] + (double)m[12 + r];
}

// Pads a world-space box outward so that every point the object-space primitive
// test can report (float rounding in From*ray and in the test itself) lies inside.
inline void pad_box(Box& b) {
    for (int a = 0; a < 3; a++) {
        float m = std::max(std::fabs(b.lo[a]), std::fabs(b.hi[a]));
        float e = 4e-6f * (m + (b.hi[a] - b.lo[a])) + 1e-30f;
        b.lo[a] -= e; b.hi[a] += e;
    }
}

struct Builder {
    std::vector<Box> pbox;
    std::vector<float> pcen;          // 3 per prim
    std::vector<Prim> prims;
    std::vector<uint32_t> idx;
    std::vector<B2> nodes;
    std::atomic<uint32_t> node_alloc{0};
    std::atomic<int> threads_free{0};
    std::atomic<uint32_t> max_depth{0};

    uint32_t alloc2() { return node_alloc.fetch_add(2); }

    void make_leaf(uint32_t n, uint32_t first, uint32_t count, const Box& b) {
        nodes[n].b = b; nodes[n].left = nodes[n].right = 0; nodes[n].first = first; nodes[n].count = count;
    }

    void build(uint32_t n, uint32_t first, uint32_t count, uint32_t depth) {
        Box b, cb;
        b.reset(); cb.reset();
        for (uint32_t i = first; i < first + count; i++) { b.grow(pbox[idx[i]]); cb.grow(&pcen[3 * (size_t)idx[i]]); }
        uint32_t d = max_depth.load();
        while (depth > d && !max_depth.compare_exchange_weak(d, depth)) {}
        if (count <= MAX_LEAF) { make_leaf(n, first, count, b); return; }

        // binned SAH over the centroid bounds, 3 axes
        int best_axis = -1, best_split = -1;
        double best_cost = INFINITY;
        float scale[3];
        for (int a = 0; a < 3; a++) { float e = cb.hi[a] - cb.lo[a]; scale[a] = e > 0 ? (float)BINS * (1.0f - 1e-6f) / e : 0.0f; }
        if (depth < 60) {
            Box bins[3][BINS]; uint32_t cnt[3][BINS];
            for (int a = 0; a < 3; a++) for (int k = 0; k < BINS; k++) { bins[a][k].reset(); cnt[a][k] = 0; }
            for (uint32_t i = first; i < first + count; i++) {
                uint32_t p = idx[i];
                for (int a = 0; a < 3; a++) {
                    int k = (int)((pcen[3 * (size_t)p + a] - cb.lo[a]) * scale[a]);
                    k = k < 0 ? 0 : (k >= BINS ? BINS - 1 : k);
                    bins[a][k].grow(pbox[p]); cnt[a][k]++;
                }
            }
            for (int a = 0; a < 3; a++) {
                if (scale[a] == 0.0f) continue;
                double la[BINS]; uint32_t ln[BINS];
                Box acc; acc.reset(); uint32_t c = 0;
                for (int k = 0; k < BINS - 1; k++) { acc.grow(bins[a][k]); c += cnt[a][k]; la[k] = acc.area(); ln[k] = c; }
                acc.reset(); c = 0;
                for (int k = BINS - 1; k > 0; k--) {
                    acc.grow(bins[a][k]); c += cnt[a][k];
                    if (ln[k - 1] == 0 || c == 0) continue;
                    double cost = la[k - 1] * ln[k - 1] + acc.area() * c;
                    if (cost < best_cost) { best_cost = cost; best_axis = a; best_split = k; }
                }
            }
        }
        uint32_t mid;
        if (best_axis >= 0) {
            int a = best_axis; float lo = cb.lo[a], sc = scale[a]; int sp = best_split;
            auto it = std::partition(idx.begin() + first, idx.begin() + first + count, [&](uint32_t p) {
                int k = (int)((pcen[3 * (size_t)p + a] - lo) * sc);
                k = k < 0 ? 0 : (k >= BINS ? BINS - 1 : k);
                return k < sp;
            });
            mid = (uint32_t)(it - idx.begin());
        } else {
            mid = first;
        }
        if (mid == first || mid == first + count) {
            // coincident centroids (or depth guard): object-median split along the widest axis
            int a = 0;
            for (int k = 1; k < 3; k++) if (b.hi[k] - b.lo[k] > b.hi[a] - b.lo[a]) a = k;
            mid = first + count / 2;
            std::nth_element(idx.begin() + first, idx.begin() + mid, idx.begin() + first + count,
                             [&](uint32_t x, uint32_t y) { return pcen[3 * (size_t)x + a] < pcen[3 * (size_t)y + a]; });
        }
        uint32_t c = alloc2();
        nodes[n].b = b; nodes[n].left = c; nodes[n].right = c + 1; nodes[n].first = 0; nodes[n].count = 0;
        uint32_t nl = mid - first, nr = count - nl;
        bool spawned = false;
        std::thread th;
        if (std::min(nl, nr) > 32768) {
            int f = threads_free.load();
            while (f > 0 && !threads_free.compare_exchange_weak(f, f - 1)) {}
            if (f > 0) { spawned = true; th = std::thread([=]() { build(c, first, nl, depth + 1); threads_free.fetch_add(1); }); }
        }
        if (!spawned) build(c, first, nl, depth + 1);
        build(c + 1, mid, nr, depth + 1);
        if (spawned) th.join();
    }
};

inline uint32_t exponent_byte(double extent) {
    // smallest e with 255 * 2^e >= extent, as a biased byte (2^e = float bits byte<<23)
    if (!(extent > 0.0)) return 1;    // 2^-126
    int e;
    double m = std::frexp(extent / 255.0, &e);   // extent/255 = m * 2^e, m in [0.5,1)
    if (m == 0.5) e -= 1;                        // exact power of two
    int biased = e + 127;
    if (biased < 1) biased = 1;
    if (biased > 254) biased = 254;
    return (uint32_t)biased;
}

}  // namespace

int build_wide_bvh(const ptb_scene_view* v, WideBuild* out, std::string* err) {
    auto t0 = std::chrono::steady_clock::now();
    *out = WideBuild();
    if (std::getenv("PTB_WIDE_MAXLEAF")) MAX_LEAF = (uint32_t)std::atoi(std::getenv("PTB_WIDE_MAXLEAF"));
    Builder B;

    // ---- gather bounded primitives in world space ----
    for (uint32_t s = 0; s < v->ShapeCount; s++) {
        const ptb_packed_shape& sh = v->Shapes[s];
        const float* to = sh.Transform.To;
        if (sh.Type == 1) { out->planes.push_back(s); continue; }
        if (sh.Type == 2 || sh.Type == 3) {
            Box b; b.reset();
            for (int c = 0; c < 8; c++) {
                float p[3] = {(c & 1) ? 1.f : -1.f, (c & 2) ? 1.f : -1.f, (c & 4) ? 1.f : -1.f};
                double w[3]; xform_point(to, p, w);
                float wf[3] = {(float)w[0], (float)w[1], (float)w[2]};
                b.grow(wf);
            }
            pad_box(b);
            B.pbox.push_back(b); B.prims.push_back({s, 0u, (uint32_t)sh.Type, 0u});
            continue;
        }
        if (sh.Type != 0) continue;
        if (sh.MeshRootNodeIndex >= v->MeshNodeCount) { *err = "shape mesh root out of range"; return PTB_E_INVALID; }
        std::vector<uint32_t> stack{sh.MeshRootNodeIndex};
        while (!stack.empty()) {
            uint32_t ni = stack.back(); stack.pop_back();
            if (ni >= v->MeshNodeCount) { *err = "mesh node index out of range"; return PTB_E_INVALID; }
            const ptb_packed_mesh_node& n = v->MeshNodes[ni];
            if (n.FaceEndIndex > 0) {
                if (n.FaceEndIndex > v->MeshFaceCount || n.FaceBeginOrNodeIndex > n.FaceEndIndex) { *err = "mesh face range out of range"; return PTB_E_INVALID; }
                for (uint32_t f = n.FaceBeginOrNodeIndex; f < n.FaceEndIndex; f++) {
                    const ptb_packed_mesh_face& F = v->MeshFaces[f];
                    Box b; b.reset();
                    const float* P[3] = {F.Position0, F.Position1, F.Position2};
                    for (int k = 0; k < 3; k++) {
                        double w[3]; xform_point(to, P[k], w);
                        float wf[3] = {(float)w[0], (float)w[1], (float)w[2]};
                        b.grow(wf);
                    }
                    if (!(std::isfinite(b.lo[0]) && std::isfinite(b.lo[1]) && std::isfinite(b.lo[2]) &&
                          std::isfinite(b.hi[0]) && std::isfinite(b.hi[1]) && std::isfinite(b.hi[2]))) continue;
                    pad_box(b);
                    B.pbox.push_back(b); B.prims.push_back({s, f, 0u, 0u});
                }
            } else {
                stack.push_back(n.FaceBeginOrNodeIndex + 1);
                stack.push_back(n.FaceBeginOrNodeIndex);
            }
        }
    }
    const size_t N = B.prims.size();
    out->prim_count = (uint32_t)N;
    if (N + out->planes.size() >= (1ull << 27)) { *err = "more than 2^27 primitives"; return PTB_E_UNSUPPORTED; }
    {
        // tag = rank << 3 | identity << 2 | kind.  rank = position in the canonical (shape, face) order, so a
        // 64-bit key {hit time bits, tag} orders hits exactly like the canonical tie rule; identity = the shape's
        // transform is the identity matrix (object ray = world ray + 0.0f, no matrix fetch); kind = shape type & 3.
        std::vector<std::pair<uint64_t, uint32_t>> order;
        order.reserve(N + out->planes.size());
        for (size_t i = 0; i < N; i++) order.push_back({((uint64_t)B.prims[i].shape << 32) | B.prims[i].face, (uint32_t)i});
        for (size_t i = 0; i < out->planes.size(); i++) order.push_back({(uint64_t)out->planes[i] << 32, (uint32_t)(N + i)});
        std::sort(order.begin(), order.end());
        auto is_identity = [&](uint32_t shape) {
            static const float I[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
            return !std::memcmp(v->Shapes[shape].Transform.From, I, 64) && !std::memcmp(v->Shapes[shape].Transform.To, I, 64);
        };
        std::vector<uint32_t> plane_tags(out->planes.size());
        for (size_t r = 0; r < order.size(); r++) {
            const uint32_t i = order[r].second;
            if (i < N) B.prims[i].tag = ((uint32_t)r << 3) | (is_identity(B.prims[i].shape) ? 4u : 0u) | (B.prims[i].kind & 3u);
            else plane_tags[i - N] = ((uint32_t)r << 3) | 1u;
        }
        std::vector<uint32_t> pl;
        for (size_t i = 0; i < out->planes.size(); i++) { pl.push_back(out->planes[i]); pl.push_back(plane_tags[i]); }
        out->planes.swap(pl);      // {shape, tag} pairs
    }
    if (N == 0) { out->node_count = 0; return PTB_OK; }
    if (N > 0x7FFFFFFFull) { *err = "too many primitives for the wide BVH"; return PTB_E_UNSUPPORTED; }

    B.pcen.resize(3 * N);
    B.idx.resize(N);
    for (size_t i = 0; i < N; i++) {
        B.idx[i] = (uint32_t)i;
        for (int a = 0; a < 3; a++) B.pcen[3 * i + a] = 0.5f * B.pbox[i].lo[a] + 0.5f * B.pbox[i].hi[a];
    }
    B.nodes.resize(2 * N + 2);
    B.node_alloc = 1;
    int hw = (int)std::thread::hardware_concurrency();
    B.threads_free = hw > 1 ? hw - 1 : 0;
    B.build(0, 0, (uint32_t)N, 1);
    out->bvh2_depth = B.max_depth.load();

    // ---- SAH-optimal collapse: C(n,i) = cheapest way to put the subtree of n into at most i child slots ----
    //   leaf n:            C(n,i)   = A(n) * P(n) * c_prim
    //   D(n,j)            = min_{0<k<j} C(left,k) + C(right,j-k)
    //   C(n,1)            = A(n) * c_node + D(n,8)            (n becomes a wide node of its own)
    //   C(n,i), i = 2..7  = min(D(n,i), C(n,i-1))
    // Children are allocated after their parent, so a descending index sweep is bottom-up.
    struct Dp { float c[7]; uint8_t split[8]; };      // c[i-1] = C(n,i); split[j-1] = k chosen for D(n,j) (0: inherit C(n,j-1))
    const uint32_t n_b2 = B.node_alloc.load();
    std::vector<Dp> dp(n_b2);
    {
        const double c_node = 1.0;
        const double c_prim = std::getenv("PTB_WIDE_CPRIM") ? std::atof(std::getenv("PTB_WIDE_CPRIM")) : 0.4;
        for (uint32_t ni = n_b2; ni-- > 0;) {
            const B2& n = B.nodes[ni];
            Dp& d = dp[ni];
            std::memset(&d, 0, sizeof d);
            const double area = n.b.area();
            if (n.count > 0) { for (int i = 0; i < 7; i++) d.c[i] = (float)(area * n.count * c_prim); continue; }
            if (ni != 0 && n.left == 0) continue;     // unused slot of the pair allocator
            const Dp& l = dp[n.left]; const Dp& r = dp[n.right];
            double D[9]; int K[9];
            for (int j = 2; j <= 8; j++) {
                D[j] = INFINITY; K[j] = 1;
                for (int k = 1; k < j; k++) {
                    if (k > 7 || j - k > 7) continue;
                    const double v = (double)l.c[k - 1] + (double)r.c[j - k - 1];
                    if (v < D[j]) { D[j] = v; K[j] = k; }
                }
            }
            d.c[0] = (float)(area * c_node + D[8]);
            d.split[7] = (uint8_t)K[8];
            for (int i = 2; i <= 7; i++) {
                if (D[i] < d.c[i - 2]) { d.c[i - 1] = (float)D[i]; d.split[i - 1] = (uint8_t)K[i]; }
                else { d.c[i - 1] = d.c[i - 2]; d.split[i - 1] = 0; }
            }
        }
    }

    // ---- assign slots, quantise, emit breadth-first ----
    struct Task { uint32_t b2, index, depth; };
    std::vector<Task> queue;
    queue.push_back({0u, 0u, 1u});
    uint32_t next_node = 1;
    out->nodes.reserve((N / 2 + 16) * 20);
    out->prims.reserve(N * 12);
    out->nodes.resize(20);
    const double root_area = std::max(B.nodes[0].b.area(), 1e-300);
    double sah = 0.0;
    uint32_t depth_max = 0;
    for (size_t qi = 0; qi < queue.size(); qi++) {
        Task t = queue[qi];
        depth_max = std::max(depth_max, t.depth);
        const B2& self = B.nodes[t.b2];
        uint32_t ch[8]; int nch = 0;
        if (self.count > 0) { ch[nch++] = t.b2; }
        else {
            // children of this wide node = the optimal 8-slot forest below the BVH2 node
            struct Item { uint32_t n; int slots; };
            Item work[16]; int nw = 0;
            const int k8 = dp[t.b2].split[7];
            work[nw++] = {self.right, 8 - k8};
            work[nw++] = {self.left, k8};
            while (nw > 0) {
                Item it = work[--nw];
                const B2& node = B.nodes[it.n];
                int i = it.slots;
                if (node.count > 0 || i == 1) { ch[nch++] = it.n; continue; }
                while (i > 1 && dp[it.n].split[i - 1] == 0) i--;          // C(n,i) inherited from C(n,i-1)
                if (i == 1) { ch[nch++] = it.n; continue; }
                const int k = dp[it.n].split[i - 1];
                work[nw++] = {node.right, i - k};
                work[nw++] = {node.left, k};
            }
        }
        sah += self.b.area() / root_area;
        // octant slot assignment: greedy on cost(c, s) = (centroid_c - centroid_node) . sign_s
        int slot_of[8]; bool slot_used[8] = {false}; bool child_done[8] = {false};
        float cost[8][8];
        float nc[3];
        for (int a = 0; a < 3; a++) nc[a] = 0.5f * self.b.lo[a] + 0.5f * self.b.hi[a];
        for (int c = 0; c < nch; c++) {
            const Box& cb = B.nodes[ch[c]].b;
            float d[3];
            for (int a = 0; a < 3; a++) d[a] = (0.5f * cb.lo[a] + 0.5f * cb.hi[a]) - nc[a];
            for (int s = 0; s < 8; s++) cost[c][s] = ((s & 1) ? d[0] : -d[0]) + ((s & 2) ? d[1] : -d[1]) + ((s & 4) ? d[2] : -d[2]);
        }
        for (int it = 0; it < nch; it++) {
            int bc = -1, bs = -1; float bv = -INFINITY;
            for (int c = 0; c < nch; c++) if (!child_done[c]) for (int s = 0; s < 8; s++) if (!slot_used[s] && cost[c][s] > bv) { bv = cost[c][s]; bc = c; bs = s; }
            child_done[bc] = true; slot_used[bs] = true; slot_of[bc] = bs;
        }
        int child_in_slot[8];
        for (int s = 0; s < 8; s++) child_in_slot[s] = -1;
        for (int c = 0; c < nch; c++) child_in_slot[slot_of[c]] = c;

        uint32_t w[20];
        std::memset(w, 0, sizeof w);
        uint32_t ebyte[3]; double escale[3];
        for (int a = 0; a < 3; a++) {
            ebyte[a] = exponent_byte((double)self.b.hi[a] - (double)self.b.lo[a]);
            escale[a] = std::ldexp(1.0, (int)ebyte[a] - 127);
            std::memcpy(&w[a], &self.b.lo[a], 4);
        }
        uint32_t imask = 0;
        uint32_t valid24 = 0; uint8_t q[6][8];
        std::memset(q, 0, sizeof q);
        const uint32_t child_base = next_node, prim_base = (uint32_t)(out->prims.size() / 12);
        for (int s = 0; s < 8; s++) {
            int c = child_in_slot[s];
            if (c < 0) continue;
            const B2& cn = B.nodes[ch[c]];
            for (int a = 0; a < 3; a++) {
                double lo = std::floor(((double)cn.b.lo[a] - (double)self.b.lo[a]) / escale[a]);
                double hi = std::ceil(((double)cn.b.hi[a] - (double)self.b.lo[a]) / escale[a]);
                lo = std::min(255.0, std::max(0.0, lo)); hi = std::min(255.0, std::max(0.0, hi));
                q[a][s] = (uint8_t)lo; q[3 + a][s] = (uint8_t)hi;
            }
            if (cn.count == 0) {
                imask |= 1u << s;
                queue.push_back({ch[c], next_node++, t.depth + 1});
            } else {
                valid24 |= ((1u << cn.count) - 1u) << (3 * s);     // slot s owns bits 3s..3s+2; primitives are stored compacted in bit order
                for (uint32_t k = 0; k < cn.count; k++) {
                    const Prim& p = B.prims[B.idx[cn.first + k]];
                    float rec[12];
                    std::memset(rec, 0, sizeof rec);
                    if (p.kind == 0) {
                        const ptb_packed_mesh_face& F = v->MeshFaces[p.face];
                        std::memcpy(&rec[0], F.Position0, 12); std::memcpy(&rec[4], F.Position1, 12); std::memcpy(&rec[8], F.Position2, 12);
                    }
                    std::memcpy(&rec[3], &p.face, 4); std::memcpy(&rec[7], &p.shape, 4); std::memcpy(&rec[11], &p.tag, 4);
                    out->prims.insert(out->prims.end(), rec, rec + 12);
                }
                sah += cn.b.area() / root_area * cn.count;
            }
        }
        w[3] = ebyte[0] | (ebyte[1] << 8) | (ebyte[2] << 16) | (imask << 24);
        w[4] = child_base; w[5] = prim_base;
        w[6] = valid24; w[7] = 0;
        for (int a = 0; a < 6; a++) std::memcpy(&w[8 + 2 * a], q[a], 8);   // qlo x,y,z then qhi x,y,z
        if (out->nodes.size() < ((size_t)next_node) * 20) out->nodes.resize((size_t)next_node * 20);
        std::memcpy(&out->nodes[(size_t)t.index * 20], w, sizeof w);
    }
    out->node_count = next_node;
    out->depth = depth_max;
    out->sah_cost = sah;
    out->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (depth_max > 30) { *err = "wide BVH deeper than the traversal stack"; return PTB_E_UNSUPPORTED; }
    return PTB_OK;
}

}  // namespace ptb
