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
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#if defined(__SSE2__) && !defined(PTB_NO_SSE)
#include <immintrin.h>
#endif

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

// Accumulators for the passes of the build: the same min / max as Box::grow, operand order included (std::min(acc, x)
// is "x < acc ? x : acc", which is _mm_min_ps(x, acc)), four lanes at a time where SSE is there.
#if defined(__SSE2__) && !defined(PTB_NO_SSE)
struct BoxAcc {                       // lanes 0-2; grow(Box) reads one float past the Box (the array is padded)
    __m128 lo, hi;
    void reset() { lo = _mm_set1_ps(INFINITY); hi = _mm_set1_ps(-INFINITY); }
    void grow(const Box& o) {
        const float* f = reinterpret_cast<const float*>(&o);
        lo = _mm_min_ps(_mm_loadu_ps(f), lo); hi = _mm_max_ps(_mm_loadu_ps(f + 3), hi);
    }
    void grow(const BoxAcc& o) { lo = _mm_min_ps(o.lo, lo); hi = _mm_max_ps(o.hi, hi); }
    Box get() const {
        float a[4], b[4]; _mm_storeu_ps(a, lo); _mm_storeu_ps(b, hi);
        Box r; for (int k = 0; k < 3; k++) { r.lo[k] = a[k]; r.hi[k] = b[k]; } return r;
    }
};
struct PointAcc {                     // lanes 0-2; grow reads four floats
    __m128 lo, hi;
    void reset() { lo = _mm_set1_ps(INFINITY); hi = _mm_set1_ps(-INFINITY); }
    void grow(const float* p) { __m128 v = _mm_loadu_ps(p); lo = _mm_min_ps(v, lo); hi = _mm_max_ps(v, hi); }
    void grow(const PointAcc& o) { lo = _mm_min_ps(o.lo, lo); hi = _mm_max_ps(o.hi, hi); }
    Box get() const {
        float a[4], b[4]; _mm_storeu_ps(a, lo); _mm_storeu_ps(b, hi);
        Box r; for (int k = 0; k < 3; k++) { r.lo[k] = a[k]; r.hi[k] = b[k]; } return r;
    }
};
#else
struct BoxAcc {
    Box b;
    void reset() { b.reset(); }
    void grow(const Box& o) { b.grow(o); }
    void grow(const BoxAcc& o) { b.grow(o.b); }
    Box get() const { return b; }
};
struct PointAcc {
    Box b;
    void reset() { b.reset(); }
    void grow(const float* p) { b.grow(p); }
    void grow(const PointAcc& o) { b.grow(o.b); }
    Box get() const { return b; }
};
#endif

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

// Collapse costs of one BVH2 node (see build_wide_bvh): c[i-1] = C(n,i); split[j-1] = k chosen for D(n,j) (0: inherit C(n,j-1)).
struct Dp { float c[7]; uint8_t split[8]; };

// fn() on up to `useful` host threads (the caller's included); fn takes its work from a shared counter.
template <typename F> void run_on_threads(size_t useful, F&& fn) {
    const size_t hw = std::max(1u, std::thread::hardware_concurrency());
    const size_t T = std::max<size_t>(1, std::min(hw, useful));
    std::vector<std::thread> pool;
    for (size_t t = 1; t < T; t++) pool.emplace_back(fn);
    fn();
    for (auto& th : pool) th.join();
}

// Storage the build threads fill: not cleared first (2N nodes of 40 + 36 bytes are 0.7 GB for 10 M primitives), the
// pages are first touched by the thread that builds the subtree.
template <typename T> struct Uninit {
    T* p = nullptr;
    ~Uninit() { std::free(p); }
    bool alloc(size_t n) { std::free(p); p = (T*)std::malloc(std::max<size_t>(1, n) * sizeof(T)); return p != nullptr; }
    T& operator[](size_t i) { return p[i]; }
    const T& operator[](size_t i) const { return p[i]; }
};

struct Builder {
    static constexpr uint32_t PIECEWISE = 262144;    // nodes with more primitives cut their passes into pieces
    static constexpr uint32_t MAX_PIECES = 32;
    static constexpr uint32_t SUBTREE_TASK = 8192;   // both sides larger: the left subtree may go to another thread
    static constexpr uint32_t ARENA = 4096;          // node slots a thread takes from the shared counter at a time
    Uninit<Dp> dp;
    double c_node = 1.0, c_prim = 0.4;
    std::vector<Box> pbox;
    std::vector<float> pcen;          // 3 per prim
    std::vector<Prim> prims;
    std::vector<uint32_t> idx;
    Uninit<B2> nodes;
    size_t node_capacity = 0;
    std::atomic<uint32_t> node_alloc{0};
    std::atomic<int> threads_free{0};
    std::atomic<uint32_t> max_depth{0};

    // Node pairs come from per-thread runs of ARENA slots: one shared counter bumped per pair would be the most
    // contended cache line of the build.  Numbering therefore depends on the thread schedule; nothing reads it
    // (the wide tree is emitted by walking from the root).
    struct Arena { uint32_t next = 0, end = 0; };
    uint32_t alloc2(Arena& a) {
        if (a.next + 2 > a.end) { a.next = node_alloc.fetch_add(ARENA); a.end = a.next + ARENA; }
        uint32_t c = a.next; a.next += 2; return c;
    }

    void make_leaf(uint32_t n, uint32_t first, uint32_t count, const Box& b) {
        nodes[n].b = b; nodes[n].left = nodes[n].right = 0; nodes[n].first = first; nodes[n].count = count;
        Dp& d = dp[n];
        std::memset(&d, 0, sizeof d);
        for (int i = 0; i < 7; i++) d.c[i] = (float)(b.area() * count * c_prim);
    }

    // SAH-optimal collapse (recurrence in build_wide_bvh), for a node whose children are done.
    void collapse_costs(uint32_t ni) {
        const B2& n = nodes[ni];
        Dp& d = dp[ni];
        std::memset(&d, 0, sizeof d);
        const double area = n.b.area();
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

    int take_threads(int wanted) {
        int f = threads_free.load();
        while (f > 0) {
            int take = std::min(f, wanted);
            if (threads_free.compare_exchange_weak(f, f - take)) return take;
        }
        return 0;
    }

    // body(piece, first, last) over [first, first + count) in `pieces` contiguous parts, one thread each
    template <typename F> void in_pieces(uint32_t first, uint32_t count, unsigned pieces, F const& body) {
        if (pieces <= 1) { body(0u, first, first + count); return; }
        std::vector<std::thread> pool;
        for (unsigned t = 1; t < pieces; t++)
            pool.emplace_back([&, t]() { body(t, first + (uint32_t)((uint64_t)count * t / pieces), first + (uint32_t)((uint64_t)count * (t + 1) / pieces)); });
        body(0u, first, first + (uint32_t)((uint64_t)count / pieces));
        for (auto& th : pool) th.join();
    }

    void build(Arena& arena, uint32_t n, uint32_t first, uint32_t count, uint32_t depth) {
        // Large nodes cut their two passes into pieces; partial boxes and counts merge exactly (min / max, integers).
        const int extra = count >= PIECEWISE ? take_threads((int)std::min<uint32_t>(MAX_PIECES - 1, count / (PIECEWISE / 4))) : 0;
        const unsigned pieces = 1 + extra;
        Box b, cb;
        {
            BoxAcc pb[MAX_PIECES]; PointAcc pc[MAX_PIECES];
            in_pieces(first, count, pieces, [&](unsigned t, uint32_t i0, uint32_t i1) {
                BoxAcc ab; PointAcc ac; ab.reset(); ac.reset();
                for (uint32_t i = i0; i < i1; i++) { ab.grow(pbox[idx[i]]); ac.grow(&pcen[3 * (size_t)idx[i]]); }
                pb[t] = ab; pc[t] = ac;
            });
            for (unsigned t = 1; t < pieces; t++) { pb[0].grow(pb[t]); pc[0].grow(pc[t]); }
            b = pb[0].get(); cb = pc[0].get();
        }
        uint32_t d = max_depth.load();
        while (depth > d && !max_depth.compare_exchange_weak(d, depth)) {}
        if (count <= MAX_LEAF) { if (extra) threads_free.fetch_add(extra); make_leaf(n, first, count, b); return; }

        // binned SAH over the centroid bounds, 3 axes
        int best_axis = -1, best_split = -1;
        double best_cost = INFINITY;
        float scale[3];
        for (int a = 0; a < 3; a++) { float e = cb.hi[a] - cb.lo[a]; scale[a] = e > 0 ? (float)BINS * (1.0f - 1e-6f) / e : 0.0f; }
        if (depth < 60) {
            struct Bins { BoxAcc box[3][BINS]; uint32_t cnt[3][BINS]; };
            Bins own;
            std::vector<Bins> others(pieces - 1);
            in_pieces(first, count, pieces, [&](unsigned t, uint32_t i0, uint32_t i1) {
                Bins& B_ = t == 0 ? own : others[t - 1];
                for (int a = 0; a < 3; a++) for (int k = 0; k < BINS; k++) { B_.box[a][k].reset(); B_.cnt[a][k] = 0; }
                for (uint32_t i = i0; i < i1; i++) {
                    uint32_t p = idx[i];
                    for (int a = 0; a < 3; a++) {
                        int k = (int)((pcen[3 * (size_t)p + a] - cb.lo[a]) * scale[a]);
                        k = k < 0 ? 0 : (k >= BINS ? BINS - 1 : k);
                        B_.box[a][k].grow(pbox[p]); B_.cnt[a][k]++;
                    }
                }
            });
            for (const Bins& o : others)
                for (int a = 0; a < 3; a++) for (int k = 0; k < BINS; k++) { own.box[a][k].grow(o.box[a][k]); own.cnt[a][k] += o.cnt[a][k]; }
            BoxAcc (&bins)[3][BINS] = own.box; uint32_t (&cnt)[3][BINS] = own.cnt;
            for (int a = 0; a < 3; a++) {
                if (scale[a] == 0.0f) continue;
                double la[BINS]; uint32_t ln[BINS];
                BoxAcc acc; acc.reset(); uint32_t c = 0;
                for (int k = 0; k < BINS - 1; k++) { acc.grow(bins[a][k]); c += cnt[a][k]; la[k] = acc.get().area(); ln[k] = c; }
                acc.reset(); c = 0;
                for (int k = BINS - 1; k > 0; k--) {
                    acc.grow(bins[a][k]); c += cnt[a][k];
                    if (ln[k - 1] == 0 || c == 0) continue;
                    double cost = la[k - 1] * ln[k - 1] + acc.get().area() * c;
                    if (cost < best_cost) { best_cost = cost; best_axis = a; best_split = k; }
                }
            }
        }
        if (extra) threads_free.fetch_add(extra);
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
        uint32_t c = alloc2(arena);
        nodes[n].b = b; nodes[n].left = c; nodes[n].right = c + 1; nodes[n].first = 0; nodes[n].count = 0;
        uint32_t nl = mid - first, nr = count - nl;
        bool spawned = false;
        std::thread th;
        if (std::min(nl, nr) > SUBTREE_TASK && take_threads(1) == 1) {
            spawned = true;
            th = std::thread([=]() { Arena own; build(own, c, first, nl, depth + 1); threads_free.fetch_add(1); });
        }
        if (!spawned) build(arena, c, first, nl, depth + 1);
        build(arena, c + 1, mid, nr, depth + 1);
        if (spawned) th.join();
        collapse_costs(n);
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
    const bool timing = std::getenv("PTB_WIDE_TIMING") != nullptr;
    auto lap = [&, last = t0](const char* what) mutable {
        if (!timing) return;
        auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[ptb wide build] %-10s %.3f s\n", what, std::chrono::duration<double>(now - last).count());
        last = now;
    };

    // ---- gather bounded primitives in world space ----
    // The primitive ORDER is part of the result (it seeds the partition order below): shapes in index order, the faces
    // of a mesh in the order its BVH2 leaves are reached depth-first, left child first.  The walk over the nodes only
    // lists those leaf runs; boxes are then computed for pieces of the list on all host threads, each piece writing
    // where it would have landed in a sequential pass.
    struct Run { uint32_t shape, begin, end; };          // faces [begin, end) of one mesh leaf; begin == end: an analytic shape
    std::vector<Run> runs;
    {
        std::vector<uint32_t> stack;
        for (uint32_t s = 0; s < v->ShapeCount; s++) {
            const ptb_packed_shape& sh = v->Shapes[s];
            if (sh.Type == 1) { out->planes.push_back(s); continue; }
            if (sh.Type == 2 || sh.Type == 3) { runs.push_back({s, 0u, 0u}); continue; }
            if (sh.Type != 0) continue;
            if (sh.MeshRootNodeIndex >= v->MeshNodeCount) { *err = "shape mesh root out of range"; return PTB_E_INVALID; }
            stack.assign(1, sh.MeshRootNodeIndex);
            while (!stack.empty()) {
                uint32_t ni = stack.back(); stack.pop_back();
                if (ni >= v->MeshNodeCount) { *err = "mesh node index out of range"; return PTB_E_INVALID; }
                const ptb_packed_mesh_node& n = v->MeshNodes[ni];
                if (n.FaceEndIndex > 0) {
                    if (n.FaceEndIndex > v->MeshFaceCount || n.FaceBeginOrNodeIndex > n.FaceEndIndex) { *err = "mesh face range out of range"; return PTB_E_INVALID; }
                    if (n.FaceBeginOrNodeIndex == n.FaceEndIndex) continue;
                    if (!runs.empty() && runs.back().shape == s && runs.back().end == n.FaceBeginOrNodeIndex && runs.back().begin != runs.back().end)
                        runs.back().end = n.FaceEndIndex;                       // leaves of one mesh follow each other
                    else
                        runs.push_back({s, n.FaceBeginOrNodeIndex, n.FaceEndIndex});
                } else {
                    stack.push_back(n.FaceBeginOrNodeIndex + 1);
                    stack.push_back(n.FaceBeginOrNodeIndex);
                }
            }
        }
    }
    lap("walk");
    {
        // pieces of at most PIECE primitives; a run longer than that is cut
        constexpr uint32_t PIECE = 65536;
        struct Piece { uint32_t shape, begin, end; size_t at; uint32_t made; };
        std::vector<Piece> pieces;
        size_t upper = 0;
        for (const Run& r : runs) {
            if (r.begin == r.end) { pieces.push_back({r.shape, 0u, 0u, upper, 0u}); upper += 1; continue; }
            for (uint32_t b = r.begin; b < r.end; b += PIECE) {
                const uint32_t e = (uint32_t)std::min<uint64_t>(r.end, (uint64_t)b + PIECE);
                pieces.push_back({r.shape, b, e, upper, 0u});
                upper += e - b;
            }
        }
        if (upper + out->planes.size() >= (1ull << 27)) { *err = "more than 2^27 primitives"; return PTB_E_UNSUPPORTED; }
        B.pbox.resize(upper + 1);                      // + 1: BoxAcc reads one float past a Box
        B.prims.resize(upper);
        std::atomic<size_t> next(0);
        auto work = [&]() {
            for (size_t k = next.fetch_add(16); k < pieces.size(); k = next.fetch_add(16))
                for (size_t q = k; q < std::min(pieces.size(), k + 16); q++) {
                    Piece& pc = pieces[q];
                    const ptb_packed_shape& sh = v->Shapes[pc.shape];
                    const float* to = sh.Transform.To;
                    size_t at = pc.at;
                    if (pc.begin == pc.end) {
                        Box b; b.reset();
                        for (int c = 0; c < 8; c++) {
                            float p[3] = {(c & 1) ? 1.f : -1.f, (c & 2) ? 1.f : -1.f, (c & 4) ? 1.f : -1.f};
                            double w[3]; xform_point(to, p, w);
                            float wf[3] = {(float)w[0], (float)w[1], (float)w[2]};
                            b.grow(wf);
                        }
                        pad_box(b);
                        B.pbox[at] = b; B.prims[at] = {pc.shape, 0u, (uint32_t)sh.Type, 0u};
                        pc.made = 1;
                        continue;
                    }
                    for (uint32_t f = pc.begin; f < pc.end; f++) {
                        const ptb_packed_mesh_face& F = v->MeshFaces[f];
                        Box b; b.reset();
                        const float* P[3] = {F.Position0, F.Position1, F.Position2};
                        for (int k3 = 0; k3 < 3; k3++) {
                            double w[3]; xform_point(to, P[k3], w);
                            float wf[3] = {(float)w[0], (float)w[1], (float)w[2]};
                            b.grow(wf);
                        }
                        if (!(std::isfinite(b.lo[0]) && std::isfinite(b.lo[1]) && std::isfinite(b.lo[2]) &&
                              std::isfinite(b.hi[0]) && std::isfinite(b.hi[1]) && std::isfinite(b.hi[2]))) continue;
                        pad_box(b);
                        B.pbox[at] = b; B.prims[at] = {pc.shape, f, 0u, 0u};
                        at++;
                    }
                    pc.made = (uint32_t)(at - pc.at);
                }
        };
        run_on_threads(pieces.size() / 16 + 1, work);
        // close the gaps left by faces without a finite box (none, normally)
        size_t total = 0;
        bool gaps = false;
        for (const Piece& pc : pieces) { if (pc.at != total) gaps = true; total += pc.made; }
        if (gaps) {
            size_t to_ = 0;
            for (const Piece& pc : pieces) {
                if (pc.at != to_) { std::move(B.pbox.begin() + pc.at, B.pbox.begin() + pc.at + pc.made, B.pbox.begin() + to_);
                                    std::move(B.prims.begin() + pc.at, B.prims.begin() + pc.at + pc.made, B.prims.begin() + to_); }
                to_ += pc.made;
            }
        }
        B.pbox.resize(total + 1);
        B.prims.resize(total);
    }
    const size_t N = B.prims.size();
    out->prim_count = (uint32_t)N;
    lap("gather");
    if (N + out->planes.size() >= (1ull << 27)) { *err = "more than 2^27 primitives"; return PTB_E_UNSUPPORTED; }
    {
        // tag = rank << 3 | identity << 2 | kind.  rank = position in the canonical (shape, face) order, so a
        // 64-bit key {hit time bits, tag} orders hits exactly like the canonical tie rule; identity = the shape's
        // transform is the identity matrix (object ray = world ray + 0.0f, no matrix fetch); kind = shape type & 3.
        std::vector<uint8_t> identity(v->ShapeCount);
        for (uint32_t s = 0; s < v->ShapeCount; s++) {
            static const float I[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
            identity[s] = !std::memcmp(v->Shapes[s].Transform.From, I, 64) && !std::memcmp(v->Shapes[s].Transform.To, I, 64);
        }
        auto key_of = [&](size_t i) { return ((uint64_t)B.prims[i].shape << 32) | B.prims[i].face; };
        bool ordered = true;
        for (size_t i = 1; i < N && ordered; i++) ordered = key_of(i - 1) <= key_of(i);
        std::vector<uint32_t> plane_tags(out->planes.size());
        if (ordered) {
            // The usual case: the gathering order IS the canonical order, up to where the planes go.  A plane shape
            // has no other primitive, so rank(primitive i) = i + planes of a lower shape index, and the other way round.
            std::vector<uint32_t> planes_before(v->ShapeCount + 1, 0u), prims_before(v->ShapeCount + 1, 0u);
            for (uint32_t s : out->planes) planes_before[s + 1]++;
            for (size_t i = 0; i < N; i++) prims_before[B.prims[i].shape + 1]++;
            for (uint32_t s = 0; s < v->ShapeCount; s++) { planes_before[s + 1] += planes_before[s]; prims_before[s + 1] += prims_before[s]; }
            std::atomic<size_t> next(0);
            constexpr size_t STEP = 1 << 16;
            run_on_threads(N / STEP + 1, [&]() {
                for (size_t b0 = next.fetch_add(STEP); b0 < N; b0 = next.fetch_add(STEP))
                    for (size_t i = b0; i < std::min(N, b0 + STEP); i++) {
                        Prim& p = B.prims[i];
                        p.tag = ((uint32_t)(i + planes_before[p.shape]) << 3) | (identity[p.shape] ? 4u : 0u) | (p.kind & 3u);
                    }
            });
            for (size_t j = 0; j < out->planes.size(); j++) plane_tags[j] = ((uint32_t)(j + prims_before[out->planes[j]]) << 3) | 1u;
        } else {
            std::vector<std::pair<uint64_t, uint32_t>> order;
            order.reserve(N + out->planes.size());
            for (size_t i = 0; i < N; i++) order.push_back({key_of(i), (uint32_t)i});
            for (size_t i = 0; i < out->planes.size(); i++) order.push_back({(uint64_t)out->planes[i] << 32, (uint32_t)(N + i)});
            std::sort(order.begin(), order.end());
            for (size_t r = 0; r < order.size(); r++) {
                const uint32_t i = order[r].second;
                if (i < N) B.prims[i].tag = ((uint32_t)r << 3) | (identity[B.prims[i].shape] ? 4u : 0u) | (B.prims[i].kind & 3u);
                else plane_tags[i - N] = ((uint32_t)r << 3) | 1u;
            }
        }
        std::vector<uint32_t> pl;
        for (size_t i = 0; i < out->planes.size(); i++) { pl.push_back(out->planes[i]); pl.push_back(plane_tags[i]); }
        out->planes.swap(pl);      // {shape, tag} pairs
    }
    lap("tags");
    if (N == 0) { out->node_count = 0; return PTB_OK; }
    if (N > 0x7FFFFFFFull) { *err = "too many primitives for the wide BVH"; return PTB_E_UNSUPPORTED; }

    B.pcen.resize(3 * N + 1);                         // PointAcc reads four floats
    B.idx.resize(N);
    {
        std::atomic<size_t> next(0);
        constexpr size_t STEP = 1 << 16;
        run_on_threads(N / STEP + 1, [&]() {
            for (size_t b0 = next.fetch_add(STEP); b0 < N; b0 = next.fetch_add(STEP))
                for (size_t i = b0; i < std::min(N, b0 + STEP); i++) {
                    B.idx[i] = (uint32_t)i;
                    for (int a = 0; a < 3; a++) B.pcen[3 * i + a] = 0.5f * B.pbox[i].lo[a] + 0.5f * B.pbox[i].hi[a];
                }
        });
    }
    // at most 2N - 1 nodes; every thread that ever builds a subtree may leave one arena partly used
    B.node_capacity = 2 * N + 2 + ((size_t)N / Builder::SUBTREE_TASK + 2) * Builder::ARENA;
    if (B.node_capacity >= 0xFFFFFFFFull || !B.nodes.alloc(B.node_capacity) || !B.dp.alloc(B.node_capacity)) { *err = "out of host memory for the BVH build"; return PTB_E_NOMEM; }
    if (std::getenv("PTB_WIDE_CPRIM")) B.c_prim = std::atof(std::getenv("PTB_WIDE_CPRIM"));
    B.node_alloc = 1;
    lap("alloc");
    int hw = (int)std::thread::hardware_concurrency();
    B.threads_free = hw > 1 ? hw - 1 : 0;
    { Builder::Arena root_arena; B.build(root_arena, 0, 0, (uint32_t)N, 1); }
    out->bvh2_depth = B.max_depth.load();
    lap("bvh2");

    // ---- SAH-optimal collapse: C(n,i) = cheapest way to put the subtree of n into at most i child slots ----
    //   leaf n:            C(n,i)   = A(n) * P(n) * c_prim
    //   D(n,j)            = min_{0<k<j} C(left,k) + C(right,j-k)
    //   C(n,1)            = A(n) * c_node + D(n,8)            (n becomes a wide node of its own)
    //   C(n,i), i = 2..7  = min(D(n,i), C(n,i-1))
    const Uninit<Dp>& dp = B.dp;                      // filled by the build, each node after its children
    lap("collapse");
    // ---- assign slots, quantise, emit breadth-first ----
    // Node t of the wide tree gets the indices child_base = (nodes emitted before its children) and prim_base =
    // (primitives emitted before its own), in breadth-first order.  Level by level: every node of the level chooses
    // its children and their slots (independent, all host threads), one prefix sum over the level hands out the
    // bases in order, then every node writes its record, its primitives and its children's entries of the next level.
    struct Task { uint32_t b2, index, depth; };
    struct Plan { uint32_t child[8]; uint32_t internal, prims; double sah; };     // child[s] = BVH2 node in slot s (or none)
    constexpr uint32_t NONE = 0xFFFFFFFFu;
    const double root_area = std::max(B.nodes[0].b.area(), 1e-300);
    double sah = 0.0;
    uint32_t depth_max = 0;
    uint32_t next_node = 1;
    size_t prim_total = 0;
    out->nodes.reserve((N / 2 + 16) * 20);
    out->prims.resize(N * 12);                         // every primitive is emitted exactly once
    out->nodes.resize(20);
    std::vector<Task> level{{0u, 0u, 1u}}, next_level;
    std::vector<Plan> plans;
    std::vector<uint32_t> child_base, prim_base;

    auto plan_node = [&](const Task& t, Plan& pl) {
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
        for (int s = 0; s < 8; s++) pl.child[s] = NONE;
        for (int c = 0; c < nch; c++) pl.child[slot_of[c]] = ch[c];
        pl.internal = 0; pl.prims = 0;
        pl.sah = self.b.area() / root_area;
        for (int s = 0; s < 8; s++) {
            if (pl.child[s] == NONE) continue;
            const B2& cn = B.nodes[pl.child[s]];
            if (cn.count == 0) pl.internal++;
            else { pl.prims += cn.count; pl.sah += cn.b.area() / root_area * cn.count; }
        }
    };

    auto write_node = [&](const Task& t, const Plan& pl, uint32_t cbase, uint32_t pbase, Task* children_out) {
        const B2& self = B.nodes[t.b2];
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
        uint32_t made_children = 0;
        float* prim_out = out->prims.data() + (size_t)pbase * 12;
        for (int s = 0; s < 8; s++) {
            if (pl.child[s] == NONE) continue;
            const B2& cn = B.nodes[pl.child[s]];
            for (int a = 0; a < 3; a++) {
                double lo = std::floor(((double)cn.b.lo[a] - (double)self.b.lo[a]) / escale[a]);
                double hi = std::ceil(((double)cn.b.hi[a] - (double)self.b.lo[a]) / escale[a]);
                lo = std::min(255.0, std::max(0.0, lo)); hi = std::min(255.0, std::max(0.0, hi));
                q[a][s] = (uint8_t)lo; q[3 + a][s] = (uint8_t)hi;
            }
            if (cn.count == 0) {
                imask |= 1u << s;
                children_out[made_children] = {pl.child[s], cbase + made_children, t.depth + 1};
                made_children++;
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
                    std::memcpy(prim_out, rec, sizeof rec);
                    prim_out += 12;
                }
            }
        }
        w[3] = ebyte[0] | (ebyte[1] << 8) | (ebyte[2] << 16) | (imask << 24);
        w[4] = cbase; w[5] = pbase;
        w[6] = valid24; w[7] = 0;
        for (int a = 0; a < 6; a++) std::memcpy(&w[8 + 2 * a], q[a], 8);   // qlo x,y,z then qhi x,y,z
        std::memcpy(&out->nodes[(size_t)t.index * 20], w, sizeof w);
    };

    while (!level.empty()) {
        const size_t L = level.size();
        depth_max = std::max(depth_max, level[0].depth);
        plans.resize(L); child_base.resize(L); prim_base.resize(L);
        constexpr size_t STEP = 256;
        {
            std::atomic<size_t> next(0);
            run_on_threads(L / STEP + 1, [&]() {
                for (size_t b0 = next.fetch_add(STEP); b0 < L; b0 = next.fetch_add(STEP))
                    for (size_t i = b0; i < std::min(L, b0 + STEP); i++) plan_node(level[i], plans[i]);
            });
        }
        const uint32_t level_child_base = next_node;
        for (size_t i = 0; i < L; i++) {
            child_base[i] = next_node; prim_base[i] = (uint32_t)prim_total;
            next_node += plans[i].internal; prim_total += plans[i].prims;
            sah += plans[i].sah;
        }
        if (prim_total > N) { *err = "internal error: wide BVH emitted a primitive twice"; return PTB_E_INVALID; }
        out->nodes.resize((size_t)next_node * 20);
        next_level.resize(next_node - level_child_base);
        {
            std::atomic<size_t> next(0);
            run_on_threads(L / STEP + 1, [&]() {
                for (size_t b0 = next.fetch_add(STEP); b0 < L; b0 = next.fetch_add(STEP))
                    for (size_t i = b0; i < std::min(L, b0 + STEP); i++)
                        write_node(level[i], plans[i], child_base[i], prim_base[i], next_level.data() + (child_base[i] - level_child_base));
            });
        }
        level.swap(next_level);
    }
    out->prims.resize(prim_total * 12);
    lap("emit");
    out->node_count = next_node;
    out->depth = depth_max;
    out->sah_cost = sah;
    out->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (depth_max > 30) { *err = "wide BVH deeper than the traversal stack"; return PTB_E_UNSUPPORTED; }
    return PTB_OK;
}

}  // namespace ptb
