// See ptb_gpu_build.h.  CUDA kernels for sm_100a; the only library call is cub's radix sort.
#include "ptb_gpu_build.h"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace ptb {
namespace {

constexpr uint32_t LEAF = 0x80000000u;      // node id flag: index into the sorted primitive order instead of the internal nodes
constexpr uint32_t NONE = 0xFFFFFFFFu;
constexpr uint32_t MAX_LEAF = 2;            // primitives per child slot, as in ptb_wide_build.cpp

struct ShapeTable {                          // one entry per shape, built on the host (tiny)
    const uint32_t* prim_base;               // [S+1] prefix sum of bounded primitives
    const uint32_t* face_begin;              // [S]   first global face of a mesh shape
    const uint32_t* rank_base;               // [S]   canonical rank of the shape's first primitive
    const uint32_t* flags;                   // [S]   type | identity << 8
    uint32_t shape_count;
};

struct Tree {
    // primitives (unsorted order)
    float4* box_lo;        // xyz = padded world lower corner, w = shape index bits
    float4* box_hi;        // xyz = upper corner,               w = face index bits
    uint32_t* tag;
    // sorted order
    const uint32_t* sorted;          // position -> primitive
    const unsigned long long* keys;  // position -> Morton key
    // internal nodes [0, n-1)
    uint32_t* left; uint32_t* right; uint32_t* cnt;     // cnt = primitives below the node
    uint32_t* parent;                // [n-1] internal, then [n] leaves
    float4* ibox_lo; float4* ibox_hi;
    uint32_t* visits;
    // SAH-optimal collapse (the forest-cost recurrence of Ylitie et al. 2017, as in ptb_wide_build.cpp):
    // dp_cost[7 n + i - 1] = C(n, i), cheapest way to put the subtree of n into at most i child slots;
    // dp_split[8 n + j - 1] = slots given to the left child in D(n, j) (0: C(n, j) inherited from C(n, j - 1)).
    float* dp_cost; uint8_t* dp_split;
    uint32_t n;
    uint32_t greedy;                 // 1: open the largest-area child instead (kept for comparison)
};

__device__ __forceinline__ void atomic_min_float(float* a, float v) {
    if (v >= 0.0f) atomicMin((int*)a, __float_as_int(v)); else atomicMax((unsigned*)a, __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_float(float* a, float v) {
    if (v >= 0.0f) atomicMax((int*)a, __float_as_int(v)); else atomicMin((unsigned*)a, __float_as_uint(v));
}

// World-space bounds of every bounded primitive, in the arithmetic of the host builder
// (transform in double, outward padding that covers the rounding of the object-space test).
// bounds6 = centroid bounds {lo xyz, hi xyz}; *invalid counts primitives with non-finite bounds.
__global__ void k_prim_bounds(ShapeTable T, const ptb_packed_shape* __restrict__ shapes, const ptb_packed_mesh_face* __restrict__ faces,
                              Tree t, float* bounds6, uint32_t* invalid) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float c[3] = {0, 0, 0};
    bool ok = false;
    if (i < t.n) {
        uint32_t lo = 0, hi = T.shape_count;               // last shape with prim_base <= i
        while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (T.prim_base[mid] <= i) lo = mid; else hi = mid; }
        const uint32_t s = lo, fl = T.flags[s], type = fl & 0xFFu, k = i - T.prim_base[s];
        const float* to = shapes[s].Transform.To;
        double m[12];
#pragma unroll
        for (int r = 0; r < 3; r++) { m[r] = to[r]; m[3 + r] = to[4 + r]; m[6 + r] = to[8 + r]; m[9 + r] = to[12 + r]; }
        float blo[3] = {INFINITY, INFINITY, INFINITY}, bhi[3] = {-INFINITY, -INFINITY, -INFINITY};
        uint32_t face = 0;
        bool nan = false;                                    // fminf/fmaxf drop a NaN operand
        auto grow = [&](float px, float py, float pz) {
#pragma unroll
            for (int r = 0; r < 3; r++) {
                const float w = (float)(m[r] * px + m[3 + r] * py + m[6 + r] * pz + m[9 + r]);
                blo[r] = fminf(blo[r], w); bhi[r] = fmaxf(bhi[r], w);
                nan = nan || !(w == w);
            }
        };
        if (type == 0u) {
            face = T.face_begin[s] + k;
            const float4* f = (const float4*)(faces + face);
            const float4 a = f[0], b = f[1], d = f[2];
            grow(a.x, a.y, a.z); grow(b.x, b.y, b.z); grow(d.x, d.y, d.z);
        } else {
            for (int q = 0; q < 8; q++) grow((q & 1) ? 1.f : -1.f, (q & 2) ? 1.f : -1.f, (q & 4) ? 1.f : -1.f);
        }
        ok = !nan;
#pragma unroll
        for (int r = 0; r < 3; r++) ok = ok && isfinite(blo[r]) && isfinite(bhi[r]);
        if (ok) {
#pragma unroll
            for (int r = 0; r < 3; r++) {
                const float mx = fmaxf(fabsf(blo[r]), fabsf(bhi[r]));
                const float e = 4e-6f * (mx + (bhi[r] - blo[r])) + 1e-30f;
                blo[r] -= e; bhi[r] += e;
                c[r] = 0.5f * blo[r] + 0.5f * bhi[r];
            }
        } else {
            blo[0] = blo[1] = blo[2] = bhi[0] = bhi[1] = bhi[2] = __int_as_float(0x7FC00000);
        }
        t.box_lo[i] = make_float4(blo[0], blo[1], blo[2], __uint_as_float(s));
        t.box_hi[i] = make_float4(bhi[0], bhi[1], bhi[2], __uint_as_float(face));
        t.tag[i] = ((T.rank_base[s] + k) << 3) | ((fl >> 8) & 1u) << 2 | (type & 3u);
    }
    // block reduction of the centroid bounds, then six atomics per block
    __shared__ float s_lo[3][8], s_hi[3][8];
    __shared__ uint32_t s_bad[8];
    float lo3[3], hi3[3];
#pragma unroll
    for (int r = 0; r < 3; r++) { lo3[r] = ok ? c[r] : INFINITY; hi3[r] = ok ? c[r] : -INFINITY; }
    uint32_t bad = (i < t.n && !ok) ? 1u : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < 3; r++) {
            lo3[r] = fminf(lo3[r], __shfl_xor_sync(0xFFFFFFFFu, lo3[r], o));
            hi3[r] = fmaxf(hi3[r], __shfl_xor_sync(0xFFFFFFFFu, hi3[r], o));
        }
        bad += __shfl_xor_sync(0xFFFFFFFFu, bad, o);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { for (int r = 0; r < 3; r++) { s_lo[r][warp] = lo3[r]; s_hi[r][warp] = hi3[r]; } s_bad[warp] = bad; }
    __syncthreads();
    if (threadIdx.x < 3) {
        const int r = threadIdx.x;
        float a = INFINITY, b = -INFINITY;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) { a = fminf(a, s_lo[r][w]); b = fmaxf(b, s_hi[r][w]); }
        if (a <= b) { atomic_min_float(&bounds6[r], a); atomic_max_float(&bounds6[3 + r], b); }
    }
    if (threadIdx.x == 32) {
        uint32_t sum = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) sum += s_bad[w];
        if (sum) atomicAdd(invalid, sum);
    }
}

__device__ __forceinline__ unsigned long long spread21(unsigned long long x) {
    x &= 0x1FFFFFull;
    x = (x | x << 32) & 0x001F00000000FFFFull;
    x = (x | x << 16) & 0x001F0000FF0000FFull;
    x = (x | x << 8) & 0x100F00F00F00F00Full;
    x = (x | x << 4) & 0x10C30C30C30C30C3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// 63-bit Morton key of the box centre inside the centroid bounds; non-finite primitives sort last.
__global__ void k_morton(Tree t, const float* __restrict__ bounds6, unsigned long long* keys, uint32_t* vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= t.n) return;
    const float4 lo = t.box_lo[i], hi = t.box_hi[i];
    unsigned long long key = ~0ull;
    if (lo.x == lo.x) {
        const float c[3] = {0.5f * lo.x + 0.5f * hi.x, 0.5f * lo.y + 0.5f * hi.y, 0.5f * lo.z + 0.5f * hi.z};
        unsigned long long q[3];
#pragma unroll
        for (int r = 0; r < 3; r++) {
            const float e = bounds6[3 + r] - bounds6[r];
            float u = e > 0.0f ? (c[r] - bounds6[r]) / e : 0.0f;
            u = fminf(fmaxf(u, 0.0f), 1.0f);
            q[r] = (unsigned long long)fminf(u * 2097152.0f, 2097151.0f);
        }
        key = spread21(q[0]) | (spread21(q[1]) << 1) | (spread21(q[2]) << 2);
    }
    keys[i] = key;
    vals[i] = i;
}

// Length of the common prefix of the keys at sorted positions i and j; equal keys are told apart by position.
__device__ __forceinline__ int delta(const unsigned long long* __restrict__ keys, uint32_t n, int i, int j) {
    if (j < 0 || j >= (int)n) return -1;
    const unsigned long long a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz(i ^ j);
    return __clzll((long long)(a ^ b));
}

// Binary radix tree over the sorted keys (Karras, "Maximizing Parallelism in the Construction of BVHs,
// Octrees, and k-d Trees", HPG 2012): internal node i covers [first, last] and splits after gamma.
__global__ void k_radix_tree(Tree t) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t n = t.n;
    if (i >= (int)n - 1) return;
    const unsigned long long* keys = t.keys;
    const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int s = lmax >> 1; s >= 1; s >>= 1) if (delta(keys, n, i, i + (l + s) * d) > dmin) l += s;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0, step = l;
    do {
        step = (step + 1) >> 1;
        if (delta(keys, n, i, i + (s + step) * d) > dnode) s += step;
    } while (step > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const uint32_t L = lo == gamma ? ((uint32_t)gamma | LEAF) : (uint32_t)gamma;
    const uint32_t R = hi == gamma + 1 ? ((uint32_t)(gamma + 1) | LEAF) : (uint32_t)(gamma + 1);
    t.left[i] = L; t.right[i] = R; t.cnt[i] = (uint32_t)(hi - lo + 1);
    t.parent[(L & LEAF) ? (n - 1) + (L & ~LEAF) : L] = (uint32_t)i;
    t.parent[(R & LEAF) ? (n - 1) + (R & ~LEAF) : R] = (uint32_t)i;
    if (i == 0) t.parent[0] = NONE;
}

__device__ __forceinline__ void node_box(const Tree& t, uint32_t id, float4& lo, float4& hi) {
    if (id & LEAF) { const uint32_t p = t.sorted[id & ~LEAF]; lo = t.box_lo[p]; hi = t.box_hi[p]; }
    else { lo = __ldcg(&t.ibox_lo[id]); hi = __ldcg(&t.ibox_hi[id]); }
}
__device__ __forceinline__ uint32_t node_count(const Tree& t, uint32_t id) { return (id & LEAF) ? 1u : t.cnt[id]; }

__device__ __forceinline__ double box_area(const float4& lo, const float4& hi) {
    const double x = (double)hi.x - lo.x, y = (double)hi.y - lo.y, z = (double)hi.z - lo.z;
    if (x < 0 || y < 0 || z < 0) return 0.0;
    return 2.0 * (x * y + y * z + z * x);
}
constexpr double C_NODE = 1.0, C_PRIM = 0.4;      // the host builder's cost model

__device__ __forceinline__ void load_dp(const Tree& t, uint32_t id, const float4& lo, const float4& hi, float* c) {
    if (id & LEAF) { const float v = (float)(box_area(lo, hi) * C_PRIM); for (int i = 0; i < 7; i++) c[i] = v; }
    else for (int i = 0; i < 7; i++) c[i] = __ldcg(&t.dp_cost[7 * (size_t)id + i]);
}

// Collapse costs of an internal node from those of its (finished) children.
__device__ __forceinline__ void node_costs(const Tree& t, uint32_t node, uint32_t L, uint32_t R, const float4& al, const float4& ah,
                                           const float4& bl, const float4& bh, double area, uint32_t count) {
    float cl[7], cr[7], cn[7];
    load_dp(t, L, al, ah, cl);
    load_dp(t, R, bl, bh, cr);
    double D[9]; int K[9];
    for (int j = 2; j <= 8; j++) {
        D[j] = INFINITY; K[j] = 1;
        for (int k = 1; k < j; k++) {
            if (k > 7 || j - k > 7) continue;
            const double v = (double)cl[k - 1] + (double)cr[j - k - 1];
            if (v < D[j]) { D[j] = v; K[j] = k; }
        }
    }
    uint8_t* sp = t.dp_split + 8 * (size_t)node;
    cn[0] = (float)(count <= MAX_LEAF ? area * count * C_PRIM : area * C_NODE + D[8]);
    sp[0] = 0; sp[7] = (uint8_t)K[8];
    for (int i = 2; i <= 7; i++) {
        if (D[i] < cn[i - 2]) { cn[i - 1] = (float)D[i]; sp[i - 1] = (uint8_t)K[i]; }
        else { cn[i - 1] = cn[i - 2]; sp[i - 1] = 0; }
    }
    for (int i = 0; i < 7; i++) __stcg(&t.dp_cost[7 * (size_t)node + i], cn[i]);
}

// Bottom-up bounds and collapse costs: one thread per leaf climbs; the second thread to arrive at a node computes it.
// ibox_lo.w carries the height of the subtree (levels, a leaf counts 1) for the depth diagnostic.
__global__ void k_fit_bounds(Tree t, uint32_t* max_depth) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= t.n) return;
    uint32_t node = t.parent[(t.n - 1) + i];
    uint32_t height = 1;
    while (node != NONE) {
        __threadfence();
        if (atomicAdd(&t.visits[node], 1u) == 0u) return;
        const uint32_t L = t.left[node], R = t.right[node];
        float4 al, ah, bl, bh;
        node_box(t, L, al, ah);
        node_box(t, R, bl, bh);
        const uint32_t hl = (L & LEAF) ? 1u : __float_as_uint(al.w), hr = (R & LEAF) ? 1u : __float_as_uint(bl.w);
        height = 1u + max(hl, hr);
        const float4 nl = make_float4(fminf(al.x, bl.x), fminf(al.y, bl.y), fminf(al.z, bl.z), __uint_as_float(height));
        const float4 nh = make_float4(fmaxf(ah.x, bh.x), fmaxf(ah.y, bh.y), fmaxf(ah.z, bh.z), 0.0f);
        __stcg(&t.ibox_lo[node], nl);
        __stcg(&t.ibox_hi[node], nh);
        node_costs(t, node, L, R, al, ah, bl, bh, box_area(nl, nh), t.cnt[node]);
        node = t.parent[node];
    }
    *max_depth = height;                // only the thread that completes the root gets here
}

__device__ __forceinline__ uint32_t exponent_byte(double extent) {     // smallest e with 255 * 2^e >= extent, biased
    if (!(extent > 0.0)) return 1u;
    int e;
    const double m = frexp(extent / 255.0, &e);
    if (m == 0.5) e -= 1;
    int biased = e + 127;
    biased = biased < 1 ? 1 : (biased > 254 ? 254 : biased);
    return (uint32_t)biased;
}

// ---- PLOC: parallel locally-ordered clustering (Meister & Bittner, "Parallel Locally-Ordered Clustering for
// Bounding Volume Hierarchy Construction", TVCG 2018).  The clusters start as the primitives in Morton order; per
// round every cluster looks PLOC_RADIUS positions left and right for the neighbour with which it makes the smallest
// box, mutual choices merge into a new node, and the survivors are compacted in order.  Same node arrays as the
// radix tree (children are created before parents, so collapse costs are computed at merge time).
constexpr int PLOC_RADIUS_MAX = 128;
constexpr int PLOC_BLOCK = 256;

struct Ploc {
    uint32_t* id; float4* lo; float4* hi;              // current clusters (compacted)
    uint32_t* id2; float4* lo2; float4* hi2;           // next round
    uint32_t* nn; uint32_t* keep; uint32_t* offset;    // nearest neighbour, survivor flag, exclusive scan of the flags
    uint32_t* node_counter; uint32_t* m_out;
};

__global__ void k_ploc_init(Tree t, Ploc p) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= t.n) return;
    const uint32_t prim = t.sorted[i];
    p.id[i] = i | LEAF; p.lo[i] = t.box_lo[prim]; p.hi[i] = t.box_hi[prim];
}

__global__ void __launch_bounds__(PLOC_BLOCK) k_ploc_nearest(Ploc p, uint32_t m, int PLOC_RADIUS) {
    __shared__ float4 s_lo[PLOC_BLOCK + 2 * PLOC_RADIUS_MAX], s_hi[PLOC_BLOCK + 2 * PLOC_RADIUS_MAX];
    const int base = (int)(blockIdx.x * PLOC_BLOCK) - PLOC_RADIUS;
    for (int k = threadIdx.x; k < PLOC_BLOCK + 2 * PLOC_RADIUS; k += PLOC_BLOCK) {
        const int g = base + k;
        if (g >= 0 && g < (int)m) { s_lo[k] = p.lo[g]; s_hi[k] = p.hi[g]; }
    }
    __syncthreads();
    const int i = (int)(blockIdx.x * PLOC_BLOCK + threadIdx.x);
    if (i >= (int)m) return;
    const float4 a = s_lo[threadIdx.x + PLOC_RADIUS], b = s_hi[threadIdx.x + PLOC_RADIUS];
    // total order on unordered pairs {i, j}: merged area, then "j is i's pair partner (i ^ 1)", then (min, max):
    // the smallest pair overall is a mutual choice, so every round merges, and equal boxes pair up at once
    float best = INFINITY; int bj = -1; bool bpartner = false;
    const int j0 = max(i - PLOC_RADIUS, 0), j1 = min(i + PLOC_RADIUS, (int)m - 1);
    for (int j = j0; j <= j1; j++) {
        if (j == i) continue;
        const float4 c = s_lo[j - base], d = s_hi[j - base];
        const float x = fmaxf(b.x, d.x) - fminf(a.x, c.x), y = fmaxf(b.y, d.y) - fminf(a.y, c.y), z = fmaxf(b.z, d.z) - fminf(a.z, c.z);
        const float area = x * y + y * z + z * x;
        const bool partner = j == (i ^ 1);
        bool better = area < best || bj < 0;
        if (!better && area == best) {
            if (partner != bpartner) better = partner;
            else {
                const int n0 = min(i, j), n1 = max(i, j), o0 = min(i, bj), o1 = max(i, bj);
                better = n0 < o0 || (n0 == o0 && n1 < o1);
            }
        }
        if (better) { best = area; bj = j; bpartner = partner; }
    }
    p.nn[i] = (uint32_t)bj;
}

__global__ void k_ploc_merge(Tree t, Ploc p, uint32_t m) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const uint32_t j = p.nn[i];
    const bool mutual = p.nn[j] == i;
    if (mutual && i > j) { p.keep[i] = 0u; return; }
    p.keep[i] = 1u;
    if (!mutual) return;
    const uint32_t node = atomicAdd(p.node_counter, 1u);
    const uint32_t L = p.id[i], R = p.id[j];
    const float4 al = p.lo[i], ah = p.hi[i], bl = p.lo[j], bh = p.hi[j];
    const uint32_t hl = (L & LEAF) ? 1u : __float_as_uint(t.ibox_lo[L].w), hr = (R & LEAF) ? 1u : __float_as_uint(t.ibox_lo[R].w);
    const uint32_t count = node_count(t, L) + node_count(t, R);
    const float4 nl = make_float4(fminf(al.x, bl.x), fminf(al.y, bl.y), fminf(al.z, bl.z), __uint_as_float(1u + max(hl, hr)));
    const float4 nh = make_float4(fmaxf(ah.x, bh.x), fmaxf(ah.y, bh.y), fmaxf(ah.z, bh.z), 0.0f);
    t.left[node] = L; t.right[node] = R; t.cnt[node] = count;
    t.ibox_lo[node] = nl; t.ibox_hi[node] = nh;
    node_costs(t, node, L, R, al, ah, bl, bh, box_area(nl, nh), count);
    p.id[i] = node; p.lo[i] = nl; p.hi[i] = nh;      // only this thread reads slot i's box in this kernel
}

__global__ void k_ploc_compact(Ploc p, uint32_t m) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    if (i == m - 1) *p.m_out = p.offset[i] + p.keep[i];
    if (!p.keep[i]) return;
    const uint32_t d = p.offset[i];
    p.id2[d] = p.id[i]; p.lo2[d] = p.lo[i]; p.hi2[d] = p.hi[i];
}

struct Collapse {
    uint32_t* wide_src;          // wide node index -> radix-tree node it was made from
    uint32_t* nodes;             // 20 words per wide node
    float4*   prims;             // 3 x float4 per leaf primitive
    uint32_t* counters;          // [0] wide nodes allocated, [1] primitives emitted, [2] overflow flag
    double*   sah;               // [0] accumulated SAH cost (relative to the root area)
    uint32_t  node_capacity;
};

// One thread per wide node of the current level: open the largest-area child until 8 slots are
// used (a subtree of <= MAX_LEAF primitives is one leaf slot), place children in octant order,
// quantise, allocate the next level's nodes and this node's primitive records.
__global__ void k_collapse_level(Tree t, Collapse c, const ptb_packed_mesh_face* __restrict__ faces, uint32_t level_begin, uint32_t level_end) {
    const uint32_t w = level_begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= level_end) return;
    const uint32_t src = c.wide_src[w];
    uint32_t item[8]; float4 ilo[8], ihi[8]; double area[8]; uint32_t cnt[8];
    int m = 0;
    float4 slo, shi;
    node_box(t, src, slo, shi);
    auto put = [&](int k, uint32_t id) { item[k] = id; node_box(t, id, ilo[k], ihi[k]); area[k] = box_area(ilo[k], ihi[k]); cnt[k] = node_count(t, id); };
    if (node_count(t, src) <= MAX_LEAF) { put(0, src); m = 1; }
    else if (t.greedy) {
        put(0, t.left[src]); put(1, t.right[src]); m = 2;
        while (m < 8) {
            int best = -1; double ba = -1.0;
            for (int k = 0; k < m; k++) if (cnt[k] > MAX_LEAF && area[k] > ba) { ba = area[k]; best = k; }
            if (best < 0) break;
            const uint32_t id = item[best];
            put(best, t.left[id]); put(m, t.right[id]); m++;
        }
    } else {
        // the optimal 8-slot forest below src
        uint32_t wn[16]; int ws[16]; int nw = 0;
        const int k8 = t.dp_split[8 * (size_t)src + 7];
        wn[nw] = t.right[src]; ws[nw++] = 8 - k8;
        wn[nw] = t.left[src]; ws[nw++] = k8;
        while (nw > 0) {
            const uint32_t id = wn[--nw]; int i = ws[nw];
            if ((id & LEAF) || i == 1) { put(m++, id); continue; }
            const uint8_t* sp = t.dp_split + 8 * (size_t)id;
            while (i > 1 && sp[i - 1] == 0) i--;
            if (i == 1) { put(m++, id); continue; }
            const int k = sp[i - 1];
            wn[nw] = t.right[id]; ws[nw++] = i - k;
            wn[nw] = t.left[id]; ws[nw++] = k;
        }
    }
    // octant slot assignment: greedy on (child centre - node centre) . octant sign
    float dx[8], dy[8], dz[8];
    const float ncx = 0.5f * slo.x + 0.5f * shi.x, ncy = 0.5f * slo.y + 0.5f * shi.y, ncz = 0.5f * slo.z + 0.5f * shi.z;
    for (int k = 0; k < m; k++) {
        dx[k] = (0.5f * ilo[k].x + 0.5f * ihi[k].x) - ncx; dy[k] = (0.5f * ilo[k].y + 0.5f * ihi[k].y) - ncy; dz[k] = (0.5f * ilo[k].z + 0.5f * ihi[k].z) - ncz;
    }
    int child_in_slot[8];
    for (int s = 0; s < 8; s++) child_in_slot[s] = -1;
    uint32_t done = 0, used = 0;
    for (int it = 0; it < m; it++) {
        int bc = -1, bs = -1; float bv = -INFINITY;
        for (int k = 0; k < m; k++) {
            if (done >> k & 1u) continue;
            for (int s = 0; s < 8; s++) {
                if (used >> s & 1u) continue;
                const float v = ((s & 1) ? dx[k] : -dx[k]) + ((s & 2) ? dy[k] : -dy[k]) + ((s & 4) ? dz[k] : -dz[k]);
                if (v > bv || bc < 0) { bv = v; bc = k; bs = s; }
            }
        }
        done |= 1u << bc; used |= 1u << bs; child_in_slot[bs] = bc;
    }
    uint32_t inner = 0, leaf_prims = 0;
    for (int s = 0; s < 8; s++) { const int k = child_in_slot[s]; if (k < 0) continue; if (cnt[k] > MAX_LEAF) inner++; else leaf_prims += cnt[k]; }
    const uint32_t child_base = inner ? atomicAdd(&c.counters[0], inner) : 0u;
    const uint32_t prim_base = leaf_prims ? atomicAdd(&c.counters[1], leaf_prims) : 0u;
    if (inner && child_base + inner > c.node_capacity) { atomicExch(&c.counters[2], 1u); return; }

    uint32_t word[20];
    for (int k = 0; k < 20; k++) word[k] = 0;
    const float plo[3] = {slo.x, slo.y, slo.z}, phi[3] = {shi.x, shi.y, shi.z};
    uint32_t ebyte[3]; double escale[3];
    for (int a = 0; a < 3; a++) {
        ebyte[a] = exponent_byte((double)phi[a] - (double)plo[a]);
        escale[a] = ldexp(1.0, (int)ebyte[a] - 127);
        word[a] = __float_as_uint(plo[a]);
    }
    uint32_t imask = 0, valid24 = 0, next_child = child_base, next_prim = prim_base;
    double sah = box_area(slo, shi);
    for (int s = 0; s < 8; s++) {
        const int k = child_in_slot[s];
        if (k < 0) continue;
        const float clo[3] = {ilo[k].x, ilo[k].y, ilo[k].z}, chi[3] = {ihi[k].x, ihi[k].y, ihi[k].z};
        for (int a = 0; a < 3; a++) {
            double lo = floor(((double)clo[a] - (double)plo[a]) / escale[a]);
            double hi = ceil(((double)chi[a] - (double)plo[a]) / escale[a]);
            lo = fmin(255.0, fmax(0.0, lo)); hi = fmin(255.0, fmax(0.0, hi));
            word[8 + 2 * a + (s >> 2)] |= (uint32_t)lo << (8 * (s & 3));
            word[14 + 2 * a + (s >> 2)] |= (uint32_t)hi << (8 * (s & 3));
        }
        if (cnt[k] > MAX_LEAF) {
            imask |= 1u << s;
            c.wide_src[next_child++] = item[k];
        } else {
            valid24 |= ((1u << cnt[k]) - 1u) << (3 * s);
            static_assert(MAX_LEAF == 2, "a leaf slot is one leaf or a node with two leaf children");
            const uint32_t pos[2] = {((item[k] & LEAF) ? item[k] : t.left[item[k]]) & ~LEAF, ((item[k] & LEAF) ? 0u : t.right[item[k]]) & ~LEAF};
            for (uint32_t q = 0; q < cnt[k]; q++) {
                const uint32_t p = t.sorted[pos[q]];
                const uint32_t shape = __float_as_uint(t.box_lo[p].w), face = __float_as_uint(t.box_hi[p].w), tag = t.tag[p];
                float4 r0 = make_float4(0, 0, 0, 0), r1 = r0, r2 = r0;
                if ((tag & 3u) == 0u) { const float4* f = (const float4*)(faces + face); r0 = f[0]; r1 = f[1]; r2 = f[2]; }
                r0.w = __uint_as_float(face); r1.w = __uint_as_float(shape); r2.w = __uint_as_float(tag);
                float4* out = c.prims + 3 * (size_t)next_prim++;
                out[0] = r0; out[1] = r1; out[2] = r2;
            }
            sah += area[k] * cnt[k];
        }
    }
    word[3] = ebyte[0] | (ebyte[1] << 8) | (ebyte[2] << 16) | (imask << 24);
    word[4] = child_base; word[5] = prim_base; word[6] = valid24;
    uint4* out = (uint4*)(c.nodes + 20 * (size_t)w);
#pragma unroll
    for (int k = 0; k < 5; k++) out[k] = make_uint4(word[4 * k], word[4 * k + 1], word[4 * k + 2], word[4 * k + 3]);
    atomicAdd(c.sah, sah);
}

// Scratch memory of one build: one cudaMalloc sized from the primitive count, handed out by bumping a pointer
// (a build makes ~35 arrays; that many cudaMalloc/cudaFree pairs cost more than the kernels at 1 M triangles).
// Requests that do not fit fall back to their own allocation.  Frees everything it handed out.
struct Scratch {
    char* pool = nullptr; size_t pool_bytes = 0, used = 0;
    std::vector<void*> ptrs;
    cudaError_t reserve(size_t bytes) {
        cudaError_t e = cudaMalloc((void**)&pool, bytes);
        if (e != cudaSuccess) { pool = nullptr; cudaGetLastError(); return cudaSuccess; }     // no pool: individual allocations
        pool_bytes = bytes;
        return cudaSuccess;
    }
    template <class T> cudaError_t get(T** p, size_t count) {
        const size_t bytes = ((count ? count : 1) * sizeof(T) + 255) & ~(size_t)255;
        if (pool && used + bytes <= pool_bytes) { *p = (T*)(pool + used); used += bytes; return cudaSuccess; }
        *p = nullptr;
        cudaError_t e = cudaMalloc((void**)p, bytes);
        if (e == cudaSuccess) ptrs.push_back(*p);
        return e;
    }
    ~Scratch() { for (void* p : ptrs) cudaFree(p); if (pool) cudaFree(pool); }
};

}  // namespace

#define GB(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { *err = std::string(#x) + ": " + cudaGetErrorString(e_); if (out->nodes) cudaFree(out->nodes); if (out->prims) cudaFree(out->prims); out->nodes = nullptr; out->prims = nullptr; return PTB_E_CUDA; } } while (0)

int build_wide_bvh_gpu(const ptb_scene_view* v, const ptb_packed_shape* d_shapes, const ptb_packed_mesh_face* d_faces,
                       cudaStream_t st, int algorithm, GpuWideBuild* out, std::string* err) {
    const auto t0 = std::chrono::steady_clock::now();
    *out = GpuWideBuild();

    // ---- per-shape table (host, O(shapes x mesh depth)) ----
    const uint32_t S = v->ShapeCount;
    std::vector<uint32_t> prim_base(S + 1, 0), face_begin(S, 0), rank_base(S, 0), flags(S, 0), plane_list;
    static const float I16[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    uint64_t prims = 0, rank = 0;
    for (uint32_t s = 0; s < S; s++) {
        const ptb_packed_shape& sh = v->Shapes[s];
        const bool identity = !std::memcmp(sh.Transform.From, I16, 64) && !std::memcmp(sh.Transform.To, I16, 64);
        flags[s] = ((uint32_t)sh.Type & 0xFFu) | (identity ? 0x100u : 0u);
        prim_base[s] = (uint32_t)prims; rank_base[s] = (uint32_t)rank;
        uint32_t count = 0;
        if (sh.Type == 1) { plane_list.push_back(s); plane_list.push_back(((uint32_t)rank << 3) | 1u); rank += 1; }
        else if (sh.Type == 2 || sh.Type == 3) count = 1;
        else if (sh.Type == 0) {
            // the faces of a mesh are one contiguous range: first face of the leftmost leaf .. end of the rightmost
            // (the packer appends each mesh's faces in node order, src/scene/scene.cpp:1265-1330)
            if (sh.MeshRootNodeIndex >= v->MeshNodeCount) { *err = "shape mesh root out of range"; return PTB_E_INVALID; }
            uint32_t a = sh.MeshRootNodeIndex, b = sh.MeshRootNodeIndex;
            for (int guard = 0; v->MeshNodes[a].FaceEndIndex == 0; guard++) {
                a = v->MeshNodes[a].FaceBeginOrNodeIndex;
                if (a >= v->MeshNodeCount || guard > 256) { *err = "mesh node chain out of range"; return PTB_E_INVALID; }
            }
            for (int guard = 0; v->MeshNodes[b].FaceEndIndex == 0; guard++) {
                b = v->MeshNodes[b].FaceBeginOrNodeIndex + 1;
                if (b >= v->MeshNodeCount || guard > 256) { *err = "mesh node chain out of range"; return PTB_E_INVALID; }
            }
            const uint32_t fb = v->MeshNodes[a].FaceBeginOrNodeIndex, fe = v->MeshNodes[b].FaceEndIndex;
            if (fe > v->MeshFaceCount || fb > fe) { *err = "mesh face range out of range"; return PTB_E_INVALID; }
            face_begin[s] = fb; count = fe - fb;
        }
        prims += count; rank += count;
        if (prims + plane_list.size() / 2 >= (1ull << 27)) { *err = "more than 2^27 primitives"; return PTB_E_UNSUPPORTED; }
    }
    prim_base[S] = (uint32_t)prims;
    out->planes = plane_list;
    const uint32_t N = (uint32_t)prims;
    if (N == 0) { out->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); return PTB_OK; }

    Scratch mem;
    // per primitive: boxes 32 + tag 4 + keys 16 + values 8 + tree 92 + collapse 84 (+ PLOC clusters 84) bytes, plus sort/scan scratch
    mem.reserve((size_t)N * (algorithm == 0 ? 240 : 324) + ((size_t)S * 20) + (32u << 20));
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    GB(cudaEventCreate(&ev0)); GB(cudaEventCreate(&ev1));
    struct EvGuard { cudaEvent_t a, b; ~EvGuard() { cudaEventDestroy(a); cudaEventDestroy(b); } } evg{ev0, ev1};
    GB(cudaEventRecord(ev0, st));

    uint32_t *d_prim_base, *d_face_begin, *d_rank_base, *d_flags;
    GB(mem.get(&d_prim_base, S + 1)); GB(mem.get(&d_face_begin, S)); GB(mem.get(&d_rank_base, S)); GB(mem.get(&d_flags, S));
    GB(cudaMemcpyAsync(d_prim_base, prim_base.data(), (S + 1) * 4, cudaMemcpyHostToDevice, st));
    GB(cudaMemcpyAsync(d_face_begin, face_begin.data(), S * 4, cudaMemcpyHostToDevice, st));
    GB(cudaMemcpyAsync(d_rank_base, rank_base.data(), S * 4, cudaMemcpyHostToDevice, st));
    GB(cudaMemcpyAsync(d_flags, flags.data(), S * 4, cudaMemcpyHostToDevice, st));
    ShapeTable T{d_prim_base, d_face_begin, d_rank_base, d_flags, S};

    Tree t;
    std::memset(&t, 0, sizeof t);
    t.n = N;
    unsigned long long *keys_a, *keys_b;
    uint32_t *vals_a, *vals_b, *d_small;
    float* d_bounds;
    GB(mem.get(&t.box_lo, N)); GB(mem.get(&t.box_hi, N)); GB(mem.get(&t.tag, N));
    GB(mem.get(&keys_a, N)); GB(mem.get(&keys_b, N)); GB(mem.get(&vals_a, N)); GB(mem.get(&vals_b, N));
    GB(mem.get(&d_bounds, 8)); GB(mem.get(&d_small, 8));
    const float binit[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
    GB(cudaMemcpyAsync(d_bounds, binit, sizeof binit, cudaMemcpyHostToDevice, st));
    GB(cudaMemsetAsync(d_small, 0, 8 * 4, st));           // [0] invalid primitives, [1] radix-tree height

    const uint32_t B = 256, G = (N + B - 1) / B;
    k_prim_bounds<<<G, B, 0, st>>>(T, d_shapes, d_faces, t, d_bounds, d_small);
    k_morton<<<G, B, 0, st>>>(t, d_bounds, keys_a, vals_a);
    size_t temp_bytes = 0;
    cub::DoubleBuffer<unsigned long long> dk(keys_a, keys_b);
    cub::DoubleBuffer<uint32_t> dv(vals_a, vals_b);
    GB(cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, dk, dv, (int)N, 0, 64, st));
    char* d_temp;
    GB(mem.get(&d_temp, temp_bytes));
    GB(cub::DeviceRadixSort::SortPairs(d_temp, temp_bytes, dk, dv, (int)N, 0, 64, st));
    uint32_t invalid = 0;
    GB(cudaMemcpyAsync(&invalid, d_small, 4, cudaMemcpyDeviceToHost, st));
    GB(cudaStreamSynchronize(st));
    const uint32_t n = N - invalid;                        // non-finite primitives sorted to the end are left out
    out->prim_count = n;
    if (n == 0) { out->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); return PTB_OK; }
    t.n = n; t.keys = dk.Current(); t.sorted = dv.Current();

    const uint32_t ni = n > 1 ? n - 1 : 1;
    GB(mem.get(&t.left, ni)); GB(mem.get(&t.right, ni)); GB(mem.get(&t.cnt, ni));
    GB(mem.get(&t.parent, (size_t)ni + n)); GB(mem.get(&t.ibox_lo, ni)); GB(mem.get(&t.ibox_hi, ni)); GB(mem.get(&t.visits, ni));
    GB(mem.get(&t.dp_cost, (size_t)ni * 7)); GB(mem.get(&t.dp_split, (size_t)ni * 8));
    t.greedy = std::getenv("PTB_LBVH_GREEDY") ? 1u : 0u;
    GB(cudaMemsetAsync(t.visits, 0, (size_t)ni * 4, st));
    uint32_t root = n > 1 ? 0u : LEAF;
    if (n > 1 && algorithm == 0) {
        k_radix_tree<<<(n - 1 + B - 1) / B, B, 0, st>>>(t);
        k_fit_bounds<<<(n + B - 1) / B, B, 0, st>>>(t, d_small + 1);
    } else if (n > 1) {
        Ploc p;
        GB(mem.get(&p.id, n)); GB(mem.get(&p.lo, n)); GB(mem.get(&p.hi, n));
        GB(mem.get(&p.id2, n)); GB(mem.get(&p.lo2, n)); GB(mem.get(&p.hi2, n));
        GB(mem.get(&p.nn, n)); GB(mem.get(&p.keep, n)); GB(mem.get(&p.offset, n));
        p.node_counter = d_small + 2; p.m_out = d_small + 3;
        size_t scan_bytes = 0;
        GB(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, p.keep, p.offset, (int)n, st));
        char* d_scan;
        GB(mem.get(&d_scan, scan_bytes));
        int radius = 64;           // profiles/r01_v9_ploc_radius_sweep.txt: C4 trace 2.59 ms at 16, 2.45 at 64, 2.47 at 128; C2 indifferent
        if (const char* e = std::getenv("PTB_PLOC_RADIUS")) radius = std::max(1, std::min(PLOC_RADIUS_MAX, std::atoi(e)));
        k_ploc_init<<<(n + B - 1) / B, B, 0, st>>>(t, p);
        uint32_t m = n, rounds = 0;
        while (m > 1) {
            const uint32_t g = (m + PLOC_BLOCK - 1) / PLOC_BLOCK;
            k_ploc_nearest<<<g, PLOC_BLOCK, 0, st>>>(p, m, radius);
            k_ploc_merge<<<g, PLOC_BLOCK, 0, st>>>(t, p, m);
            GB(cub::DeviceScan::ExclusiveSum(d_scan, scan_bytes, p.keep, p.offset, (int)m, st));
            k_ploc_compact<<<g, PLOC_BLOCK, 0, st>>>(p, m);
            uint32_t next = 0;
            GB(cudaMemcpyAsync(&next, p.m_out, 4, cudaMemcpyDeviceToHost, st));
            GB(cudaStreamSynchronize(st));
            if (next == 0 || next >= m || ++rounds > 4096) { *err = "PLOC made no progress"; return PTB_E_UNSUPPORTED; }
            if (std::getenv("PTB_BUILD_VERBOSE")) fprintf(stderr, "ploc round %u: %u -> %u\n", rounds, m, next);
            m = next;
            std::swap(p.id, p.id2); std::swap(p.lo, p.lo2); std::swap(p.hi, p.hi2);
        }
        root = n - 2;                       // the last node created
        out->rounds = rounds;
    }

    Collapse c;
    c.node_capacity = n + 16;           // every wide node stands for a distinct radix-tree node
    uint32_t* big_nodes;
    double* d_sah;
    GB(mem.get(&c.wide_src, c.node_capacity)); GB(mem.get(&big_nodes, (size_t)c.node_capacity * 20)); GB(mem.get(&c.counters, 4)); GB(mem.get(&d_sah, 1));
    c.nodes = big_nodes; c.sah = d_sah;
    GB(cudaMalloc((void**)&out->prims, (size_t)n * 48));
    c.prims = (float4*)out->prims;
    const uint32_t cinit[4] = {1u, 0u, 0u, 0u};
    GB(cudaMemcpyAsync(c.wide_src, &root, 4, cudaMemcpyHostToDevice, st));
    GB(cudaMemcpyAsync(c.counters, cinit, sizeof cinit, cudaMemcpyHostToDevice, st));
    GB(cudaMemsetAsync(d_sah, 0, 8, st));
    uint32_t level_begin = 0, level_end = 1, depth = 0, host_counters[4] = {1, 0, 0, 0};
    while (level_begin < level_end) {
        depth++;
        const uint32_t count = level_end - level_begin;
        k_collapse_level<<<(count + 127) / 128, 128, 0, st>>>(t, c, d_faces, level_begin, level_end);
        GB(cudaMemcpyAsync(host_counters, c.counters, sizeof host_counters, cudaMemcpyDeviceToHost, st));
        GB(cudaStreamSynchronize(st));
        if (host_counters[2]) { *err = "wide node buffer overflow"; cudaFree(out->prims); out->prims = nullptr; return PTB_E_UNSUPPORTED; }
        level_begin = level_end; level_end = host_counters[0];
        if (depth > 64) break;
    }
    if (depth > 30) { *err = "wide BVH deeper than the traversal stack"; cudaFree(out->prims); out->prims = nullptr; return PTB_E_UNSUPPORTED; }
    if (host_counters[1] != n) { *err = "collapse emitted a different number of primitives than it was given"; cudaFree(out->prims); out->prims = nullptr; return PTB_E_UNSUPPORTED; }
    out->node_count = host_counters[0];
    out->depth = depth;
    GB(cudaMalloc((void**)&out->nodes, (size_t)out->node_count * 80));
    GB(cudaMemcpyAsync(out->nodes, big_nodes, (size_t)out->node_count * 80, cudaMemcpyDeviceToDevice, st));
    double sah = 0.0, root_area = 1.0;
    float4 rl, rh;
    uint32_t height = 1;
    GB(cudaMemcpyAsync(&sah, d_sah, 8, cudaMemcpyDeviceToHost, st));
    if (n > 1) { GB(cudaMemcpyAsync(&rl, t.ibox_lo + root, 16, cudaMemcpyDeviceToHost, st)); GB(cudaMemcpyAsync(&rh, t.ibox_hi + root, 16, cudaMemcpyDeviceToHost, st)); }
    GB(cudaEventRecord(ev1, st));
    GB(cudaStreamSynchronize(st));
    if (n > 1) {
        const double x = (double)rh.x - rl.x, y = (double)rh.y - rl.y, z = (double)rh.z - rl.z;
        root_area = std::max(2.0 * (x * y + y * z + z * x), 1e-300);
        out->sah_cost = sah / root_area;
        std::memcpy(&height, &rl.w, 4);
    }
    out->bvh2_depth = height;
    cudaEventElapsedTime(&out->device_ms, ev0, ev1);
    out->build_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return PTB_OK;
}

}  // namespace ptb
