// sm_100a kernels of the path-tracing integrator and their launchers.
//
//   k_restart   K3  basic_scatter.glsl main, Restart=1  (clear film, start one path per pixel)
//   k_trace     K1  basic_trace.glsl                    (closest hit per slot)
//   k_scatter   K2  basic_scatter.glsl main, Restart=0  (shade, accumulate, regenerate)
//   k_mega      K1+K2 fused over `rounds` with the path state in registers
//   k_resolve   K4  resolve.glsl
//   k_raycast   K1/K5 on caller rays (parity harness, preview_render.glsl:96-109)
//
// Compiled with --fmad=false (see ptb_vec.cuh).
#include "ptb_kernels.h"
#include <cstdio>
#include <cstdlib>
#include "ptb_resolve.cuh"
#include "ptb_wide.cuh"
#include "ptb_preview.cuh"

namespace ptb {

enum { MODE_BVH2_FAITHFUL = 0, MODE_BVH2_CANONICAL = 1, MODE_WIDE = 2 };

template <int MODE>
__device__ __forceinline__ HitId closest_hit(const DeviceScene& s, const Ray& r) {
    if (MODE == MODE_BVH2_FAITHFUL) return closest_hit_bvh2<false>(s, r);
    if (MODE == MODE_BVH2_CANONICAL) return closest_hit_bvh2<true>(s, r);
    return closest_hit_wide(s, r);
}

static int mode_of(unsigned traversal, unsigned tie) {
    if (traversal == PTB_TRAVERSAL_WIDE) return MODE_WIDE;
    return tie == PTB_TIE_CANONICAL ? MODE_BVH2_CANONICAL : MODE_BVH2_FAITHFUL;
}

// slot -> pixel (tile-major 16x16, basic_scatter.glsl:324-328)
__device__ __forceinline__ bool slot_pixel(const RenderConsts& pc, u32 slot, u32& x, u32& y) {
    const u32 tile = slot >> 8, l = slot & 255u;
    u32 tx, ty;
    udivmod(tile, pc.tiles_x, pc.tiles_x_magic, ty, tx);
    x = tx * 16u + (l & 15u);
    y = ty * 16u + (l >> 4);
    return x < pc.width && y < pc.height;        // slots beyond the last tile row have y >= height
}

// ---- state records <-> register structs ----
__device__ __forceinline__ Ray load_ray(const WavefrontState& st, u32 slot) {
    const float4 r = st.ray[slot];
    Ray ray;
    ray.origin = mk3(r.x, r.y, r.z);
    ray.velocity = unpack_unit_vector(__float_as_uint(r.w));
    ray.duration = PTB_HIT_TIME_LIMIT;
    return ray;
}
__device__ __forceinline__ void store_ray(const WavefrontState& st, u32 slot, const Ray& ray) {
    st.ray[slot] = make_float4(ray.origin.x, ray.origin.y, ray.origin.z, __uint_as_float(pack_unit_vector(ray.velocity)));
}
__device__ __forceinline__ PathState load_path(const WavefrontState& st, u32 slot) {
    const uint4 a = st.path0[slot];
    const float4 t = st.path1[slot], p = st.path2[slot], c = st.path3[slot];
    PathState path;
    path.pixel = a.x; path.lambda0 = __uint_as_float(a.y);
    unpack_active(a.z, a.w, path.active);
    path.throughput = mk4(t.x, t.y, t.z, t.w);
    path.probability = mk4(p.x, p.y, p.z, p.w);
    path.sample = mk3(c.x, c.y, c.z);
    path.length = __float_as_uint(c.w);
    return path;
}
__device__ __forceinline__ void store_path(const WavefrontState& st, u32 slot, const PathState& path) {
    u32 w01, w23;
    pack_active(path.active, w01, w23);
    st.path0[slot] = make_uint4(path.pixel, __float_as_uint(path.lambda0), w01, w23);
    st.path1[slot] = make_float4(path.throughput.x, path.throughput.y, path.throughput.z, path.throughput.w);
    st.path2[slot] = make_float4(path.probability.x, path.probability.y, path.probability.z, path.probability.w);
    st.path3[slot] = make_float4(path.sample.x, path.sample.y, path.sample.z, __uint_as_float(path.length));
}

// The same conversions on records already in registers (the staged scatter kernel reads them from shared memory).
__device__ __forceinline__ Ray ray_of(const float4 r) {
    Ray ray;
    ray.origin = mk3(r.x, r.y, r.z);
    ray.velocity = unpack_unit_vector(__float_as_uint(r.w));
    ray.duration = PTB_HIT_TIME_LIMIT;
    return ray;
}
__device__ __forceinline__ float4 record_of(const Ray& ray) {
    return make_float4(ray.origin.x, ray.origin.y, ray.origin.z, __uint_as_float(pack_unit_vector(ray.velocity)));
}
__device__ __forceinline__ PathState path_of(const uint4 a, const float4 t, const float4 p, const float4 c) {
    PathState path;
    path.pixel = a.x; path.lambda0 = __uint_as_float(a.y);
    unpack_active(a.z, a.w, path.active);
    path.throughput = mk4(t.x, t.y, t.z, t.w);
    path.probability = mk4(p.x, p.y, p.z, p.w);
    path.sample = mk3(c.x, c.y, c.z);
    path.length = __float_as_uint(c.w);
    return path;
}

// Compact closest-hit record written by the trace kernels.
__device__ __forceinline__ void store_hit(const DeviceScene& s, const WavefrontState& st, u32 slot, const HitId& h) {
    if (h.shape == PTB_SHAPE_NONE) {
        st.hit0[slot] = make_uint4(__float_as_uint(PTB_HIT_TIME_LIMIT), 0xFFFFFFFFu, 0xFFFFFFFFu, 0u);
        return;
    }
    const u32 material = __ldg(&s.shapes[h.shape].MaterialIndex);
    st.hit0[slot] = make_uint4(__float_as_uint(h.t), (h.shape << 16) | material, h.prim, h.shape_type);   // 16-bit ids: basic.glsl.inc:150
    st.hit1[slot] = make_uint4(__float_as_uint(h.coords.x), __float_as_uint(h.coords.y), __float_as_uint(h.coords.z), 0u);
}
// The wavefront hit as scatter sees it (basic.glsl.inc:99-131): attributes are derived here and
// pass through the reference's 24-byte quantisation (oct-encoded normal/tangent).
__device__ __forceinline__ PackedHit load_packed_hit(const DeviceScene& s, const WavefrontState& st, u32 slot) {
    const uint4 h0 = st.hit0[slot];
    if (h0.y == 0xFFFFFFFFu) return packed_miss();
    const uint4 h1 = st.hit1[slot];
    HitId h;
    h.t = __uint_as_float(h0.x); h.shape = h0.y >> 16; h.prim = h0.z; h.shape_type = h0.w;
    h.coords = mk3(__uint_as_float(h1.x), __uint_as_float(h1.y), __uint_as_float(h1.z));
    const Surface sf = finish_hit(s, h);
    return pack_hit(h, sf);
}

// ---- per-face vertex attributes, gathered once per mesh upload ----
// finish_hit needs the packed normal and uv of a hit face's three vertices.  Through the reference's arrays that is a
// dependent chain - face record (48 B, for three indices) -> three 8-byte vertex records anywhere in the vertex array - i.e.
// five 32-byte sectors and two memory round trips per hit; on the 10 M-triangle scenes, whose 520 MB of faces and vertices
// do not fit in L2, that is most of what the scatter kernels wait for.  One 32-byte record per face, indexed by the face
// index the hit already carries, makes it one sector and one round trip.  Same bits, so results are unchanged.
__global__ void __launch_bounds__(256) k_face_attrs(const ptb_packed_mesh_face* __restrict__ faces, const ptb_packed_mesh_vertex* __restrict__ vertices,
                                                     uint4* __restrict__ out, u32 n) {
    const u32 i = blockIdx.x * 256u + threadIdx.x;
    if (i >= n) return;
    const ptb_packed_mesh_face* f = &faces[i];
    const uint2 v0 = *(const uint2*)&vertices[f->VertexIndex0], v1 = *(const uint2*)&vertices[f->VertexIndex1], v2 = *(const uint2*)&vertices[f->VertexIndex2];
    out[2 * (size_t)i] = make_uint4(v0.x, v0.y, v1.x, v1.y);
    out[2 * (size_t)i + 1] = make_uint4(v2.x, v2.y, 0u, 0u);
}
cudaError_t launch_face_attrs(const ptb_packed_mesh_face* faces, const ptb_packed_mesh_vertex* vertices, uint4* out, unsigned n, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    k_face_attrs<<<(n + 255u) / 256u, 256, 0, stream>>>(faces, vertices, out, n);
    return cudaGetLastError();
}

// ---- K3: restart ----
__global__ void __launch_bounds__(256) k_restart(const DeviceScene s, const RenderConsts pc, const WavefrontState st, float4* __restrict__ accum) {
    const u32 slot = blockIdx.x * 256u + threadIdx.x;
    u32 x, y;
    if (!slot_pixel(pc, slot, x, y)) return;
    Rng g;
    g.state = y * 65537u + x + pc.random_seed * 277803737u;
    PathState path; Ray ray;
    generate_new_path(s, pc, g, x, y, path, ray);
    store_ray(st, slot, ray);
    store_path(st, slot, path);
    accum[y * pc.width + x] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (threadIdx.x == 0) st.tile_paths[blockIdx.x] = 0u;
    if (slot < 4) st.ray_counter[slot] = 0u;        // ray fetch [0,1] and scatter tile fetch [2,3] counters, one of each per pipeline
}

// ---- K1: trace ----
template <int MODE>
__global__ void __launch_bounds__(128) k_trace(const DeviceScene s, const RenderConsts pc, const WavefrontState st) {
    const u32 slot = pc.slot_base + blockIdx.x * 128u + threadIdx.x;
    if (slot == pc.slot_base) st.ray_counter[2] = 0u;       // tile fetch counter of the staged scatter kernel that follows
    u32 x, y;
    if (!slot_pixel(pc, slot, x, y)) return;
    const Ray ray = load_ray(st, slot);
    const HitId h = closest_hit<MODE>(s, ray);
    store_hit(s, st, slot, h);
}


// ---- K1, persistent warp-cooperative variant -----------------------------------------------
// B200 has no RT cores: closest-hit queries are software traversal on the SMs, and their cost is
// dominated by SIMT divergence (rays of a warp finish after very different numbers of steps;
// v1 above runs at ~8 of 32 lanes, profiles/r01_v1_k_trace_wide.txt).  Here the grid is sized
// to the machine (resident warps only) and every warp loops
//     refill  : lanes whose ray is done fetch new rays from a global counter (ballot + prefix rank),
//     node    : every active lane visits one node (all active lanes have node work here),
//     prims   : lanes with leaf primitives test them (vote-driven loop),
//     pop     : lanes out of node work pop their stack or retire and store their hit.
// The first entries of each lane's stack live in shared memory, lane-interleaved (bank = lane, so
// any mix of stack depths is conflict-free); deeper entries spill to local memory.
#ifndef PTB_SMEM_STACK
#define PTB_SMEM_STACK 10
#endif
#define PTB_TRACE_WARPS 4
#ifndef PTB_TRACE_MIN_BLOCKS
#define PTB_TRACE_MIN_BLOCKS 6
#endif

struct WarpStack {
    uint2* smem;                                   // entry k of this lane at smem[k * 32]
    uint2 spill[PTB_WIDE_STACK - PTB_SMEM_STACK];
    int sp;
    __device__ __forceinline__ void push(uint2 v) {
        if (sp < PTB_SMEM_STACK) smem[sp * 32] = v;
        else if (sp < PTB_WIDE_STACK) spill[sp - PTB_SMEM_STACK] = v;
        else return;
        sp++;
    }
    __device__ __forceinline__ uint2 pop() {
        --sp;
        return sp < PTB_SMEM_STACK ? smem[sp * 32] : spill[sp - PTB_SMEM_STACK];
    }
    __device__ __forceinline__ bool empty() const { return sp == 0; }
};

// Ray source / hit sink of the integrator: the wavefront state, slot order.
struct WavefrontJob {
    DeviceScene s; RenderConsts pc; WavefrontState st;
    __device__ __forceinline__ u32 total() const { return pc.slot_count; }
    __device__ __forceinline__ unsigned* counter() const { return st.ray_counter; }
    __device__ __forceinline__ bool load(u32 i, Ray& r) const { u32 x, y; if (!slot_pixel(pc, pc.slot_base + i, x, y)) return false; r = load_ray(st, pc.slot_base + i); return true; }
    __device__ __forceinline__ void store(u32 i, const HitId& h) const { store_hit(s, st, pc.slot_base + i, h); }
    __device__ __forceinline__ void begin() const { if (blockIdx.x == 0 && threadIdx.x == 0) st.ray_counter[2] = 0u; }   // tile fetch counter of the staged scatter kernel that follows
};
// Caller rays (parity harness).
struct RaycastJob {
    DeviceScene s; const ptb_ray* rays; ptb_hit* out; u32 n; unsigned* ctr;
    __device__ __forceinline__ u32 total() const { return n; }
    __device__ __forceinline__ unsigned* counter() const { return ctr; }
    __device__ __forceinline__ bool load(u32 i, Ray& r) const {
        const ptb_ray in = rays[i];
        r.origin = mk3(in.ox, in.oy, in.oz); r.velocity = mk3(in.dx, in.dy, in.dz); r.duration = in.duration;
        return true;
    }
    __device__ __forceinline__ void store(u32 i, const HitId& h) const;
    __device__ __forceinline__ void begin() const {}
};

template <class Job, int REFILL>
__global__ void __launch_bounds__(32 * PTB_TRACE_WARPS, PTB_TRACE_MIN_BLOCKS) k_trace_persistent(const Job job) {
    __shared__ uint2 stack_mem[PTB_TRACE_WARPS][PTB_SMEM_STACK][32];
    job.begin();
    const DeviceScene& s = job.s;
    const u32 lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const u32 lt_mask = (1u << lane) - 1u;
    const u32 total = job.total();

    bool active = false, exhausted = false;
    u32 index = 0, cur_shape = PTB_SHAPE_NONE;
    Ray ray, obj;
    WideCtx c;
    HitId h;
    uint2 ng = make_uint2(0u, 0u);
    WarpStack stack;
    stack.smem = &stack_mem[warp][0][lane];
    stack.sp = 0;
    ray.origin = ray.velocity = mk3(0, 0, 0); ray.duration = 0; obj = ray; c = wide_ctx(ray);
    h.t = 0; h.shape = PTB_SHAPE_NONE; h.prim = 0; h.shape_type = 0; h.coords = mk3(0, 0, 0);

    for (;;) {
        // ---- refill ----
        const u32 idle = __ballot_sync(0xFFFFFFFFu, !active);
        if (!exhausted && (__popc(idle) >= REFILL || idle == 0xFFFFFFFFu)) {
            const u32 want = __popc(idle);
            u32 base = 0;
            if (lane == 0) base = atomicAdd(job.counter(), want);
            base = __shfl_sync(0xFFFFFFFFu, base, 0);
            if (base + want >= total) exhausted = true;
            if (!active) {
                index = base + __popc(idle & lt_mask);
                if (index < total && job.load(index, ray)) {
                    h = wide_begin(s, ray);
                    if (s.wide_tlas_root == 0xFFFFFFFFu) {
                        job.store(index, h);
                    } else {
                        c = wide_ctx(ray);
                        ng = make_uint2(0u, 0x80000000u);
                        stack.sp = 0;
                        cur_shape = PTB_SHAPE_NONE;
                        active = true;
                    }
                }
            }
        }
        if (!__any_sync(0xFFFFFFFFu, active)) {
            if (exhausted) break;
            continue;
        }

        // ---- node: every active lane has node work here ----
        PrimGroup pg;
        pg.base = pg.bits = pg.valid = 0u;
        if (active) wide_node_step(s, c, h.t, ng, pg, stack);

        // ---- leaf primitives ----
        while (__any_sync(0xFFFFFFFFu, pg.bits != 0u)) {
            if (pg.bits != 0u) wide_prim_step(s, ray, pg, cur_shape, obj, h);
        }

        // ---- pop / retire ----
        if (active && ng.y <= 0x00FFFFFFu) {
            if (stack.empty()) { job.store(index, h); active = false; }
            else ng = stack.pop();
        }
    }
}

// ---- K1, cooperative leaf tests ------------------------------------------------------------
// In k_trace_persistent the leaf-primitive loop runs at ~5 of 32 lanes and takes half of the kernel's
// time (profiles/r01_v2_k_trace_persistent.txt): per node step only a few lanes have primitives to
// test, and the rest of the warp waits for them.  Here primitive tests are decoupled from the ray that
// asked for them.  After a node step a lane appends its hit leaf primitives as ONE group record
// {owner lane, primitive base, hit bits, valid bits} to a per-warp queue in shared memory and goes on;
// whenever the queue holds 32 primitives the whole warp tests one each: the group sizes are scanned,
// lane i finds the group that owns the i-th primitive (bit mask of group starts + find-leading-one),
// reads the owner's ray and current reach from shared memory, runs the exact reference test, and
// publishes an accepted hit with a 64-bit atomicMin on the owner's key {t bits, canonical rank}.  That
// key order IS the canonical tie rule (lowest t, then shape, then face: the builder ranks primitives
// in that order), so the result does not depend on which lane tested what, or when.
//   SPEC    a lane keeps traversing while its primitives wait in the queue (its reach is stale until
//           the flush: more node steps, full lanes) or stalls until they have been tested;
//   WAIT_T  a partial flush happens when this many lanes cannot go on without one.
// Refill: when REFILL lanes are idle and the warp's pool is empty, the whole warp fetches 32 consecutive rays from
// the global counter, decodes them, tests the unbounded shapes and parks the prepared records in shared memory;
// idle lanes then pick their next ray up with a few shared-memory loads (refill work at 32 lanes instead of ~10).
// Measured and rejected (profiles/r01_v3_ab_builds.txt): prefetch.global.L1 of the next node and of the queued
// primitives right after a node step (+13 % time: extra address arithmetic, L1 churn); 64 registers / 8 blocks per
// SM; fetch chunks of 64 / 128 rays (tail imbalance, +8 % at 128); flushing on a group count instead of the exact
// primitive count; __fns for the k-th pending bit.
#ifndef PTB_COOP_FLUSH_PRIMS
#define PTB_COOP_FLUSH_PRIMS 24         // 24: measured best (16..32 within 2 %, profiles/r01_v3_ab_builds.txt)
#endif
#define PTB_COOP_QUEUE 64           // groups; >= 31 (left by a flush) + 32 (one per lane and round), power of two
#ifndef PTB_COOP_MIN_BLOCKS
#define PTB_COOP_MIN_BLOCKS 6
#endif
// 1: a finished lane keeps its hit in shared memory and stores it when it picks up its next ray (>= REFILL lanes act together
// there; retiring lanes are ~3.5 per iteration).  Together with PTB_COOP_ATOMIC32: C2 trace 0.605 -> 0.598 ms, whole job +1.3 %
// (profiles/r02_f_ab_defer_store_atomic32.txt; each alone: -0.8 % / +0.3 %).
#ifndef PTB_COOP_DEFER_STORE
#define PTB_COOP_DEFER_STORE 1
#endif
// 1: the closest-hit key {t bits, canonical rank} is kept as two 32-bit words and published with two native shared-memory
// atomicMin (ATOMS.MIN) instead of one 64-bit atomicMin, which compiles to a compare-and-swap spin loop (ATOMS.CAST.SPIN.64).
#ifndef PTB_COOP_ATOMIC32
#define PTB_COOP_ATOMIC32 1
#endif
#if PTB_COOP_ATOMIC32
#define PTB_COOP_REACH(W, l) u2f((W).tkey[l])
#else
#define PTB_COOP_REACH(W, l) u2f((u32)((W).key[l] >> 32))
#endif

struct CoopWarp {
    uint2 stack[PTB_SMEM_STACK][32];
#if PTB_COOP_ATOMIC32
    u32 tkey[32], rkey[32];           // closest hit so far: t bits (+0 canonical) and tag; tag 0xFFFFFFFF = none
#else
    unsigned long long key[32];       // closest hit so far: {t bits (+0 canonical), tag}; tag 0xFFFFFFFF = none
#endif
    float ray[6][32];                 // world-space origin, velocity of each lane's ray
    u32 detail[7][32];                // t, shape, face, shape type, coords of the hit `key` names
    uint4 group[PTB_COOP_QUEUE];      // {owner lane << 27 | primitive base, pending bits, valid bits, -}
    u32 pool[15][32];                 // the next rays, prepared by the whole warp: origin, velocity, key, slot, plane hit
    u32 starts[32];                   // flush scratch: queue position (relative to the head) of the group starting at primitive i
};

// Tests the first min(32, queued) primitives of the group queue, one per lane.  ghead is updated.
template <bool STATS = false>
__device__ __forceinline__ void coop_flush(const DeviceScene& s, CoopWarp& W, u32 lane, u32& ghead, u32 gtail, u32& qprims, u32* tests = nullptr) {
    __syncwarp();
    const u32 groups = min(gtail - ghead, 32u);
    uint4 g = make_uint4(0u, 0u, 0u, 0u);
    if (lane < groups) g = W.group[(ghead + lane) & (PTB_COOP_QUEUE - 1)];
    const u32 n = (u32)__popc(g.y);
    u32 incl = n;
#pragma unroll
    for (int d = 1; d < 32; d *= 2) { const u32 t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= (u32)d) incl += t; }
    const u32 excl = incl - n;
    const bool starts_here = n != 0u && excl < 32u;
    const u32 start_mask = __reduce_or_sync(0xFFFFFFFFu, starts_here ? 1u << excl : 0u);
    if (starts_here) W.starts[excl] = lane;
    const u32 take = min(__shfl_sync(0xFFFFFFFFu, incl, 31), 32u);      // primitives held by the first 32 groups
    __syncwarp();

    bool won = false;
    u32 owner = 0;
#if PTB_COOP_ATOMIC32
    u32 tb = 0, t_seen = 0, rank_tag = 0;
#else
    unsigned long long key = 0;
#endif
    HitId h;
    h.t = 0; h.shape = PTB_SHAPE_NONE; h.prim = 0xFFFFFFFFu; h.shape_type = 0; h.coords = mk3(0, 0, 0);
    if (lane < take) {
        if (STATS) { tests[0]++; if (lane == 0) tests[1]++; }
        const u32 start = highest_bit(start_mask & (0xFFFFFFFFu >> (31u - lane)));
        const uint4 gg = W.group[(ghead + W.starts[start]) & (PTB_COOP_QUEUE - 1)];
        u32 bits = gg.y;
        for (u32 k = lane - start; k != 0u; k--) bits &= bits - 1u;          // this lane takes the k-th lowest pending bit
        const u32 b = (u32)__ffs((int)bits) - 1u;
        owner = gg.x >> 27;
        const float4* tp = s.wide_tris + 3 * (size_t)((gg.x & 0x07FFFFFFu) + (u32)__popc(gg.z & ((1u << b) - 1u)));
        const float4 v0 = PTB_LDG(&tp[0]), v1 = PTB_LDG(&tp[1]), v2 = PTB_LDG(&tp[2]);
        const u32 face = f2u(v0.w), shape = f2u(v1.w), tag = f2u(v2.w);
        Ray w;
        w.origin = mk3(W.ray[0][owner], W.ray[1][owner], W.ray[2][owner]);
        w.velocity = mk3(W.ray[3][owner], W.ray[4][owner], W.ray[5][owner]);
        h.t = w.duration = PTB_COOP_REACH(W, owner);
        Ray o;
        if (tag & 4u) {
            // identity transform: ((1*x + 0*y) + 0*z) + 0 of mat_mul_point / mat_mul_vector is x + 0.0f for every
            // finite ray (the zero products only matter for x = -0, which ends as +0 either way): no matrix fetch
            o.origin = mk3(w.origin.x + 0.0f, w.origin.y + 0.0f, w.origin.z + 0.0f);
            o.velocity = mk3(w.velocity.x + 0.0f, w.velocity.y + 0.0f, w.velocity.z + 0.0f);
            o.duration = w.duration;
        } else {
            o = object_ray(s, w, shape);
        }
        if ((tag & 3u) == 0u) test_triangle<true>(o, mk3(v0.x, v0.y, v0.z), mk3(v1.x, v1.y, v1.z), mk3(v2.x, v2.y, v2.z), shape, face, h);
        else test_analytic<true>(o, tag & 3u, shape, h);
#if PTB_COOP_ATOMIC32
        if (h.shape != PTB_SHAPE_NONE) {
            tb = f2u(h.t);
            if (tb == 0x80000000u) tb = 0u;         // -0 == +0 in the reference's comparisons
            t_seen = f2u(w.duration);
            rank_tag = tag;
            atomicMin(&W.tkey[owner], tb);
            won = true;
        }
    }
    __syncwarp();
    // the lowest t is known; candidates at that t settle the tie by rank.  When t improved, the rank of the old hit goes first:
    // the first lane that wrote this t read a larger one before (t_seen > best), so at least one candidate sees the improvement.
    const u32 best = won ? W.tkey[owner] : 0u;
    won = won && tb == best;
    if (won && best < t_seen) W.rkey[owner] = 0xFFFFFFFFu;
    __syncwarp();
    if (won) atomicMin(&W.rkey[owner], rank_tag);
    __syncwarp();
    if (won && W.rkey[owner] == rank_tag) {
#else
        if (h.shape != PTB_SHAPE_NONE) {
            u32 tb = f2u(h.t);
            if (tb == 0x80000000u) tb = 0u;         // -0 == +0 in the reference's comparisons
            key = ((unsigned long long)tb << 32) | tag;
            atomicMin(&W.key[owner], key);
            won = true;
        }
    }
    __syncwarp();
    if (won && W.key[owner] == key) {
#endif
        W.detail[0][owner] = f2u(h.t); W.detail[1][owner] = h.shape; W.detail[2][owner] = h.prim; W.detail[3][owner] = h.shape_type;
        W.detail[4][owner] = f2u(h.coords.x); W.detail[5][owner] = f2u(h.coords.y); W.detail[6][owner] = f2u(h.coords.z);
    }
    // consumed groups leave the queue (they are a prefix of it); at most one group is consumed in part
    const u32 used = excl >= take ? 0u : min(n, take - excl);
    if (used != 0u && used != n) {
        u32 bits = g.y;
        for (u32 k = used; k != 0u; k--) bits &= bits - 1u;
        W.group[(ghead + lane) & (PTB_COOP_QUEUE - 1)].y = bits;
    }
    ghead += (u32)__popc(__ballot_sync(0xFFFFFFFFu, n != 0u && used == n));
    qprims -= min(qprims, take);
    __syncwarp();
}

// STATS: the same kernel with per-lane work counters, summed into job-independent device memory at exit (one extra,
// untimed launch from ptb_renderer_trace_stats: what bench.py's own-layout byte model and lane occupancy are measured with).
//   [0] node steps (80-byte BVH8 nodes fetched + tested)   [1] leaf primitive tests (48-byte records)
//   [2] warp loop iterations   [3] flushes (warp-wide primitive test rounds)   [4] rays retired
template <class Job, int REFILL, bool SPEC, int WAIT_T, bool STATS = false>
__global__ void __launch_bounds__(32 * PTB_TRACE_WARPS, PTB_COOP_MIN_BLOCKS) k_trace_coop(const Job job, unsigned long long* __restrict__ stats = nullptr) {
    __shared__ CoopWarp warp_mem[PTB_TRACE_WARPS];
    job.begin();
    const DeviceScene& s = job.s;
    const u32 lane = threadIdx.x & 31u;
    CoopWarp& W = warp_mem[threadIdx.x >> 5];
    const u32 lt_mask = (1u << lane) - 1u;
    const u32 total = job.total();

    bool active = false, drained = false;      // drained: the global counter has passed `total`
    u32 pool_next = 0, pool_end = 0;           // this warp's chunk of ray indices (warp-uniform)
    u32 index = 0;
    u32 ghead = 0, gtail = 0, qprims = 0;      // group queue sequence numbers and queued primitive count (warp-uniform)
    u32 last_seq = 0;                          // one past this lane's newest group: its tests are pending while last_seq > ghead
#if PTB_COOP_DEFER_STORE
    bool unsaved = false;                      // this lane's finished ray still has its hit in W.detail: stored when the lane picks up its next ray
#endif
    WideCtx c;
    uint2 ng = make_uint2(0u, 0u);
    WarpStack stack;
    stack.smem = &W.stack[0][lane];
    stack.sp = 0;
    { Ray z; z.origin = z.velocity = mk3(0, 0, 0); z.duration = 0; c = wide_ctx(z); }
    u32 n_nodes = 0, n_iter = 0, n_retired = 0, n_tests[2] = {0u, 0u};

    for (;;) {
        if (STATS && lane == 0) n_iter++;
        // ---- refill: whole-warp preparation of the next 32 rays into a shared-memory pool, cheap pick-up by idle lanes ----
        const u32 idle = __ballot_sync(0xFFFFFFFFu, !active);
        if (__popc(idle) >= REFILL || idle == 0xFFFFFFFFu) {
            if (pool_next == pool_end && !drained) {
                u32 base = 0;
                if (lane == 0) base = atomicAdd(job.counter(), 32u);
                base = __shfl_sync(0xFFFFFFFFu, base, 0);
                if (base + 32u >= total) drained = true;
                // every lane loads one ray, tests the unbounded shapes (planes) and publishes the prepared record
                const u32 my = base + lane;
                Ray ray;
                ray.origin = ray.velocity = mk3(0, 0, 0); ray.duration = 0;
                bool valid = my < total && job.load(my, ray);
                HitId h;
                h.t = ray.duration; h.shape = PTB_SHAPE_NONE; h.prim = 0xFFFFFFFFu; h.shape_type = 0; h.coords = mk3(0, 0, 0);
                u32 tag = 0xFFFFFFFFu;
                if (valid) {
                    for (u32 i = 0; i < s.wide_plane_count; i++) {
                        const uint2 pl = PTB_LDG((const uint2*)s.wide_planes + i);
                        const u32 before = h.shape;
                        test_analytic<true>(object_ray(s, ray, pl.x), PTB_SHAPE_PLANE, pl.x, h);
                        if (h.shape != before) tag = pl.y;
                    }
                    if (s.wide_tlas_root == 0xFFFFFFFFu) { job.store(my, h); valid = false; }
                }
                const u32 vmask = __ballot_sync(0xFFFFFFFFu, valid);
                __syncwarp();                                   // the previous pool has been consumed by all lanes
                if (valid) {
                    const u32 e = (u32)__popc(vmask & lt_mask);
                    u32 tb = f2u(h.t);
                    if (tb == 0x80000000u) tb = 0u;
                    W.pool[0][e] = f2u(ray.origin.x); W.pool[1][e] = f2u(ray.origin.y); W.pool[2][e] = f2u(ray.origin.z);
                    W.pool[3][e] = f2u(ray.velocity.x); W.pool[4][e] = f2u(ray.velocity.y); W.pool[5][e] = f2u(ray.velocity.z);
                    W.pool[6][e] = tag; W.pool[7][e] = tb; W.pool[8][e] = my;
                    W.pool[9][e] = f2u(h.t); W.pool[10][e] = h.shape; W.pool[11][e] = h.shape_type;
                    W.pool[12][e] = f2u(h.coords.x); W.pool[13][e] = f2u(h.coords.y); W.pool[14][e] = f2u(h.coords.z);
                }
                pool_next = 0; pool_end = (u32)__popc(vmask);
                __syncwarp();
            }
            const u32 avail = pool_end - pool_next;
            const u32 rank = (u32)__popc(idle & lt_mask);
            if (!active && rank < avail) {
                const u32 e = pool_next + rank;
#if PTB_COOP_DEFER_STORE
                if (unsaved) {
                    HitId h;
                    h.t = u2f(W.detail[0][lane]); h.shape = W.detail[1][lane]; h.prim = W.detail[2][lane]; h.shape_type = W.detail[3][lane];
                    h.coords = mk3(u2f(W.detail[4][lane]), u2f(W.detail[5][lane]), u2f(W.detail[6][lane]));
                    job.store(index, h);
                    unsaved = false;
                }
#endif
                Ray ray;
                ray.origin = mk3(u2f(W.pool[0][e]), u2f(W.pool[1][e]), u2f(W.pool[2][e]));
                ray.velocity = mk3(u2f(W.pool[3][e]), u2f(W.pool[4][e]), u2f(W.pool[5][e]));
                ray.duration = 0;
                index = W.pool[8][e];
                c = wide_ctx(ray);
                ng = make_uint2(0u, 0x80000000u);
                stack.sp = 0;
                last_seq = ghead;
                active = true;
                W.ray[0][lane] = ray.origin.x; W.ray[1][lane] = ray.origin.y; W.ray[2][lane] = ray.origin.z;
                W.ray[3][lane] = ray.velocity.x; W.ray[4][lane] = ray.velocity.y; W.ray[5][lane] = ray.velocity.z;
#if PTB_COOP_ATOMIC32
                W.tkey[lane] = W.pool[7][e]; W.rkey[lane] = W.pool[6][e];
#else
                W.key[lane] = ((unsigned long long)W.pool[7][e] << 32) | W.pool[6][e];
#endif
                const u32 hs = W.pool[10][e];
                W.detail[0][lane] = W.pool[9][e]; W.detail[1][lane] = hs; W.detail[2][lane] = hs == PTB_SHAPE_NONE ? 0xFFFFFFFFu : 0u;
                W.detail[3][lane] = W.pool[11][e];
                W.detail[4][lane] = W.pool[12][e]; W.detail[5][lane] = W.pool[13][e]; W.detail[6][lane] = W.pool[14][e];
            }
            pool_next += min((u32)__popc(idle), avail);
            __syncwarp();
            if (!__any_sync(0xFFFFFFFFu, active)) {          // only possible right after a refill attempt
                if (drained && pool_next == pool_end) break;
                continue;
            }
        }

        // ---- node ----
        PrimGroup pg;
        pg.base = pg.bits = pg.valid = 0u;
        if (active && ng.y > 0x00FFFFFFu && (SPEC || (int)(last_seq - ghead) <= 0)) {
            const float reach = PTB_COOP_REACH(W, lane);
            if (STATS) n_nodes++;
            wide_node_step(s, c, reach, ng, pg, stack);
            if (ng.y <= 0x00FFFFFFu && !stack.empty()) ng = stack.pop();
        }

        // ---- queue the hit leaf primitives, one group per lane; 32 queued primitives are tested at once ----
        const u32 has = __ballot_sync(0xFFFFFFFFu, pg.bits != 0u);
        if (has != 0u) {
            if (pg.bits != 0u) {
                const u32 q = gtail + (u32)__popc(has & lt_mask);
                W.group[q & (PTB_COOP_QUEUE - 1)] = make_uint4((lane << 27) | pg.base, pg.bits, pg.valid, 0u);
                last_seq = q + 1u;
            }
            gtail += (u32)__popc(has);
            qprims += __reduce_add_sync(0xFFFFFFFFu, (u32)__popc(pg.bits));
            while (qprims >= (u32)PTB_COOP_FLUSH_PRIMS) coop_flush<STATS>(s, W, lane, ghead, gtail, qprims, n_tests);
        }

        // ---- partial flush when too many lanes wait for their tests ----
        if (gtail != ghead) {
            const bool pending = (int)(last_seq - ghead) > 0;
            const bool stepper = active && ng.y > 0x00FFFFFFu && (SPEC || !pending);
            const bool waiting = active && pending && !stepper;
            const u32 votes = __reduce_add_sync(0xFFFFFFFFu, (stepper ? 1u : 0u) | (waiting ? 0x100u : 0u));      // both counts, one reduction
            if ((votes >> 8) >= (u32)WAIT_T || (votes & 0xFFu) == 0u) coop_flush<STATS>(s, W, lane, ghead, gtail, qprims, n_tests);
        }

        // ---- retire ----
        if (active && ng.y <= 0x00FFFFFFu && (int)(last_seq - ghead) <= 0) {
#if PTB_COOP_DEFER_STORE
            unsaved = true;                    // retiring lanes are few per iteration (3.5 of 32): the stores wait for the pick-up, where >= REFILL lanes act together
#else
            HitId h;
            h.t = u2f(W.detail[0][lane]); h.shape = W.detail[1][lane]; h.prim = W.detail[2][lane]; h.shape_type = W.detail[3][lane];
            h.coords = mk3(u2f(W.detail[4][lane]), u2f(W.detail[5][lane]), u2f(W.detail[6][lane]));
            job.store(index, h);
#endif
            active = false;
            if (STATS) n_retired++;
        }
    }
#if PTB_COOP_DEFER_STORE
    if (unsaved) {
        HitId h;
        h.t = u2f(W.detail[0][lane]); h.shape = W.detail[1][lane]; h.prim = W.detail[2][lane]; h.shape_type = W.detail[3][lane];
        h.coords = mk3(u2f(W.detail[4][lane]), u2f(W.detail[5][lane]), u2f(W.detail[6][lane]));
        job.store(index, h);
    }
#endif
    if (STATS) {
        const u32 v[5] = {n_nodes, n_tests[0], n_iter, n_tests[1], n_retired};
#pragma unroll
        for (int k = 0; k < 5; k++) {
            const u32 t = __reduce_add_sync(0xFFFFFFFFu, v[k]);
            if (lane == 0 && t) atomicAdd(&stats[k], (unsigned long long)t);
        }
    }
}

// ---- K2: scatter ----
#ifndef PTB_SCATTER_THREADS
#define PTB_SCATTER_THREADS 256
#endif
#ifndef PTB_SCATTER_SORT
#define PTB_SCATTER_SORT 1
#endif
#ifndef PTB_SCATTER_MIN_BLOCKS
#define PTB_SCATTER_MIN_BLOCKS (1024 / PTB_SCATTER_THREADS)
#endif
// L2 prefetch distance of the direct scatter kernel, in blocks (0 = off).  A block's seven state records are seven
// contiguous 4 KB runs; one thread asks the L2 for the runs of the block that will start about half a wave of
// resident blocks later (148 SMs x 4 blocks / 2), with one bulk-prefetch instruction per run (TMA unit, no registers,
// no shared memory).  The kernel's loads then find their lines in L2 instead of waiting on HBM.
#ifndef PTB_SCATTER_PREFETCH_BLOCKS
#define PTB_SCATTER_PREFETCH_BLOCKS 296
#endif
__device__ __forceinline__ void prefetch_l2_bulk(const void* p, u32 bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p), "r"(bytes) : "memory");
}
// FEATURES: compile-time scene-feature mask (PTB_FEAT_*).  The fully general kernel is ~300 KB of
// SASS; with mixed materials its divergent warps thrash the instruction cache (ncu: 45 no_instruction
// stall cycles per issue on C3, profiles/r01_c3_scatter_generic.txt).  Two remedies, both here:
//   * k_scatter<F>       one thread per slot in slot order; the launcher picks the smallest F covering
//                        the scene (a diffuse-only scene runs a 5k-instruction kernel);
//   * material queues    k_classify sorts the slots of a round into {miss, diffuse, metal, other} lists
//                        and k_scatter_list<F> runs each list with the kernel specialised for it, so
//                        a warp only meets one material's code (the wavefront queues of the north star).
__device__ __forceinline__ bool scatter_slot(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st,
                                             float4* __restrict__ accum, u32 slot, u32 x, u32 y) {
    PathState path = load_path(st, slot);
    Ray ray = load_ray(st, slot);
    const PackedHit p = load_packed_hit(s, st, slot);
    const SurfaceHit hit = unpack_hit(p, ray);
    const bool finished = scatter_invocation(s, pc, x, y, path, ray, hit, accum);
    store_ray(st, slot, ray);
    store_path(st, slot, path);
    return finished;
}

template <u32 FEATURES>
__global__ void __launch_bounds__(PTB_SCATTER_THREADS, PTB_SCATTER_MIN_BLOCKS)
k_scatter(const DeviceScene scene, const RenderConsts pc, const WavefrontState st, float4* __restrict__ accum) {
    DeviceScene s = scene;
    s.features = FEATURES;               // a constant from here on: absent features fold away
    const u32 first = pc.slot_base + blockIdx.x * (u32)PTB_SCATTER_THREADS;
    u32 slot = first + threadIdx.x;
    u32 x, y;
    bool valid = slot_pixel(pc, slot, x, y);
#if PTB_SCATTER_PREFETCH_BLOCKS > 0
    if (threadIdx.x == 0) {
        const u32 ahead = first + (u32)PTB_SCATTER_PREFETCH_BLOCKS * (u32)PTB_SCATTER_THREADS;
        if (ahead + (u32)PTB_SCATTER_THREADS <= pc.slot_base + pc.slot_count) {
            const u32 bytes = (u32)PTB_SCATTER_THREADS * 16u;
            prefetch_l2_bulk(st.hit0 + ahead, bytes); prefetch_l2_bulk(st.hit1 + ahead, bytes); prefetch_l2_bulk(st.ray + ahead, bytes);
            prefetch_l2_bulk(st.path0 + ahead, bytes); prefetch_l2_bulk(st.path1 + ahead, bytes);
            prefetch_l2_bulk(st.path2 + ahead, bytes); prefetch_l2_bulk(st.path3 + ahead, bytes);
        }
    }
#endif
#if PTB_SCATTER_SORT
    // Hits and misses run different code (BSDF sampling vs sky + new camera path): the block's slots are dealt to
    // its threads hits first, so that all but one warp of a block are uniform.  The block still covers the same
    // PTB_SCATTER_THREADS consecutive slots, so its state traffic touches the same cache lines.
    __shared__ u32 order[PTB_SCATTER_THREADS];
    __shared__ u32 warp_hits[PTB_SCATTER_THREADS / 32], warp_rest[PTB_SCATTER_THREADS / 32];
    {
        const u32 lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
        const bool hit = valid && st.hit0[slot].y != 0xFFFFFFFFu;
        const u32 hb = __ballot_sync(0xFFFFFFFFu, hit);
        if (lane == 0) { warp_hits[warp] = (u32)__popc(hb); warp_rest[warp] = 32u - (u32)__popc(hb); }
        __syncthreads();
        u32 hits_before = 0, rest_before = 0, hits_total = 0;
#pragma unroll
        for (u32 w = 0; w < PTB_SCATTER_THREADS / 32; w++) {
            if (w < warp) { hits_before += warp_hits[w]; rest_before += warp_rest[w]; }
            hits_total += warp_hits[w];
        }
        const u32 below = (1u << lane) - 1u;
        const u32 pos = hit ? hits_before + (u32)__popc(hb & below) : hits_total + rest_before + (u32)__popc(~hb & below);
        order[pos] = threadIdx.x;
        __syncthreads();
        slot = first + order[threadIdx.x];
        valid = slot_pixel(pc, slot, x, y);
    }
#endif
    bool finished = false;
    if (valid) finished = scatter_slot(s, pc, st, accum, slot, x, y);
    // finished-path counters per 256-slot tile; no block barrier here, so warps retire as they finish
    const u32 c = (u32)__popc(__ballot_sync(0xFFFFFFFFu, finished));
    if ((threadIdx.x & 31u) == 0u && c != 0u) atomicAdd(&st.tile_paths[first >> 8], c);
    if (first + threadIdx.x == pc.slot_base) *st.ray_counter = 0u;       // re-arm the dynamic ray fetch of the next trace launch
}


// ---- K2, staged through shared memory by the TMA unit (ScatterMode DIRECT_STAGED: measured, NOT the default) ----
// The direct kernel above waits on memory more than it moves bytes or issues instructions (C2: 57 % of issue slots, 46 % of the
// measured HBM rate, 3.6 long-scoreboard stall cycles per issue): every block loads hit0, sorts hits first behind two barriers,
// and only then issues its 7 x 16-byte loads per slot.  This kernel tries the remedy the hardware offers: the state of a
// 256-slot tile - six contiguous 4 KB runs (ray, hit0, path0..3) - is brought into shared memory by six bulk asynchronous
// copies (cp.async.bulk + mbarrier, no registers, issued by one thread), two tiles deep: resident blocks take tile numbers from
// a counter (fetched one iteration early, so nobody waits for the atomic), and while a block shades tile i the copies of its
// tile i+1 are in flight.  Threads read the records of whichever slot the hits-first order deals them from shared memory (the
// permutation is free there) and store the new ray and path records themselves.  hit1 (hits only) stays a direct load from a
// run prefetched into L2 when the tile's copies are issued.  Same per-slot arithmetic: films and state are bit-identical
// (test_scatter_staged_equals_direct).
// RESULT (profiles/r02_g_ab_scatter_staged*.txt, C2): 0.180 ms against 0.151 ms for k_scatter, at the same 4 blocks per SM; storing
// through shared memory with bulk copies as well (a proxy fence and a third barrier per tile): 0.176 ms.  The 48 KB of staging
// per block take 200 KB of the SM's 256 KB away from L1, which the dependent fetches of finish_hit (shape transforms, face,
// vertices, material words: 56 % L1 hits in k_scatter) and the film pixels need more than the state loads need hiding - those
// already find their lines in L2 thanks to the bulk L2 prefetch 296 blocks ahead.  k_scatter stays the default.
#define PTB_STAGE_RECORDS 6u            // ray, hit0, path0, path1, path2, path3
#define PTB_STAGE_BYTES (PTB_STAGE_RECORDS * 256u * 16u)

__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, u32 parity) {
    u32 done;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src, u32 bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst_smem)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_store(void* dst, const void* src_smem, u32 bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" :: "l"(dst), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <u32 FEATURES>
__global__ void __launch_bounds__(256, 4)
k_scatter_staged(const DeviceScene scene, const RenderConsts pc, const WavefrontState st, float4* __restrict__ accum) {
    extern __shared__ __align__(128) unsigned char stage_mem[];          // 2 stages x 6 records x 256 slots x 16 B
    __shared__ unsigned long long full[2];
    __shared__ u32 next_tile[2];
    __shared__ u32 order[256];
    __shared__ u32 warp_hits[8], warp_rest[8];
    DeviceScene s = scene;
    s.features = FEATURES;
    const u32 tiles = pc.slot_count >> 8;
    unsigned* const tile_counter = st.ray_counter + 2;                   // zeroed by the trace kernel of this round
    const u32 lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;

    auto issue = [&](u32 tile, u32 k) {                                   // thread 0: the six runs of `tile` into stage k
        const size_t first = (size_t)pc.slot_base + (size_t)tile * 256u;
        uint4* dst = (uint4*)(stage_mem + (size_t)k * PTB_STAGE_BYTES);
        mbar_expect_tx(&full[k], PTB_STAGE_BYTES);
        bulk_load(dst + 0 * 256, st.ray + first, 4096u, &full[k]);
        bulk_load(dst + 1 * 256, st.hit0 + first, 4096u, &full[k]);
        bulk_load(dst + 2 * 256, st.path0 + first, 4096u, &full[k]);
        bulk_load(dst + 3 * 256, st.path1 + first, 4096u, &full[k]);
        bulk_load(dst + 4 * 256, st.path2 + first, 4096u, &full[k]);
        bulk_load(dst + 5 * 256, st.path3 + first, 4096u, &full[k]);
        prefetch_l2_bulk(st.hit1 + first, 4096u);
    };

    u32 ahead = 0;                                                         // thread 0: the tile number fetched one iteration early
    if (threadIdx.x == 0) {
        mbar_init(&full[0], 1u);
        mbar_init(&full[1], 1u);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (blockIdx.x == 0) *st.ray_counter = 0u;                        // re-arm the dynamic ray fetch of the next trace launch
        const u32 a = atomicAdd(tile_counter, 1u);
        next_tile[0] = a;
        if (a < tiles) issue(a, 0u);
        ahead = atomicAdd(tile_counter, 1u);                              // not needed before the first barrier of iteration 0
    }
    __syncthreads();

    for (u32 it = 0;; it++) {
        const u32 k = it & 1u;
        const u32 tile = next_tile[k];
        if (tile >= tiles) break;
        mbar_wait(&full[k], (it >> 1) & 1u);

        const uint4* rec = (const uint4*)(stage_mem + (size_t)k * PTB_STAGE_BYTES);
        const u32 first = pc.slot_base + tile * 256u;
        u32 x, y;
        u32 j = threadIdx.x;
        {
            // hits first (see k_scatter): all but one warp of the block run uniform code
            const bool hit = slot_pixel(pc, first + j, x, y) && rec[1 * 256 + j].y != 0xFFFFFFFFu;
            const u32 hb = __ballot_sync(0xFFFFFFFFu, hit);
            if (lane == 0) { warp_hits[warp] = (u32)__popc(hb); warp_rest[warp] = 32u - (u32)__popc(hb); }
            __syncthreads();
            // every thread is past the tile before this one: its stage is free, and the copies of the next tile go there now.
            // The tile number after that is fetched here as well and read one iteration later, so nobody waits for the atomic.
            if (threadIdx.x == 0) {
                next_tile[k ^ 1u] = ahead;
                if (ahead < tiles) issue(ahead, k ^ 1u);
                ahead = atomicAdd(tile_counter, 1u);
            }
            u32 hits_before = 0, rest_before = 0, hits_total = 0;
#pragma unroll
            for (u32 w = 0; w < 8; w++) {
                if (w < warp) { hits_before += warp_hits[w]; rest_before += warp_rest[w]; }
                hits_total += warp_hits[w];
            }
            const u32 below = (1u << lane) - 1u;
            const u32 pos = hit ? hits_before + (u32)__popc(hb & below) : hits_total + rest_before + (u32)__popc(~hb & below);
            order[pos] = threadIdx.x;
            __syncthreads();
            j = order[threadIdx.x];
        }
        const u32 slot = first + j;
        bool finished = false;
        if (slot_pixel(pc, slot, x, y)) {
            const uint4 h0 = rec[1 * 256 + j];
            Ray ray = ray_of(*(const float4*)&rec[0 * 256 + j]);
            PathState path = path_of(rec[2 * 256 + j], *(const float4*)&rec[3 * 256 + j], *(const float4*)&rec[4 * 256 + j], *(const float4*)&rec[5 * 256 + j]);
            PackedHit p = packed_miss();
            if (h0.y != 0xFFFFFFFFu) {
                const uint4 h1 = st.hit1[slot];
                HitId h;
                h.t = __uint_as_float(h0.x); h.shape = h0.y >> 16; h.prim = h0.z; h.shape_type = h0.w;
                h.coords = mk3(__uint_as_float(h1.x), __uint_as_float(h1.y), __uint_as_float(h1.z));
                const Surface sf = finish_hit(s, h);
                p = pack_hit(h, sf);
            }
            const SurfaceHit hit = unpack_hit(p, ray);
            finished = scatter_invocation(s, pc, x, y, path, ray, hit, accum);
            store_ray(st, slot, ray);
            store_path(st, slot, path);
        }
        // finished-path counters per 256-slot tile; no block barrier here, so warps go on to the next tile as they finish
        const u32 c = (u32)__popc(__ballot_sync(0xFFFFFFFFu, finished));
        if (lane == 0u && c != 0u) atomicAdd(&st.tile_paths[first >> 8], c);
    }
}

// ---- material queues ----
// Class of a slot for this round: what its scatter invocation can execute.
//   0 miss, 1 diffuse hit, 2 metal hit (all three: no enclosing medium), 3 everything else.
#define PTB_CLASSES 4
__global__ void __launch_bounds__(256) k_classify(const DeviceScene s, const RenderConsts pc, const WavefrontState st,
                                                   u32* __restrict__ lists, u32* __restrict__ counts) {
    __shared__ u32 warp_count[8][PTB_CLASSES];
    __shared__ u32 base[PTB_CLASSES];
    const u32 slot = pc.slot_base + blockIdx.x * 256u + threadIdx.x;
    const u32 lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    u32 x, y, cls = PTB_CLASSES;
    if (slot_pixel(pc, slot, x, y)) {
        const u32 sm = st.hit0[slot].y;
        const uint4 p0 = st.path0[slot];
        const bool media = (p0.z & p0.w) != 0xFFFFFFFFu || (s.features & PTB_FEAT_SCENE_SCATTER) != 0u;
        if (media) cls = 3u;
        else if (sm == 0xFFFFFFFFu) cls = 0u;
        else {
            const u32 type = __ldg(&s.material_data[32u * (sm & 0xFFFFu)]);
            cls = type == 0u ? 1u : type == 1u ? 2u : 3u;
        }
    }
    u32 rank = 0;
#pragma unroll
    for (u32 c = 0; c < PTB_CLASSES; c++) {
        const u32 b = __ballot_sync(0xFFFFFFFFu, cls == c);
        if (cls == c) rank = (u32)__popc(b & ((1u << lane) - 1u));
        if (lane == 0) warp_count[warp][c] = (u32)__popc(b);
    }
    __syncthreads();
    if (threadIdx.x < PTB_CLASSES) {
        u32 total = 0;
        for (u32 w = 0; w < 8; w++) { const u32 n = warp_count[w][threadIdx.x]; warp_count[w][threadIdx.x] = total; total += n; }
        base[threadIdx.x] = total ? atomicAdd(&counts[threadIdx.x], total) : 0u;
    }
    __syncthreads();
    if (cls < PTB_CLASSES) lists[(size_t)cls * pc.slot_count + base[cls] + warp_count[warp][cls] + rank] = slot;
    if (slot == pc.slot_base) *st.ray_counter = 0u;       // re-arm the dynamic ray fetch of the next trace launch
}

template <u32 FEATURES>
__global__ void __launch_bounds__(PTB_SCATTER_THREADS, PTB_SCATTER_MIN_BLOCKS)
k_scatter_list(const DeviceScene scene, const RenderConsts pc, const WavefrontState st, float4* __restrict__ accum,
               const u32* __restrict__ list, const u32* __restrict__ count) {
    DeviceScene s = scene;
    s.features = FEATURES;
    const u32 n = *count;
    const u32 stride = gridDim.x * (u32)PTB_SCATTER_THREADS;
    for (u32 i = blockIdx.x * (u32)PTB_SCATTER_THREADS + threadIdx.x; (i & ~31u) < n; i += stride) {      // warp-uniform trip count
        bool finished = false;
        u32 slot = 0xFFFFFFFFu;
        if (i < n) {
            slot = list[i];
            u32 x, y;
            slot_pixel(pc, slot, x, y);
            finished = scatter_slot(s, pc, st, accum, slot, x, y);
        }
        // finished-path counters: lanes of a warp mostly share a tile (lists keep slot order within a block)
        const u32 tile = finished ? slot >> 8 : 0xFFFFFFFFu;
        const u32 peers = __match_any_sync(0xFFFFFFFFu, tile);
        if (finished && (threadIdx.x & 31u) == (u32)__ffs((int)peers) - 1u) atomicAdd(&st.tile_paths[tile], (u32)__popc(peers));
    }
}

// ---- megakernel: rounds x {trace, scatter}, state in registers ----
template <int MODE>
__global__ void __launch_bounds__(128) k_mega(const DeviceScene s, RenderConsts pc, const WavefrontState st, float4* __restrict__ accum,
                                               u32 rounds, u32 seed_stride) {
    const u32 slot = blockIdx.x * 128u + threadIdx.x;
    u32 x, y;
    u32 finished = 0;
    if (slot_pixel(pc, slot, x, y)) {
        PathState path = load_path(st, slot);
        Ray ray = load_ray(st, slot);
        HitId h; Surface sf;
        h.shape = PTB_SHAPE_NONE; h.prim = 0xFFFFFFFFu; h.shape_type = 0; h.t = PTB_HIT_TIME_LIMIT;
        for (u32 k = 0; k < rounds; k++) {
            h = closest_hit<MODE>(s, ray);
            PackedHit p = packed_miss();
            if (h.shape != PTB_SHAPE_NONE) { sf = finish_hit(s, h); p = pack_hit(h, sf); }
            const SurfaceHit hit = unpack_hit(p, ray);
            if (scatter_invocation(s, pc, x, y, path, ray, hit, accum)) finished++;
            // the same quantisation the wavefront state applies between rounds
            ray.velocity = unpack_unit_vector(pack_unit_vector(ray.velocity));
            u32 w01, w23;
            pack_active(path.active, w01, w23);
            unpack_active(w01, w23, path.active);
            pc.random_seed += seed_stride;
        }
        store_ray(st, slot, ray);
        store_path(st, slot, path);
    }
    finished = __reduce_add_sync(0xFFFFFFFFu, finished);
    if ((threadIdx.x & 31u) == 0 && finished) atomicAdd(&st.tile_paths[slot >> 8], finished);
}

// ---- K4: resolve ----
__global__ void __launch_bounds__(256) k_resolve(const float4* __restrict__ accum, u32 count, ptb_resolve_params params,
                                                  uchar4* __restrict__ out_srgb8, float* __restrict__ out_linear) {
    const u32 i = blockIdx.x * 256u + threadIdx.x;
    if (i >= count) return;
    const f3 c = resolve_pixel(accum[i], params.Brightness, params.ToneMappingMode, params.ToneMappingWhiteLevel);
    if (out_srgb8) out_srgb8[i] = make_uchar4((unsigned char)linear_to_srgb8(c.x), (unsigned char)linear_to_srgb8(c.y), (unsigned char)linear_to_srgb8(c.z), 255);
    if (out_linear) { out_linear[3 * (size_t)i] = c.x; out_linear[3 * (size_t)i + 1] = c.y; out_linear[3 * (size_t)i + 2] = c.z; }
}

// ---- ray cast on caller rays ----
__device__ __forceinline__ void write_ray_hit(const DeviceScene& s, const HitId& h, ptb_hit* __restrict__ out, size_t i) {
    ptb_hit o;
    memset(&o, 0, sizeof o);
    o.t = h.t; o.shape = h.shape; o.prim = 0xFFFFFFFFu;
    if (h.shape != PTB_SHAPE_NONE) {
        const Surface sf = finish_hit(s, h);
        o.prim = h.prim; o.material = sf.material; o.shape_type = h.shape_type;
        o.nx = sf.normal.x; o.ny = sf.normal.y; o.nz = sf.normal.z;
        o.tx = sf.tangent_x.x; o.ty = sf.tangent_x.y; o.tz = sf.tangent_x.z;
        o.u = sf.uv.x; o.v = sf.uv.y;
        o.bx = h.coords.x; o.by = h.coords.y; o.bz = h.coords.z;
    }
    out[i] = o;
}
__device__ __forceinline__ void RaycastJob::store(u32 i, const HitId& h) const { write_ray_hit(s, h, out, i); }

template <int MODE>
__global__ void __launch_bounds__(128) k_raycast(const DeviceScene s, const ptb_ray* __restrict__ rays, size_t n, ptb_hit* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * 128u + threadIdx.x;
    if (i >= n) return;
    const ptb_ray in = rays[i];
    Ray r;
    r.origin = mk3(in.ox, in.oy, in.oz); r.velocity = mk3(in.dx, in.dy, in.dz); r.duration = in.duration;
    write_ray_hit(s, closest_hit<MODE>(s, r), out, i);
}

// Debug/test: the attributes scatter derives from the stored hits, in the reference's packed form.
__global__ void __launch_bounds__(256) k_hit_attributes(const DeviceScene s, const RenderConsts pc, const WavefrontState st, uint4* __restrict__ out) {
    const u32 slot = blockIdx.x * 256u + threadIdx.x;
    u32 x, y;
    if (!slot_pixel(pc, slot, x, y)) return;
    const PackedHit p = load_packed_hit(s, st, slot);
    out[slot] = make_uint4(p.normal, p.tangent, __float_as_uint(p.u), __float_as_uint(p.v));
}

// ---- preview visualiser (preview_render.glsl): one thread per pixel, image + shape ids to device buffers ----
template <int MODE>
__global__ void __launch_bounds__(128) k_preview(const DeviceScene s, const PreviewConsts p, float4* __restrict__ rgba, u32* __restrict__ shapes,
                                                  u32* __restrict__ hit_shape) {
    const u32 i = blockIdx.x * 128u + threadIdx.x;
    if (i >= p.width * p.height) return;
    const u32 px = i % p.width, py = i / p.width;
    u32 shape;
    const f3 c = preview_pixel<MODE>(s, p, px, py, &shape);
    rgba[i] = make_float4(c.x, c.y, c.z, 1.0f);
    if (shapes) shapes[i] = shape;
    if (px == p.mouse_x && py == p.mouse_y) *hit_shape = shape;      // the pick query (preview_render.glsl:172-173)
}

// ---- launchers ----
cudaError_t launch_restart(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum, cudaStream_t stream) {
    k_restart<<<pc.slot_count / 256u, 256, 0, stream>>>(s, pc, st, accum);
    return cudaGetLastError();
}

// Persistent grid: exactly the blocks that are resident at once (SMs x occupancy) on the CURRENT device.  Cached per
// kernel instantiation and per device (one process may drive several GPUs: ptb_group_create_local).
#define PTB_MAX_DEVICES 64
template <class K>
static unsigned persistent_grid(K kernel, int threads, unsigned* cache /* [PTB_MAX_DEVICES] */) {
    int dev = 0, sms = 148, per_sm = 1;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < PTB_MAX_DEVICES && cache[dev]) return cache[dev];
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0);
    if (per_sm < 1) per_sm = 1;
    const unsigned grid = (unsigned)(sms * per_sm);
    if (dev >= 0 && dev < PTB_MAX_DEVICES) cache[dev] = grid;      // racing threads write the same value
    return grid;
}

// Variants of the wide trace kernel (TraceVariant of the render parameters; 0 = one thread per slot):
//   1-4   k_trace_persistent, refill threshold (idle lanes that trigger a fetch) 8 / 4 / 16 / 32
//   5-    k_trace_coop {refill, speculate, waiting lanes that force a partial flush}
// Parking leaf work on the stack when few lanes have any (vote-driven postponing) was measured in
// round 1 and did not pay on B200; it was removed.  Sweep: profiles/r01_v3_variant_sweep.txt.
#define PTB_PERSISTENT_VARIANTS(X) X(1, 8) X(2, 4) X(3, 16) X(4, 32)
#define PTB_COOP_VARIANTS(X) X(5, 8, false, 8) X(6, 8, false, 16) X(7, 4, true, 8) X(8, 2, true, 8) X(9, 8, false, 4) \
                             X(10, 8, true, 8) X(11, 8, true, 16) X(12, 8, true, 32) X(13, 8, true, 4) X(14, 4, true, 16) X(15, 12, true, 8) \
                             X(16, 16, true, 8) X(17, 6, true, 8) X(18, 6, true, 12) X(19, 1, true, 8)

template <class Job>
static cudaError_t launch_persistent(const Job& job, unsigned variant, cudaStream_t stream) {
    switch (variant) {
#define X(ID, R) case ID: { static unsigned cache[PTB_MAX_DEVICES]; const unsigned grid = persistent_grid(k_trace_persistent<Job, R>, 32 * PTB_TRACE_WARPS, cache); \
                            k_trace_persistent<Job, R><<<grid, 32 * PTB_TRACE_WARPS, 0, stream>>>(job); break; }
        PTB_PERSISTENT_VARIANTS(X)
#undef X
#define X(ID, R, S, W) case ID: { static unsigned cache[PTB_MAX_DEVICES]; const unsigned grid = persistent_grid(k_trace_coop<Job, R, S, W>, 32 * PTB_TRACE_WARPS, cache); \
                                  k_trace_coop<Job, R, S, W><<<grid, 32 * PTB_TRACE_WARPS, 0, stream>>>(job); break; }
        PTB_COOP_VARIANTS(X)
#undef X
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_trace(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, unsigned traversal, unsigned tie,
                         unsigned variant, cudaStream_t stream) {
    const int mode = mode_of(traversal, tie);
    if (mode == MODE_WIDE && variant != 0) {
        WavefrontJob job;
        job.s = s; job.pc = pc; job.st = st;
        return launch_persistent(job, variant, stream);
    }
    const unsigned grid = pc.slot_count / 128u;
    switch (mode) {
        case MODE_BVH2_FAITHFUL: k_trace<MODE_BVH2_FAITHFUL><<<grid, 128, 0, stream>>>(s, pc, st); break;
        case MODE_BVH2_CANONICAL: k_trace<MODE_BVH2_CANONICAL><<<grid, 128, 0, stream>>>(s, pc, st); break;
        default: k_trace<MODE_WIDE><<<grid, 128, 0, stream>>>(s, pc, st); break;
    }
    return cudaGetLastError();
}

cudaError_t launch_trace_stats(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, unsigned long long* stats, cudaStream_t stream) {
    WavefrontJob job;
    job.s = s; job.pc = pc; job.st = st;
    static unsigned cache[PTB_MAX_DEVICES];
    auto kernel = k_trace_coop<WavefrontJob, 4, true, 16, true>;          // = PTB_TRACE_VARIANT_DEFAULT (14) + counters
    const unsigned grid = persistent_grid(kernel, 32 * PTB_TRACE_WARPS, cache);
    kernel<<<grid, 32 * PTB_TRACE_WARPS, 0, stream>>>(job, stats);
    return cudaGetLastError();
}

cudaError_t launch_hit_attributes(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, uint4* out, cudaStream_t stream) {
    k_hit_attributes<<<pc.slot_count / 256u, 256, 0, stream>>>(s, pc, st, out);
    return cudaGetLastError();
}

template <u32 F>
static cudaError_t launch_staged(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum, cudaStream_t stream) {
    static unsigned cache[PTB_MAX_DEVICES];
    const size_t smem = 2u * (size_t)PTB_STAGE_BYTES;
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= PTB_MAX_DEVICES || !cache[dev]) {
        cudaError_t e = cudaFuncSetAttribute(k_scatter_staged<F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
#ifdef PTB_STAGED_CARVEOUT
        cudaFuncSetAttribute(k_scatter_staged<F>, cudaFuncAttributePreferredSharedMemoryCarveout, PTB_STAGED_CARVEOUT);
#endif
        int sms = 148, per_sm = 1;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_scatter_staged<F>, 256, smem);
        if (per_sm < 1) per_sm = 1;
        if (getenv("PTB_DEBUG_OCCUPANCY")) fprintf(stderr, "k_scatter_staged<%u>: %d blocks/SM, %d SMs, %zu B dynamic smem\n", F, per_sm, sms, smem);
        if (dev >= 0 && dev < PTB_MAX_DEVICES) cache[dev] = (unsigned)(sms * per_sm);
    }
    const unsigned resident = (dev >= 0 && dev < PTB_MAX_DEVICES) ? cache[dev] : 148u;
    const unsigned tiles = pc.slot_count >> 8;
    k_scatter_staged<F><<<tiles < resident ? tiles : resident, 256, smem, stream>>>(s, pc, st, accum);
    return cudaGetLastError();
}

cudaError_t launch_scatter(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum, bool staged, cudaStream_t stream) {
    const unsigned grid = pc.slot_count / (unsigned)PTB_SCATTER_THREADS;
    const u32 opaque = PTB_FEAT_DIFFUSE | PTB_FEAT_METAL | PTB_FEAT_TEXTURES | PTB_FEAT_SKY_TEXTURE;
    if (staged) {
        if ((s.features & ~PTB_FEAT_DIFFUSE) == 0u) return launch_staged<PTB_FEAT_DIFFUSE>(s, pc, st, accum, stream);
        if ((s.features & ~opaque) == 0u) return launch_staged<opaque>(s, pc, st, accum, stream);
        return launch_staged<PTB_FEAT_ALL>(s, pc, st, accum, stream);
    }
    if ((s.features & ~PTB_FEAT_DIFFUSE) == 0u) k_scatter<PTB_FEAT_DIFFUSE><<<grid, PTB_SCATTER_THREADS, 0, stream>>>(s, pc, st, accum);
    else if ((s.features & ~opaque) == 0u) k_scatter<opaque><<<grid, PTB_SCATTER_THREADS, 0, stream>>>(s, pc, st, accum);
    else k_scatter<PTB_FEAT_ALL><<<grid, PTB_SCATTER_THREADS, 0, stream>>>(s, pc, st, accum);
    return cudaGetLastError();
}

// Material queues: classify, then one specialised kernel per class.  `queue` = PTB_CLASSES lists of slot_count
// entries followed by PTB_CLASSES counters.
size_t scatter_queue_bytes(unsigned slot_count) { return ((size_t)PTB_CLASSES * slot_count + PTB_CLASSES) * sizeof(u32); }

template <u32 F>
static void launch_list(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum, const u32* list, const u32* count,
                        cudaStream_t stream) {
    static unsigned cache[PTB_MAX_DEVICES];
    const unsigned grid = persistent_grid(k_scatter_list<F>, PTB_SCATTER_THREADS, cache);
    k_scatter_list<F><<<grid, PTB_SCATTER_THREADS, 0, stream>>>(s, pc, st, accum, list, count);
}

cudaError_t launch_scatter_queues(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum, unsigned* queue,
                                  cudaStream_t stream) {
    u32* lists = queue;
    u32* counts = queue + (size_t)PTB_CLASSES * pc.slot_count;
    cudaError_t e = cudaMemsetAsync(counts, 0, PTB_CLASSES * sizeof(u32), stream);
    if (e != cudaSuccess) return e;
    k_classify<<<pc.slot_count / 256u, 256, 0, stream>>>(s, pc, st, lists, counts);
    const size_t S = pc.slot_count;
    launch_list<PTB_FEAT_SKY_TEXTURE>(s, pc, st, accum, lists + 0 * S, counts + 0, stream);
    if (s.features & PTB_FEAT_DIFFUSE) launch_list<PTB_FEAT_DIFFUSE | PTB_FEAT_TEXTURES>(s, pc, st, accum, lists + 1 * S, counts + 1, stream);
    if (s.features & PTB_FEAT_METAL) launch_list<PTB_FEAT_METAL | PTB_FEAT_TEXTURES>(s, pc, st, accum, lists + 2 * S, counts + 2, stream);
    if (s.features & (PTB_FEAT_TRANSLUCENT | PTB_FEAT_SCENE_SCATTER)) launch_list<PTB_FEAT_ALL>(s, pc, st, accum, lists + 3 * S, counts + 3, stream);
    return cudaGetLastError();
}
unsigned scatter_queue_launches(const DeviceScene& s) {
    return 2u + ((s.features & PTB_FEAT_DIFFUSE) ? 1u : 0u) + ((s.features & PTB_FEAT_METAL) ? 1u : 0u)
              + ((s.features & (PTB_FEAT_TRANSLUCENT | PTB_FEAT_SCENE_SCATTER)) ? 1u : 0u);
}

cudaError_t launch_megakernel(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum,
                              unsigned traversal, unsigned tie, unsigned rounds, unsigned seed_stride, cudaStream_t stream) {
    const unsigned grid = pc.slot_count / 128u;
    switch (mode_of(traversal, tie)) {
        case MODE_BVH2_FAITHFUL: k_mega<MODE_BVH2_FAITHFUL><<<grid, 128, 0, stream>>>(s, pc, st, accum, rounds, seed_stride); break;
        case MODE_BVH2_CANONICAL: k_mega<MODE_BVH2_CANONICAL><<<grid, 128, 0, stream>>>(s, pc, st, accum, rounds, seed_stride); break;
        default: k_mega<MODE_WIDE><<<grid, 128, 0, stream>>>(s, pc, st, accum, rounds, seed_stride); break;
    }
    return cudaGetLastError();
}

cudaError_t launch_preview(const DeviceScene& s, const PreviewConsts& p, unsigned traversal, unsigned tie, float4* rgba, unsigned* shapes,
                           unsigned* hit_shape, cudaStream_t stream) {
    const unsigned n = p.width * p.height;
    if (n == 0) return cudaSuccess;
    const unsigned grid = (n + 127u) / 128u;
    switch (mode_of(traversal, tie)) {
        case MODE_BVH2_FAITHFUL: k_preview<MODE_BVH2_FAITHFUL><<<grid, 128, 0, stream>>>(s, p, rgba, shapes, hit_shape); break;
        case MODE_BVH2_CANONICAL: k_preview<MODE_BVH2_CANONICAL><<<grid, 128, 0, stream>>>(s, p, rgba, shapes, hit_shape); break;
        default: k_preview<MODE_WIDE><<<grid, 128, 0, stream>>>(s, p, rgba, shapes, hit_shape); break;
    }
    return cudaGetLastError();
}

cudaError_t launch_resolve(const float4* accum, unsigned width, unsigned height, ptb_resolve_params params,
                           uchar4* out_srgb8, float* out_linear_rgb, cudaStream_t stream) {
    const unsigned count = width * height;
    k_resolve<<<(count + 255u) / 256u, 256, 0, stream>>>(accum, count, params, out_srgb8, out_linear_rgb);
    return cudaGetLastError();
}

cudaError_t launch_raycast(const DeviceScene& s, const ptb_ray* rays, size_t n, unsigned traversal, unsigned tie, unsigned variant,
                           unsigned* counter, ptb_hit* out, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    const int mode = mode_of(traversal, tie);
    if (mode == MODE_WIDE && variant != 0 && counter && n < 0xF0000000ull) {
        cudaError_t e = cudaMemsetAsync(counter, 0, sizeof(unsigned), stream);
        if (e != cudaSuccess) return e;
        RaycastJob job;
        job.s = s; job.rays = rays; job.out = out; job.n = (u32)n; job.ctr = counter;
        return launch_persistent(job, variant, stream);
    }
    const unsigned grid = (unsigned)((n + 127) / 128);
    switch (mode) {
        case MODE_BVH2_FAITHFUL: k_raycast<MODE_BVH2_FAITHFUL><<<grid, 128, 0, stream>>>(s, rays, n, out); break;
        case MODE_BVH2_CANONICAL: k_raycast<MODE_BVH2_CANONICAL><<<grid, 128, 0, stream>>>(s, rays, n, out); break;
        default: k_raycast<MODE_WIDE><<<grid, 128, 0, stream>>>(s, rays, n, out); break;
    }
    return cudaGetLastError();
}

}  // namespace ptb
