// Software traversal of the compressed 8-wide BVH (built by ptb_wide_build.cpp).
//
// This replaces the reference's nested BVH2 walks (src/scene/scene.glsl.inc
// :336-399 per mesh, :468-520 over shapes) on the hot path.  One level, world
// space, 80-byte nodes read as five 16-byte loads, child boxes quantised to
// 8 bits per plane, children visited in ray-octant order from a bit mask, and a
// stack of 8-byte {base, mask} groups.  Leaves hold up to three primitives, each
// 48 bytes (three 16-byte loads) carrying the reference's OBJECT-space vertex
// positions, global face index and shape index: the exact primitive test is the
// reference's arithmetic on the object-space ray (ptb_trace.cuh), so a hit's t,
// face and shape are bit-identical to the reference's; the tree only decides
// which primitives get tested, and its slab test is made conservative below.
//
// Node (20 words):
//   w0-2  p.xyz        origin of the quantisation grid (node lower corner)
//   w3    ex | ey<<8 | ez<<16 | imask<<24      biased exponents; imask = slots holding inner nodes
//   w4    child_base   index of the first inner child (inner children are contiguous in slot order)
//   w5    prim_base    index of the first leaf primitive (leaf primitives are contiguous in slot order)
//   w6    valid24      leaf primitives: slot s owns bits 3s..3s+2, one bit per primitive; the primitives
//                      are stored compacted, so primitive of bit b = prim_base + popc(valid24 below b)
//   w7    unused
//   w8-13 qlo x,y,z    8 bytes each
//   w14-19 qhi x,y,z
//
// The building blocks (ray context, node step, primitive test) are shared by the
// per-thread query closest_hit_wide() — used by the megakernel and the ray-cast
// harness — and by the persistent warp-cooperative trace kernel in ptb_kernels.cu.
#pragma once
#include "ptb_trace.cuh"

namespace ptb {

#define PTB_WIDE_STACK 32      // tree depth <= 30, checked by the builder

PTB_HD u32 bytes_sign_extend(u32 x) {     // per byte: 0xFF if its top bit is set, else 0
#if defined(__CUDA_ARCH__)
    u32 r;
    asm("prmt.b32 %0, %1, 0, 0xBA98;" : "=r"(r) : "r"(x));
    return r;
#else
    u32 r = 0;
    for (int i = 0; i < 4; i++) if (x & (0x80u << (8 * i))) r |= 0xFFu << (8 * i);
    return r;
#endif
}
PTB_HD u32 highest_bit(u32 x) {
#if defined(__CUDA_ARCH__)
    return 31u - (u32)__clz((int)x);
#else
    return 31u - (u32)__builtin_clz(x);
#endif
}
PTB_HD u32 popcount(u32 x) {
#if defined(__CUDA_ARCH__)
    return (u32)__popc(x);
#else
    return (u32)__builtin_popcount(x);
#endif
}
// Byte j of w as the float 1 + q * 2^-15, built with one byte permute (ALU pipe)
// instead of an int->float conversion (XU pipe, quarter rate): bytes = 3F 80 qq 00.
// `one_bits` must hold 0x3F800000 in a REGISTER whose value ptxas cannot see (it is read
// from constant memory once per kernel): SASS PRMT takes only one non-register operand,
// and it has to be the selector, otherwise every PRMT drags a MOV of its selector along.
#if defined(__CUDACC__)
__constant__ u32 k_one_bits = 0x3F800000u;
#endif
PTB_HD u32 load_one_bits() {
#if defined(__CUDA_ARCH__)
    return k_one_bits;
#else
    return 0x3F800000u;
#endif
}
template <int J>
PTB_HD float byte_as_unit_float(u32 w, u32 one_bits) {
#if defined(__CUDA_ARCH__)
    u32 r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(w), "r"(one_bits), "n"(0x7604 | (J << 4)));
    return __uint_as_float(r);
#else
    return u2f(one_bits | (((w >> (8 * J)) & 0xFFu) << 8));
#endif
}

// Reciprocal for the culling structure only (<= 1 ulp; the slab bias below budgets for it).
PTB_HD float fast_rcp(float x) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
#else
    return 1.0f / x;
#endif
}

// Per-ray constants of the traversal.
struct WideCtx {
    f3 origin;
    f3 inv_dir;          // 1/d with |d| clamped away from zero (sign kept, -0 counts as negative)
    bool nx, ny, nz;
    u32 octinv;
    u32 swap_masks;      // 0x55 / 0x33 / 0x0F in bytes 0 / 1 / 2 where the octant flips that index bit, else 0
    u32 one_bits;        // 0x3F800000, opaque to the compiler (see byte_as_unit_float)
};

PTB_HD WideCtx wide_ctx(const Ray& r) {
    // The tree is a culling structure: it only has to be conservative (the reference's own
    // slab test divides by zero instead, common.glsl.inc:157-158).
    const float tiny = 8.271806e-25f;   // 2^-80
    const f3 d = r.velocity;
    WideCtx c;
    c.origin = r.origin;
    c.inv_dir = mk3(fast_rcp(fabsf(d.x) > tiny ? d.x : copysignf(tiny, d.x)),
                    fast_rcp(fabsf(d.y) > tiny ? d.y : copysignf(tiny, d.y)),
                    fast_rcp(fabsf(d.z) > tiny ? d.z : copysignf(tiny, d.z)));
    c.nx = c.inv_dir.x < 0.0f; c.ny = c.inv_dir.y < 0.0f; c.nz = c.inv_dir.z < 0.0f;
    c.octinv = 7u - ((c.nx ? 1u : 0u) | (c.ny ? 2u : 0u) | (c.nz ? 4u : 0u));
    c.swap_masks = ((c.octinv & 1u) ? 0x55u : 0u) | ((c.octinv & 2u) ? 0x3300u : 0u) | ((c.octinv & 4u) ? 0x0F0000u : 0u);
    c.one_bits = load_one_bits();
    return c;
}

// Slab test of one quantised child box.  Plane distance along the ray: t = q*a + b with
// a = 2^e/d, b = (p - o)/d, evaluated as fma(1 + q*2^-15, A, B) with A = 32768 a, B = b - A.
// The lo/hi biases carry the rounding-error bound of that evaluation (including the 1-ulp
// reciprocal), so the test can only err towards "hit".
template <int J>
PTB_HD u32 wide_test_child(u32 nearx, u32 farx, u32 neary, u32 fary, u32 nearz, u32 farz,
                           float Ax, float Ay, float Az, float Bxl, float Bxh, float Byl, float Byh, float Bzl, float Bzh, float reach, u32 one) {
    const float t0x = fmaf(byte_as_unit_float<J>(nearx, one), Ax, Bxl), t1x = fmaf(byte_as_unit_float<J>(farx, one), Ax, Bxh);
    const float t0y = fmaf(byte_as_unit_float<J>(neary, one), Ay, Byl), t1y = fmaf(byte_as_unit_float<J>(fary, one), Ay, Byh);
    const float t0z = fmaf(byte_as_unit_float<J>(nearz, one), Az, Bzl), t1z = fmaf(byte_as_unit_float<J>(farz, one), Az, Bzh);
    const float tmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.0f));
    const float tmax = fminf(fminf(t1x, t1y), fminf(t1z, reach));
    return tmin <= tmax ? 1u : 0u;
}

// Leaf primitives hit in one node: bit b set = primitive prim_base + popc(valid & below(b)).
struct PrimGroup { u32 base, bits, valid; };

// Visits the nearest pending child of node group `ng`: pushes the remainder of the group,
// loads the child node, tests its 8 children against [0, reach], and returns the child's own
// node group (in ng) and its hit leaf primitives (in pg).
template <class Stack>
PTB_HD void wide_node_step(const DeviceScene& s, const WideCtx& c, float reach, uint2& ng, PrimGroup& pg, Stack& stack) {
    const u32 hits = ng.y;
    const u32 bit = highest_bit(hits);
    ng.y &= ~(1u << bit);
    if (ng.y > 0x00FFFFFFu) stack.push(ng);
    const u32 slot = (bit - 24u) ^ c.octinv;
    const u32 rel = popcount(hits & ~(0xFFFFFFFFu << slot));
    const uint4* np = s.wide_nodes + 5 * (size_t)(ng.x + rel);
    const uint4 n0 = PTB_LDG(&np[0]), n1 = PTB_LDG(&np[1]), n2 = PTB_LDG(&np[2]), n3 = PTB_LDG(&np[3]), n4 = PTB_LDG(&np[4]);

    // A = 2^(e+15) / d ;  B = (p - o)/d - A ;  error bound 2^-21 (|b| + |A|)
    const float Ax = u2f(((n0.w & 0xFFu) + 15u) << 23) * c.inv_dir.x;
    const float Ay = u2f((((n0.w >> 8) & 0xFFu) + 15u) << 23) * c.inv_dir.y;
    const float Az = u2f((((n0.w >> 16) & 0xFFu) + 15u) << 23) * c.inv_dir.z;
    const float bx = (u2f(n0.x) - c.origin.x) * c.inv_dir.x;
    const float by = (u2f(n0.y) - c.origin.y) * c.inv_dir.y;
    const float bz = (u2f(n0.z) - c.origin.z) * c.inv_dir.z;
    const float ex = (fabsf(bx) + fabsf(Ax)) * 4.7683716e-7f;
    const float ey = (fabsf(by) + fabsf(Ay)) * 4.7683716e-7f;
    const float ez = (fabsf(bz) + fabsf(Az)) * 4.7683716e-7f;
    const float Bx = bx - Ax, By = by - Ay, Bz = bz - Az;
    const float Bxl = Bx - ex, Bxh = Bx + ex, Byl = By - ey, Byh = By + ey, Bzl = Bz - ez, Bzh = Bz + ez;

    u32 hit8 = 0;      // bit = slot
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const u32 qlx = half ? n2.y : n2.x, qly = half ? n2.w : n2.z, qlz = half ? n3.y : n3.x;
        const u32 qhx = half ? n3.w : n3.z, qhy = half ? n4.y : n4.x, qhz = half ? n4.w : n4.z;
        const u32 nearx = c.nx ? qhx : qlx, farx = c.nx ? qlx : qhx;
        const u32 neary = c.ny ? qhy : qly, fary = c.ny ? qly : qhy;
        const u32 nearz = c.nz ? qhz : qlz, farz = c.nz ? qlz : qhz;
        hit8 |= wide_test_child<0>(nearx, farx, neary, fary, nearz, farz, Ax, Ay, Az, Bxl, Bxh, Byl, Byh, Bzl, Bzh, reach, c.one_bits) << (4 * half + 0);
        hit8 |= wide_test_child<1>(nearx, farx, neary, fary, nearz, farz, Ax, Ay, Az, Bxl, Bxh, Byl, Byh, Bzl, Bzh, reach, c.one_bits) << (4 * half + 1);
        hit8 |= wide_test_child<2>(nearx, farx, neary, fary, nearz, farz, Ax, Ay, Az, Bxl, Bxh, Byl, Byh, Bzl, Bzh, reach, c.one_bits) << (4 * half + 2);
        hit8 |= wide_test_child<3>(nearx, farx, neary, fary, nearz, farz, Ax, Ay, Az, Bxl, Bxh, Byl, Byh, Bzl, Bzh, reach, c.one_bits) << (4 * half + 3);
    }
    const u32 imask = n0.w >> 24;

    // inner children: slot s goes to bit 24 + (s ^ octinv), so the highest set bit is the child
    // nearest along the ray's octant (three conditional swaps = XOR on the bit index)
    u32 x = hit8 & imask, t;
    t = ((x >> 1) ^ x) & c.swap_masks;         x = x ^ t ^ (t << 1);      // x < 256, so only the mask's low byte matters
    t = ((x >> 2) ^ x) & (c.swap_masks >> 8);  x = x ^ t ^ (t << 2);
    t = ((x >> 4) ^ x) & (c.swap_masks >> 16); x = x ^ t ^ (t << 4);
    ng.x = n1.x;
    ng.y = (x << 24) | imask;

    // leaf children: spread bit s to bit 3s, widen to the slot's 3 bits, keep existing primitives
    u32 y = hit8 & ~imask;
    y = (y | (y << 8)) & 0x0000F00Fu;
    y = (y | (y << 4)) & 0x000C30C3u;
    y = (y | (y << 2)) & 0x00249249u;
    pg.base = n1.y;
    pg.valid = n1.z;
    pg.bits = (y * 7u) & n1.z;
}

// Takes the next primitive out of the group and tests it: exact reference test on the
// object-space ray of its shape (cached across consecutive primitives of the same shape).
PTB_HD void wide_prim_step(const DeviceScene& s, const Ray& world, PrimGroup& pg, u32& cur_shape, Ray& obj, HitId& h) {
    const u32 b = highest_bit(pg.bits);
    pg.bits &= ~(1u << b);
    const u32 index = pg.base + popcount(pg.valid & ((1u << b) - 1u));
    const float4* tp = s.wide_tris + 3 * (size_t)index;
    const float4 v0 = PTB_LDG(&tp[0]), v1 = PTB_LDG(&tp[1]), v2 = PTB_LDG(&tp[2]);
    const u32 face = f2u(v0.w), shape = f2u(v1.w), kind = f2u(v2.w) & 3u;
    if (shape != cur_shape) { obj = object_ray(s, world, shape); cur_shape = shape; }
    if (kind == 0u) test_triangle<true>(obj, mk3(v0.x, v0.y, v0.z), mk3(v1.x, v1.y, v1.z), mk3(v2.x, v2.y, v2.z), shape, face, h);
    else test_analytic<true>(obj, kind, shape, h);
}

// Unbounded shapes (planes) are not in the tree: tested first, which also shortens the reach.
PTB_HD HitId wide_begin(const DeviceScene& s, const Ray& r) {
    HitId h;
    h.t = r.duration; h.shape = PTB_SHAPE_NONE; h.prim = 0xFFFFFFFFu; h.shape_type = 0; h.coords = mk3(0, 0, 0);
    for (u32 i = 0; i < s.wide_plane_count; i++) {
        const u32 shape = PTB_LDG(&s.wide_planes[2 * i]);
        const Ray o = object_ray(s, r, shape);
        test_analytic<true>(o, PTB_SHAPE_PLANE, shape, h);
    }
    return h;
}

// Per-thread stack in local memory (megakernel, ray-cast harness, host emulation).
struct LocalStack {
    uint2 e[PTB_WIDE_STACK];
    int sp;
    PTB_HD void push(uint2 v) { if (sp < PTB_WIDE_STACK) e[sp++] = v; }
    PTB_HD uint2 pop() { return e[--sp]; }
    PTB_HD bool empty() const { return sp == 0; }
};

// Canonical closest hit: lexicographically lowest (t, shape, face) among all
// primitives the reference tests would accept with t <= ray.duration.
PTB_HD HitId closest_hit_wide(const DeviceScene& s, const Ray& r, Complexity* cx = nullptr) {
    HitId h = wide_begin(s, r);
    if (s.wide_tlas_root == 0xFFFFFFFFu) return h;
    const WideCtx c = wide_ctx(r);
    LocalStack stack;
    stack.sp = 0;
    uint2 ng = make_uint2(0u, 0x80000000u);     // root pseudo-group: one hit, slot 7^octinv, no inner mask
    u32 cur_shape = PTB_SHAPE_NONE;
    Ray obj = r;
    while (true) {
        PrimGroup pg;
        wide_node_step(s, c, h.t, ng, pg, stack);
        if (cx) { cx->scene++; cx->mesh += popcount(pg.bits); }
        while (pg.bits) wide_prim_step(s, r, pg, cur_shape, obj, h);
        if (ng.y <= 0x00FFFFFFu) {
            if (stack.empty()) break;
            ng = stack.pop();
        }
    }
    return h;
}

}  // namespace ptb
