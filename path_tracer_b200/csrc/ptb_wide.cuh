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
//   w6-7  meta[8]      0 = empty; inner: 0b001_11sss (sss = slot); leaf: unary(count)<<5 | offset
//   w8-13 qlo x,y,z    8 bytes each
//   w14-19 qhi x,y,z
#pragma once
#include "ptb_trace.cuh"

namespace ptb {

#define PTB_WIDE_STACK 32

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
PTB_HD float byte_to_float(u32 w, int j) { return (float)((w >> (8 * j)) & 0xFFu); }

// Canonical closest hit: lexicographically lowest (t, shape, face) among all
// primitives the reference tests would accept with t <= ray.duration.
PTB_HD HitId closest_hit_wide(const DeviceScene& s, const Ray& r) {
    HitId h;
    h.t = r.duration; h.shape = PTB_SHAPE_NONE; h.prim = 0xFFFFFFFFu; h.shape_type = 0; h.coords = mk3(0, 0, 0);

    // unbounded shapes first: cheap, and they shorten the reach before traversal
    for (u32 i = 0; i < s.wide_plane_count; i++) {
        u32 shape = PTB_LDG(&s.wide_planes[i]);
        Ray o = object_ray(s, r, shape);
        test_analytic<true>(o, PTB_SHAPE_PLANE, shape, h);
    }
    if (s.wide_tlas_root == 0xFFFFFFFFu) return h;

    // Reciprocal direction with |d| clamped away from zero: the tree is a culling
    // structure, it only has to be conservative (the reference's own slab test
    // divides by zero instead, common.glsl.inc:157-158).
    const float tiny = 8.271806e-25f;   // 2^-80
    f3 d = r.velocity;
    f3 id = mk3(1.0f / (fabsf(d.x) > tiny ? d.x : copysignf(tiny, d.x)),
                1.0f / (fabsf(d.y) > tiny ? d.y : copysignf(tiny, d.y)),
                1.0f / (fabsf(d.z) > tiny ? d.z : copysignf(tiny, d.z)));
    const bool nx = id.x < 0.0f, ny = id.y < 0.0f, nz = id.z < 0.0f;   // sign of the clamped reciprocal: -0 counts as negative
    const u32 octinv = 7u - ((nx ? 1u : 0u) | (ny ? 2u : 0u) | (nz ? 4u : 0u));
    const u32 octinv4 = octinv * 0x01010101u;

    uint2 stack[PTB_WIDE_STACK];
    int sp = 0;
    uint2 ng = make_uint2(0u, 0x80000000u);     // root pseudo-group: one hit, slot 7^octinv, no inner mask
    u32 cur_shape = PTB_SHAPE_NONE;
    Ray obj = r;

    while (true) {
        // ---- pop the nearest pending child of the current node group ----
        const u32 hits = ng.y;
        const u32 bit = highest_bit(hits);
        ng.y &= ~(1u << bit);
        if (ng.y > 0x00FFFFFFu && sp < PTB_WIDE_STACK) stack[sp++] = ng;
        const u32 slot = (bit - 24u) ^ octinv;
        const u32 rel = popcount(hits & ~(0xFFFFFFFFu << slot));
        const uint4* np = s.wide_nodes + 5 * (size_t)(ng.x + rel);
        const uint4 n0 = PTB_LDG(&np[0]), n1 = PTB_LDG(&np[1]), n2 = PTB_LDG(&np[2]), n3 = PTB_LDG(&np[3]), n4 = PTB_LDG(&np[4]);

        // ---- slab test of the 8 quantised child boxes ----
        const float ax = u2f((n0.w & 0xFFu) << 23) * id.x;
        const float ay = u2f(((n0.w >> 8) & 0xFFu) << 23) * id.y;
        const float az = u2f(((n0.w >> 16) & 0xFFu) << 23) * id.z;
        const float bx = (u2f(n0.x) - r.origin.x) * id.x;
        const float by = (u2f(n0.y) - r.origin.y) * id.y;
        const float bz = (u2f(n0.z) - r.origin.z) * id.z;
        // rounding-error bound of q*a+b against the exact plane distance: 2^-22 (|b| + 256 |a|)
        const float ex = (fabsf(bx) + 256.0f * fabsf(ax)) * 2.3841858e-7f;
        const float ey = (fabsf(by) + 256.0f * fabsf(ay)) * 2.3841858e-7f;
        const float ez = (fabsf(bz) + 256.0f * fabsf(az)) * 2.3841858e-7f;
        const float bxl = bx - ex, bxh = bx + ex, byl = by - ey, byh = by + ey, bzl = bz - ez, bzh = bz + ez;

        u32 hitmask = 0;
#pragma unroll
        for (int half = 0; half < 2; half++) {
            const u32 meta4 = half ? n1.w : n1.z;
            const u32 is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
            const u32 inner_mask4 = bytes_sign_extend(is_inner4 << 3);
            const u32 bit_index4 = (meta4 ^ (octinv4 & inner_mask4)) & 0x1F1F1F1Fu;
            const u32 child_bits4 = (meta4 >> 5) & 0x07070707u;
            const u32 qlx = half ? n2.y : n2.x, qly = half ? n2.w : n2.z, qlz = half ? n3.y : n3.x;
            const u32 qhx = half ? n3.w : n3.z, qhy = half ? n4.y : n4.x, qhz = half ? n4.w : n4.z;
            const u32 nearx = nx ? qhx : qlx, farx = nx ? qlx : qhx;
            const u32 neary = ny ? qhy : qly, fary = ny ? qly : qhy;
            const u32 nearz = nz ? qhz : qlz, farz = nz ? qlz : qhz;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float t0x = fmaf(byte_to_float(nearx, j), ax, bxl), t1x = fmaf(byte_to_float(farx, j), ax, bxh);
                const float t0y = fmaf(byte_to_float(neary, j), ay, byl), t1y = fmaf(byte_to_float(fary, j), ay, byh);
                const float t0z = fmaf(byte_to_float(nearz, j), az, bzl), t1z = fmaf(byte_to_float(farz, j), az, bzh);
                const float tmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.0f));
                const float tmax = fminf(fminf(t1x, t1y), fminf(t1z, h.t));
                if (tmin <= tmax) hitmask |= ((child_bits4 >> (8 * j)) & 0xFFu) << ((bit_index4 >> (8 * j)) & 0xFFu);
            }
        }
        ng.x = n1.x;
        ng.y = (hitmask & 0xFF000000u) | (n0.w >> 24);
        u32 tri_bits = hitmask & 0x00FFFFFFu;
        const u32 tri_base = n1.y;

        // ---- leaf primitives of this node ----
        while (tri_bits) {
            const u32 j = highest_bit(tri_bits);
            tri_bits &= ~(1u << j);
            const float4* tp = s.wide_tris + 3 * (size_t)(tri_base + j);
            const float4 v0 = PTB_LDG(&tp[0]), v1 = PTB_LDG(&tp[1]), v2 = PTB_LDG(&tp[2]);
            const u32 face = f2u(v0.w), shape = f2u(v1.w), kind = f2u(v2.w);
            if (shape != cur_shape) { obj = object_ray(s, r, shape); cur_shape = shape; }
            if (kind == 0u) test_triangle<true>(obj, mk3(v0.x, v0.y, v0.z), mk3(v1.x, v1.y, v1.z), mk3(v2.x, v2.y, v2.z), shape, face, h);
            else test_analytic<true>(obj, kind, shape, h);
        }

        if (ng.y <= 0x00FFFFFFu) {
            if (sp == 0) break;
            ng = stack[--sp];
        }
    }
    return h;
}

}  // namespace ptb
