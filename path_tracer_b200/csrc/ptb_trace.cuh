// Closest-hit queries: primitive tests, the reference-layout two-level BVH2
// traversal, and the closest-hit attribute fetch shared by both traversal
// backends (the wide-BVH backend is in ptb_wide.cuh).
//
// Reference behaviour: src/scene/scene.glsl.inc:304-611 (IntersectMeshFace,
// IntersectMeshNode, IntersectShape, Intersect, Trace) and the slab test of
// src/core/common.glsl.inc:153-185.  The primitive tests keep the reference's
// exact operation order (object-space ray from the un-normalised `From`
// transform, true divisions) so hit times are bit-identical; what changes is
// the memory side: 16-byte vector loads through the read-only path, both
// children of a mesh node pair fetched as one 64-byte segment.
#pragma once
#include "ptb_scene.cuh"

namespace ptb {

#define PTB_SHAPE_NONE 0xFFFFFFFFu
#define PTB_SHAPE_MESH 0u
#define PTB_SHAPE_PLANE 1u
#define PTB_SHAPE_SPHERE 2u
#define PTB_SHAPE_CUBE 3u

// Result of traversal: just enough to identify the closest primitive.
struct HitId {
    float t;
    u32 shape;        // PTB_SHAPE_NONE on miss
    u32 prim;         // global face index for meshes, 0 otherwise
    u32 shape_type;
    f3 coords;        // barycentrics (mesh) or object-space hit position
};

// Closest-hit attributes (scene.glsl.inc:102-119 minus the traversal counters).
struct Surface {
    f3 normal, tangent_x;
    f2 uv;
    u32 material;
};

// Traversal counters of the preview visualiser (scene.glsl.inc:117-118): loop iterations over shape
// nodes / mesh nodes in the reference's BVH2; node steps / primitive tests in the wide BVH.
struct Complexity { u32 scene, mesh; };

PTB_HD bool lower_id(u32 shape, u32 prim, const HitId& h) { return shape < h.shape || (shape == h.shape && prim < h.prim); }

// common.glsl.inc:153-185.  `canonical` visits boxes whose entry time equals the
// reach, so an equal-t primitive with a lower ID is still found.
template <bool CANONICAL>
PTB_HD float slab_entry(const Ray& r, float reach, f3 bmin, f3 bmax) {
    f3 t0 = (bmin - r.origin) / r.velocity;
    f3 t1 = (bmax - r.origin) / r.velocity;
    f3 lo = min3(t0, t1), hi = max3(t0, t1);
    float entry = fmaxf(fmaxf(lo.x, lo.y), lo.z);
    float exit = fminf(fminf(hi.x, hi.y), hi.z);
    if (exit < entry) return PTB_INFINITY;
    if (exit <= 0.0f) return PTB_INFINITY;
    if (CANONICAL ? (entry > reach) : (entry >= reach)) return PTB_INFINITY;
    return entry;
}

// scene.glsl.inc:304-334 — Moller-Trumbore, two-sided; equal t replaces
// (faithful) or replaces only with a lower ID (canonical).
template <bool CANONICAL>
PTB_HD void test_triangle(const Ray& r, f3 p0, f3 p1, f3 p2, u32 shape, u32 face, HitId& h) {
    f3 e1 = p1 - p0, e2 = p2 - p0;
    f3 pv = cross(r.velocity, e2);
    float det = dot(e1, pv);
    if (fabsf(det) < PTB_EPSILON) return;
    float inv = 1.0f / det;
    f3 sv = r.origin - p0;
    float u = inv * dot(sv, pv);
    if (u < 0.0f || u > 1.0f) return;
    f3 qv = cross(sv, e1);
    float v = inv * dot(r.velocity, qv);
    if (v < 0.0f || u + v > 1.0f) return;
    float t = inv * dot(e2, qv);
    if (t < 0.0f || t > h.t) return;
    if (CANONICAL && t == h.t && !lower_id(shape, face, h)) return;
    h.t = t;
    h.shape_type = PTB_SHAPE_MESH;
    h.shape = shape;
    h.prim = face;
    h.coords = mk3((1.0f - u) - v, u, v);
}

template <bool CANONICAL>
PTB_HD void test_face(const DeviceScene& s, const Ray& r, u32 face, u32 shape, HitId& h) {
    const float4* f = (const float4*)&s.mesh_faces[face];
    float4 a = PTB_LDG(&f[0]), b = PTB_LDG(&f[1]), c = PTB_LDG(&f[2]);
    test_triangle<CANONICAL>(r, mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(c.x, c.y, c.z), shape, face, h);
}

// World ray -> object ray (common.glsl.inc:84-91; direction not renormalised so
// object-space t equals world-space t).
PTB_HD Ray object_ray(const DeviceScene& s, const Ray& w, u32 shape) {
    const float4* m = (const float4*)&s.shapes[shape].Transform.From[0];
    float from[16];
#pragma unroll
    for (int c = 0; c < 4; c++) { float4 col = PTB_LDG(&m[c]); from[4 * c] = col.x; from[4 * c + 1] = col.y; from[4 * c + 2] = col.z; from[4 * c + 3] = col.w; }
    Ray o;
    o.origin = mat_mul_point(from, w.origin);
    o.velocity = mat_mul_vector(from, w.velocity);
    o.duration = w.duration;
    return o;
}

// Analytic shapes (scene.glsl.inc:413-465) on the object-space ray.
template <bool CANONICAL>
PTB_HD void test_analytic(const Ray& r, u32 type, u32 shape, HitId& h) {
    if (type == PTB_SHAPE_PLANE) {
        float t = -r.origin.z / r.velocity.z;
        if (t < 0.0f || t > h.t) return;
        if (CANONICAL && t == h.t && !lower_id(shape, 0, h)) return;
        h.t = t; h.shape_type = PTB_SHAPE_PLANE; h.shape = shape; h.prim = 0;
        h.coords = r.origin + r.velocity * t;
    } else if (type == PTB_SHAPE_SPHERE) {
        float v = dot(r.velocity, r.velocity);
        float p = dot(r.origin, r.velocity);
        float q = dot(r.origin, r.origin) - 1.0f;
        float d2 = p * p - q * v;
        if (d2 < 0.0f) return;
        float d = sqrtf(d2);
        if (d < p) return;
        float s0 = -p - d, s1 = -p + d;
        float sx = s0 < 0.0f ? s1 : s0;
        if (sx < 0.0f || sx > v * h.t) return;
        float t = sx / v;
        if (CANONICAL && t == h.t && !lower_id(shape, 0, h)) return;
        h.t = t; h.shape_type = PTB_SHAPE_SPHERE; h.shape = shape; h.prim = 0;
        h.coords = r.origin + r.velocity * t;
    } else if (type == PTB_SHAPE_CUBE) {
        f3 a = (mk3(-1, -1, -1) - r.origin) / r.velocity;
        f3 b = (mk3(+1, +1, +1) - r.origin) / r.velocity;
        f3 lo = min3(a, b), hi = max3(a, b);
        float t0 = fmaxf(fmaxf(lo.x, lo.y), lo.z);
        float t1 = fminf(fminf(hi.x, hi.y), hi.z);
        if (t1 < t0) return;
        if (t1 <= 0.0f) return;
        float t = t0 < 0.0f ? t1 : t0;
        if (!CANONICAL) { if (t >= h.t) return; }
        else { if (t > h.t) return; if (t == h.t && !lower_id(shape, 0, h)) return; }
        h.t = t; h.shape_type = PTB_SHAPE_CUBE; h.shape = shape; h.prim = 0;
        h.coords = r.origin + r.velocity * t;
    }
}

// ---- reference-layout BVH2 traversal -------------------------------------------
// Node = 32 B = two 16-byte loads: {min.xyz, word0}, {max.xyz, word1}.
struct Node2 { f3 bmin; u32 w0; f3 bmax; u32 w1; };
PTB_HD Node2 load_node2(const void* base, u32 index) {
    const float4* p = (const float4*)base + 2 * (size_t)index;
    float4 a = PTB_LDG(&p[0]), b = PTB_LDG(&p[1]);
    Node2 n;
    n.bmin = mk3(a.x, a.y, a.z); n.w0 = f2u(a.w);
    n.bmax = mk3(b.x, b.y, b.z); n.w1 = f2u(b.w);
    return n;
}

#define PTB_BVH2_STACK 64   // the reference has 32 with no guard (scene.glsl.inc:338,472); 64 + guard here

// scene.glsl.inc:336-399 — near child first, far child pushed; mesh children are adjacent.
template <bool CANONICAL>
PTB_HD void traverse_mesh_bvh2(const DeviceScene& s, const Ray& r, u32 root, u32 shape, HitId& h, Complexity* cx = nullptr) {
    u32 stack[PTB_BVH2_STACK];
    u32 depth = 0;
    Node2 node = load_node2(s.mesh_nodes, root);
    while (true) {
        if (cx) cx->mesh++;                                  // scene.glsl.inc:345
        if (node.w1 > 0) {
            for (u32 f = node.w0; f < node.w1; f++) test_face<CANONICAL>(s, r, f, shape, h);
        } else {
            u32 ia = node.w0, ib = ia + 1;
            node = load_node2(s.mesh_nodes, ia);
            Node2 nb = load_node2(s.mesh_nodes, ib);
            float ta = slab_entry<CANONICAL>(r, h.t, node.bmin, node.bmax);
            float tb = slab_entry<CANONICAL>(r, h.t, nb.bmin, nb.bmax);
            if (ta > tb) {
                if (ta < PTB_INFINITY && depth < PTB_BVH2_STACK) stack[depth++] = ia;
                node = nb;
                continue;
            }
            if (tb < PTB_INFINITY) {
                if (depth < PTB_BVH2_STACK) stack[depth++] = ib;
                continue;
            }
            if (ta < PTB_INFINITY) continue;
        }
        if (depth == 0) break;
        node = load_node2(s.mesh_nodes, stack[--depth]);
    }
}

template <bool CANONICAL>
PTB_HD void intersect_shape_bvh2(const DeviceScene& s, const Ray& world, u32 shape, HitId& h, Complexity* cx = nullptr) {
    const uint4 head = PTB_LDG((const uint4*)&s.shapes[shape]);   // type, material, mesh root, pad
    Ray r = object_ray(s, world, shape);
    if (head.x == PTB_SHAPE_MESH) traverse_mesh_bvh2<CANONICAL>(s, r, head.z, shape, h, cx);
    else test_analytic<CANONICAL>(r, head.x, shape, h);
}

// scene.glsl.inc:468-520 — top level over shapes; 16-bit child indices, leaf when 0.
template <bool CANONICAL>
PTB_HD HitId closest_hit_bvh2(const DeviceScene& s, const Ray& r, Complexity* cx = nullptr) {
    HitId h;
    h.t = r.duration; h.shape = PTB_SHAPE_NONE; h.prim = 0xFFFFFFFFu; h.shape_type = 0; h.coords = mk3(0, 0, 0);
    if (s.globals.ShapeCount == 0) return h;
    u32 stack[PTB_BVH2_STACK];
    u32 depth = 0;
    Node2 na = load_node2(s.shape_nodes, 0);
    while (true) {
        if (cx) cx->scene++;                                 // scene.glsl.inc:480
        if (na.w0 == 0) {
            intersect_shape_bvh2<CANONICAL>(s, r, na.w1, h, cx);
        } else {
            u32 ia = na.w0 & 0xFFFFu, ib = na.w0 >> 16;
            na = load_node2(s.shape_nodes, ia);
            Node2 nb = load_node2(s.shape_nodes, ib);
            float ta = slab_entry<CANONICAL>(r, h.t, na.bmin, na.bmax);
            float tb = slab_entry<CANONICAL>(r, h.t, nb.bmin, nb.bmax);
            if (ta > tb) {
                if (ta < PTB_INFINITY && depth < PTB_BVH2_STACK) stack[depth++] = ia;
                na = nb;
                continue;
            }
            if (tb < PTB_INFINITY) {
                if (depth < PTB_BVH2_STACK) stack[depth++] = ib;
                continue;
            }
            if (ta < PTB_INFINITY) continue;
        }
        if (depth == 0) break;
        na = load_node2(s.shape_nodes, stack[--depth]);
    }
    return h;
}

// ---- closest-hit attributes (scene.glsl.inc:535-608) -----------------------------
// TransformNormal = normalize((vec4(N,0) * From).xyz); TransformDirection = normalize((To * vec4(D,0)).xyz)
PTB_HD Surface finish_hit(const DeviceScene& s, const HitId& h) {
    Surface out;
    const ptb_packed_shape* sh = &s.shapes[h.shape];
    out.material = PTB_LDG(&sh->MaterialIndex);
    float to[16], from[16];
    {
        const float4* m = (const float4*)&sh->Transform.To[0];
#pragma unroll
        for (int c = 0; c < 4; c++) { float4 col = PTB_LDG(&m[c]); to[4 * c] = col.x; to[4 * c + 1] = col.y; to[4 * c + 2] = col.z; to[4 * c + 3] = col.w; }
#pragma unroll
        for (int c = 0; c < 4; c++) { float4 col = PTB_LDG(&m[4 + c]); from[4 * c] = col.x; from[4 * c + 1] = col.y; from[4 * c + 2] = col.z; from[4 * c + 3] = col.w; }
    }
    if (h.shape_type == PTB_SHAPE_MESH) {
        uint2 v0, v1, v2;
        if (s.face_attrs) {
            const uint4 a = PTB_LDG(&s.face_attrs[2 * (size_t)h.prim]), b = PTB_LDG(&s.face_attrs[2 * (size_t)h.prim + 1]);
            v0 = make_uint2(a.x, a.y); v1 = make_uint2(a.z, a.w); v2 = make_uint2(b.x, b.y);
        } else {
            const ptb_packed_mesh_face* f = &s.mesh_faces[h.prim];
            u32 i0 = PTB_LDG(&f->VertexIndex0), i1 = PTB_LDG(&f->VertexIndex1), i2 = PTB_LDG(&f->VertexIndex2);
            v0 = PTB_LDG((const uint2*)&s.mesh_vertices[i0]);
            v1 = PTB_LDG((const uint2*)&s.mesh_vertices[i1]);
            v2 = PTB_LDG((const uint2*)&s.mesh_vertices[i2]);
        }
        f3 c = h.coords;
        f3 n = safe_normalize((unpack_unit_vector(v0.x) * c.x + unpack_unit_vector(v1.x) * c.y) + unpack_unit_vector(v2.x) * c.z);
        out.normal = normalize(vector_mul_mat(n, from));
        out.tangent_x = tangent_vector(out.normal);
        f2 a = unpack_half2x16(v0.y), b = unpack_half2x16(v1.y), d = unpack_half2x16(v2.y);
        out.uv = (a * c.x + b * c.y) + d * c.z;
    } else if (h.shape_type == PTB_SHAPE_PLANE) {
        out.normal = normalize(vector_mul_mat(mk3(0, 0, 1), from));
        out.tangent_x = normalize(mat_mul_vector(to, mk3(1, 0, 0)));
        out.uv = mk2(gfract(h.coords.x), gfract(h.coords.y));
    } else if (h.shape_type == PTB_SHAPE_SPHERE) {
        f3 p = h.coords;
        float u = (atan2f(p.y, p.x) + PTB_PI) / PTB_TAU;
        float v = (p.z + 1.0f) / 2.0f;
        out.normal = normalize(vector_mul_mat(p, from));
        out.tangent_x = normalize(mat_mul_vector(to, cross(p, mk3(-p.y, p.x, 0.0f))));
        out.uv = mk2(u, v);
    } else {
        f3 p = h.coords, q = abs3(p);
        f3 n, tx;
        if (q.x >= q.y && q.x >= q.z) {
            float sg = gsign(p.x);
            n = mk3(sg, 0, 0); tx = mk3(0, sg, 0);
            out.uv = mk2(0.5f * (1.0f + p.y), 0.5f * (1.0f + p.z));
        } else if (q.y >= q.x && q.y >= q.z) {
            float sg = gsign(p.y);
            n = mk3(0, sg, 0); tx = mk3(0, 0, sg);
            out.uv = mk2(0.5f * (1.0f + p.x), 0.5f * (1.0f + p.z));
        } else {
            float sg = gsign(p.z);
            n = mk3(0, 0, sg); tx = mk3(sg, 0, 0);
            out.uv = mk2(0.5f * (1.0f + p.x), 0.5f * (1.0f + p.y));
        }
        out.normal = normalize(vector_mul_mat(n, from));
        out.tangent_x = normalize(mat_mul_vector(to, tx));
    }
    return out;
}

}  // namespace ptb
