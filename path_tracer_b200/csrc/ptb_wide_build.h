// Scene flattening: packed two-level scene (reference layout) -> one compressed
// 8-wide BVH over world space (host side, runs inside ptb_scene_update).
//
// The reference traverses a top-level BVH2 over shapes and one BVH2 per mesh
// (src/scene/scene.cpp:435-599,1402-1492; src/scene/scene.glsl.inc:336-520).
// B200 has no RT cores, so the integrator's own traversal structure is chosen for
// SM-side software traversal instead: every bounded primitive of every shape
// (mesh faces, spheres, cubes) becomes a leaf primitive of ONE BVH8 with
// 80-byte quantised nodes (layout in ptb_wide.cuh); unbounded shapes (planes)
// are kept in a short list that is tested before traversal.  Leaf primitives
// keep the reference's object-space vertex positions and its global face index,
// so the exact primitive tests — and therefore hit times and IDs — are the
// reference's.
#pragma once
#include <stdint.h>
#include <string>
#include <vector>
#include "ptb.h"

namespace ptb {

struct WideBuild {
    std::vector<uint32_t> nodes;     // 20 words (80 B) per node; node 0 is the root
    std::vector<float>    prims;     // 12 floats (48 B) per leaf primitive
    std::vector<uint32_t> planes;    // {shape index, tag} per unbounded shape
    uint32_t node_count = 0, prim_count = 0, depth = 0, bvh2_depth = 0;
    double   sah_cost = 0.0;         // SAH cost of the wide tree (node area-weighted), for diagnostics
    double   build_seconds = 0.0;
};

// Returns 0 on success; negative PTB_E_* with *err filled otherwise.
int build_wide_bvh(const ptb_scene_view* view, WideBuild* out, std::string* err);

}  // namespace ptb
