// Scene flattening on the GPU: the same compressed 8-wide BVH as ptb_wide_build.cpp
// (layout in ptb_wide.cuh), built by CUDA kernels from the packed arrays that
// ptb_scene_update has already uploaded.  This is the fast builder for large or
// frequently edited scenes (the reference's build is the
// single-threaded binned-SAH recursion of src/scene/scene.cpp:435-599 and is the
// dominant cost of "time to first image" at 10 M triangles):
//
//   world-space primitive bounds -> 63-bit Morton keys -> radix sort -> binary tree
//   (LBVH: the radix tree of Karras 2012 + bottom-up bounds;  PLOC: locally-ordered
//   agglomerative clustering, Meister & Bittner 2018) with the SAH-optimal collapse
//   costs computed bottom-up -> collapse to 8-wide, level by level -> octant slot
//   assignment + 8-bit outward-rounded quantisation.
//
// The tree is a culling structure only: leaf primitives carry the reference's
// object-space vertices, face/shape indices and the canonical rank tag, so hits
// (t, IDs) are bit-identical to the host-built SAH tree's and to the oracle's,
// whatever the tree shape.  A Morton-order tree is cheaper to build (milliseconds
// instead of seconds) and slower to traverse; ptb_scene_set_builder() chooses.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "ptb.h"

namespace ptb {

struct GpuWideBuild {
    uint32_t* nodes = nullptr;       // device, 20 words per node (exact size)
    float*    prims = nullptr;       // device, 12 floats per leaf primitive (exact size)
    std::vector<uint32_t> planes;    // host: {shape, tag} per unbounded shape
    uint32_t node_count = 0, prim_count = 0, depth = 0, bvh2_depth = 0, rounds = 0;
    double   sah_cost = 0.0;
    double   build_seconds = 0.0;    // host wall clock around the whole build (including the final synchronise)
    float    device_ms = 0.0f;       // CUDA events around the kernels
};

// algorithm: 0 = Morton-order radix tree (LBVH), 1 = PLOC on the Morton order (better trees, more rounds).
// d_shapes / d_faces: the device copies of view->Shapes / view->MeshFaces.  Work is enqueued on `st`
// and synchronised before returning.  Returns 0, or a negative PTB_E_* with *err filled; on failure
// nothing is left allocated.
int build_wide_bvh_gpu(const ptb_scene_view* view, const ptb_packed_shape* d_shapes, const ptb_packed_mesh_face* d_faces,
                       cudaStream_t st, int algorithm, GpuWideBuild* out, std::string* err);

}  // namespace ptb
