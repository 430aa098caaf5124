// Private definitions shared by the translation units that implement the C ABI
// (ptb_api.cu: contexts, scene, film, renderer; ptb_group.cu: multi-GPU groups).
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "ptb.h"
#include "ptb_kernels.h"

namespace ptb {
int api_fail(int code, const std::string& msg);
int api_cuda_fail(cudaError_t e, const char* what);
void api_set_error(const std::string& msg);
const std::string& api_error();
}
#define CU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return ptb::api_cuda_fail(e_, #x); } while (0)

namespace ptb {
template <class T>
struct DevBuf {
    T* p = nullptr; size_t n = 0;
    cudaError_t upload(const T* host, size_t count, cudaStream_t st) {
        if (count != n || !p) {
            if (p) cudaFree(p);
            p = nullptr; n = 0;
            if (count) { cudaError_t e = cudaMalloc((void**)&p, count * sizeof(T)); if (e != cudaSuccess) return e; }
            n = count;
        }
        if (count) return cudaMemcpyAsync(p, host, count * sizeof(T), cudaMemcpyHostToDevice, st);
        return cudaSuccess;
    }
    cudaError_t alloc(size_t count) {
        release();
        if (count) { cudaError_t e = cudaMalloc((void**)&p, count * sizeof(T)); if (e != cudaSuccess) return e; }
        n = count;
        return cudaSuccess;
    }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
};
}  // namespace ptb

struct ptb_context { int device; cudaStream_t stream; unsigned* counter; };

// PTB_BUILDER_AUTO: the tree built on the GPU is in use at once; a host thread builds the SAH tree from the device
// copies of the packed arrays and parks it here until the next call that uses the scene swaps it in.  Any tree gives
// bit-identical hits (leaf primitives and the canonical rank are the same), so the swap is invisible in the results.
struct TreeUpgrade {
    std::thread th;
    std::mutex mu;                       // device copies of shapes / faces / mesh nodes vs. re-upload; the hand-over below
    bool running = false;                // under mu
    uint64_t wanted = 0;                 // under mu: generation of the scene geometry (bumped by EVERY geometry change)
    uint64_t requested = 0;              // under mu: newest generation a host build was asked for (0 = none)
    std::atomic<int> ready{0};           // 1: nodes/prims below hold the tree of generation `wanted`
    std::atomic<bool> quit{false};
    uint32_t* nodes = nullptr; float* prims = nullptr;
    uint32_t node_count = 0, prim_count = 0, depth = 0, bvh2_depth = 0;
    double sah_cost = 0.0, build_seconds = 0.0;
};

struct ptb_scene {
    ptb_context* ctx;
    TreeUpgrade* upgrade;
    ptb::DevBuf<ptb_packed_texture> textures;
    ptb::DevBuf<uint32_t> materials;
    ptb::DevBuf<ptb_packed_shape> shapes;
    ptb::DevBuf<ptb_packed_shape_node> shape_nodes;
    ptb::DevBuf<ptb_packed_mesh_face> faces;
    ptb::DevBuf<ptb_packed_mesh_vertex> vertices;
    ptb::DevBuf<ptb_packed_mesh_node> mesh_nodes;
    ptb::DevBuf<uint4> face_attrs;   // per-face vertex attributes gathered on the device (DeviceScene::face_attrs)
    ptb::DevBuf<ptb_packed_camera> cameras;
    std::vector<ptb::DevBuf<float>> layers;
    ptb::DevBuf<const float*> layer_table;
    ptb::DevBuf<uint32_t> wide_nodes;
    ptb::DevBuf<float> wide_prims;
    ptb::DevBuf<uint32_t> wide_planes;
    ptb::DeviceScene ds;
    ptb_wide_info wide_info;
    unsigned raycast_variant;
    uint32_t builder;                // PTB_BUILDER_*
    uint32_t material_features;      // PTB_FEAT_DIFFUSE/METAL/TRANSLUCENT bits of the uploaded material slots
};

// (Re)builds s->face_attrs from the device copies of faces and vertices; on an allocation failure the scene keeps working
// through the indexed path (ds.face_attrs = null).  Enqueued on the context's stream.
extern "C" int ptb_scene_build_face_attrs(ptb_scene* s);

struct ptb_film {
    ptb_context* ctx;
    uint32_t width, height;
    float4* accum; bool external;
    uchar4* srgb; float* linear;
};

struct ptb_renderer {
    ptb_context* ctx; ptb_scene* scene; ptb_film* film;
    ptb_render_params params;
    uint32_t frame_index;
    ptb::RenderConsts pc;
    ptb::WavefrontState st;
    void* state_block;
    unsigned* queue[2];           // material queues of the scatter step, one per pipeline (allocated on first use)
    cudaStream_t aux;             // second pipeline
    cudaEvent_t ev_fork, ev_join;
    uint64_t rays_traced, kernel_launches;
};

