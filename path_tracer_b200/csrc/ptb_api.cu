// C ABI of the integrator (include/ptb.h): contexts, scene upload + wide-BVH
// flattening, film, renderer (reset / run / resolve) and the ray-cast harness.
// Replaces the host side of src/integrator/{basic,integrator}.cpp and the
// device-upload half of src/scene/scene.cpp:1643-2006.  There is no CPU
// execution path: without a usable CUDA device every entry point fails.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <atomic>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "ptb_internal.h"
#include "ptb_preview.cuh"
#include "ptb_wide_build.h"
#include "ptb_gpu_build.h"

using namespace ptb;

namespace {
thread_local std::string g_error;
int fail(int code, const std::string& msg) { g_error = msg; return code; }
int cuda_fail(cudaError_t e, const char* what) {
    g_error = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    return PTB_E_CUDA;
}
}  // namespace

namespace ptb {     // for the other translation units of the library (ptb_spectrum.cu, ptb_group.cu)
int api_fail(int code, const std::string& msg) { return fail(code, msg); }
int api_cuda_fail(cudaError_t e, const char* what) { return cuda_fail(e, what); }
void api_set_error(const std::string& msg) { g_error = msg; }
const std::string& api_error() { return g_error; }
cudaStream_t api_stream(ptb_context* ctx) { return ctx->stream; }
int api_device(ptb_context* ctx) { return ctx->device; }
}

static void upgrade_thread(ptb_scene* s) {
    TreeUpgrade* u = s->upgrade;
    cudaSetDevice(s->ctx->device);
    for (;;) {
        uint64_t gen;
        std::vector<ptb_packed_shape> shapes;
        std::vector<ptb_packed_mesh_face> faces;
        std::vector<ptb_packed_mesh_node> nodes;
        bool ok = true;
        {
            std::lock_guard<std::mutex> l(u->mu);
            gen = u->wanted;
            shapes.resize(s->shapes.n); faces.resize(s->faces.n); nodes.resize(s->mesh_nodes.n);
            ok = ok && (shapes.empty() || cudaMemcpy(shapes.data(), s->shapes.p, shapes.size() * sizeof shapes[0], cudaMemcpyDeviceToHost) == cudaSuccess);
            ok = ok && (faces.empty() || cudaMemcpy(faces.data(), s->faces.p, faces.size() * sizeof faces[0], cudaMemcpyDeviceToHost) == cudaSuccess);
            ok = ok && (nodes.empty() || cudaMemcpy(nodes.data(), s->mesh_nodes.p, nodes.size() * sizeof nodes[0], cudaMemcpyDeviceToHost) == cudaSuccess);
        }
        WideBuild wb;
        std::string err;
        if (ok) {
            ptb_scene_view v;
            std::memset(&v, 0, sizeof v);
            v.Shapes = shapes.data(); v.ShapeCount = (uint32_t)shapes.size();
            v.MeshFaces = faces.data(); v.MeshFaceCount = (uint32_t)faces.size();
            v.MeshNodes = nodes.data(); v.MeshNodeCount = (uint32_t)nodes.size();
            ok = build_wide_bvh(&v, &wb, &err) == PTB_OK;
        }
        uint32_t* d_nodes = nullptr; float* d_prims = nullptr;
        if (ok && !u->quit.load() && wb.node_count) {
            ok = cudaMalloc((void**)&d_nodes, wb.nodes.size() * 4) == cudaSuccess && cudaMalloc((void**)&d_prims, wb.prims.size() * 4) == cudaSuccess
              && cudaMemcpy(d_nodes, wb.nodes.data(), wb.nodes.size() * 4, cudaMemcpyHostToDevice) == cudaSuccess
              && cudaMemcpy(d_prims, wb.prims.data(), wb.prims.size() * 4, cudaMemcpyHostToDevice) == cudaSuccess;
        }
        std::lock_guard<std::mutex> l(u->mu);
        if (ok && !u->quit.load() && u->wanted == gen && u->requested == gen && wb.node_count) {
            u->nodes = d_nodes; u->prims = d_prims;
            u->node_count = wb.node_count; u->prim_count = wb.prim_count; u->depth = wb.depth; u->bvh2_depth = wb.bvh2_depth;
            u->sah_cost = wb.sah_cost; u->build_seconds = wb.build_seconds;
            u->ready.store(1);
            u->running = false;
            return;
        }
        if (d_nodes) cudaFree(d_nodes);
        if (d_prims) cudaFree(d_prims);
        // failed, nothing to build, or the geometry changed and nobody asked for a host tree of the new one
        // (another builder was selected, or the GPU build fell back to the host builder): the tree in use stays
        if (u->quit.load() || !ok || u->wanted == gen || u->requested != u->wanted) { u->running = false; return; }
        // the geometry changed while this tree was being built and a tree of the new one is wanted: build it
    }
}

// Every geometry change (whatever the builder, whether or not a GPU build follows): trees of the old geometry that
// are parked or still being built must never be installed.  Caller holds u->mu.
static void upgrade_invalidate_locked(TreeUpgrade* u) {
    u->wanted++;
    if (u->ready.exchange(0)) { cudaFree(u->nodes); cudaFree(u->prims); u->nodes = nullptr; u->prims = nullptr; }
}

// Called (host thread of the API) after a GPU build has been installed for new geometry.
static void upgrade_request(ptb_scene* s) {
    if (!s->upgrade) s->upgrade = new TreeUpgrade();
    TreeUpgrade* u = s->upgrade;
    bool start = false;
    {
        std::lock_guard<std::mutex> l(u->mu);
        if (u->wanted == 0) u->wanted = 1;
        u->requested = u->wanted;
        if (!u->running) { u->running = true; start = true; }
    }
    if (start) {
        if (u->th.joinable()) u->th.join();
        u->th = std::thread(upgrade_thread, s);
    }
}

// Swaps a finished host-built tree in.  Called at the top of every entry point that traverses the scene.
static int upgrade_poll(ptb_scene* s) {
    TreeUpgrade* u = s->upgrade;
    if (!u || !u->ready.load()) return PTB_OK;
    CU(cudaSetDevice(s->ctx->device));
    CU(cudaStreamSynchronize(s->ctx->stream));          // no kernel may still be walking the old tree
    std::lock_guard<std::mutex> l(u->mu);
    if (!u->ready.exchange(0)) return PTB_OK;
    s->wide_nodes.release(); s->wide_prims.release();
    s->wide_nodes.p = u->nodes; s->wide_nodes.n = (size_t)u->node_count * 20;
    s->wide_prims.p = u->prims; s->wide_prims.n = (size_t)u->prim_count * 12;
    u->nodes = nullptr; u->prims = nullptr;
    s->ds.wide_nodes = (const uint4*)s->wide_nodes.p;
    s->ds.wide_tris = (const float4*)s->wide_prims.p;
    s->wide_info.node_count = u->node_count; s->wide_info.prim_count = u->prim_count;
    s->wide_info.depth = u->depth; s->wide_info.bvh2_depth = u->bvh2_depth;
    s->wide_info.sah_cost = u->sah_cost; s->wide_info.build_seconds = u->build_seconds;
    s->wide_info.builder = PTB_BUILDER_SAH_HOST; s->wide_info.device_build_ms = 0.0f;
    return PTB_OK;
}

static void upgrade_shutdown(ptb_scene* s) {
    TreeUpgrade* u = s->upgrade;
    if (!u) return;
    u->quit.store(true);
    if (u->th.joinable()) u->th.join();
    if (u->nodes) cudaFree(u->nodes);
    if (u->prims) cudaFree(u->prims);
    delete u;
    s->upgrade = nullptr;
}

// PTB_SCATTER_AUTO: material queues pay when a warp would otherwise meet several materials' code.
static bool use_queues(const ptb_renderer* r) {
    if (r->params.ScatterMode == PTB_SCATTER_QUEUES) return true;
    if (r->params.ScatterMode == PTB_SCATTER_DIRECT || r->params.ScatterMode == PTB_SCATTER_DIRECT_STAGED) return false;
    const uint32_t f = r->scene->ds.features;
    const uint32_t types = f & (PTB_FEAT_DIFFUSE | PTB_FEAT_METAL | PTB_FEAT_TRANSLUCENT);
    return (types & (types - 1u)) != 0u || (f & (PTB_FEAT_TRANSLUCENT | PTB_FEAT_SCENE_SCATTER)) != 0u;
}
static int scatter_step(ptb_renderer* r, const RenderConsts& pc, const WavefrontState& wst, unsigned pipe, cudaStream_t st) {
    if (!use_queues(r)) { CU(launch_scatter(r->scene->ds, pc, wst, r->film->accum, r->params.ScatterMode == PTB_SCATTER_DIRECT_STAGED, st)); r->kernel_launches += 1; return PTB_OK; }
    if (!r->queue[pipe]) {
        cudaError_t e = cudaMalloc((void**)&r->queue[pipe], scatter_queue_bytes(r->pc.slot_count));
        if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc(material queues)");
    }
    CU(launch_scatter_queues(r->scene->ds, pc, wst, r->film->accum, r->queue[pipe], st));
    r->kernel_launches += scatter_queue_launches(r->scene->ds);
    return PTB_OK;
}
// Pipelines: the slot range of the film split at a tile boundary; each half has its own ray counter.
static unsigned pipeline_count(const ptb_renderer* r) {
    if (r->params.Pipelines == 1u || r->pc.slot_count < 512u) return 1u;
    // automatic = one: on B200 the two-stream split measured 3.5 % slower on C2/C3 and neutral on C4
    // (profiles/r01_v6_pipelines.txt): the persistent trace kernel of the other half takes the SMs, so only the
    // kernel tails overlap, and that does not pay for two half-sized launches per kernel.
    return r->params.Pipelines == 2u ? 2u : 1u;
}

extern "C" {

const char* ptb_last_error(void) { return g_error.c_str(); }
const char* ptb_version(void) { return "path_tracer_b200 0.1 (sm_100a)"; }

int ptb_context_create(int cuda_device, ptb_context** out) {
    if (!out) return fail(PTB_E_INVALID, "ptb_context_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount");
    if (cuda_device < 0 || cuda_device >= count) return fail(PTB_E_CUDA, "ptb_context_create: no such CUDA device (there is no CPU fallback)");
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, cuda_device));
    if (prop.major != 10) return fail(PTB_E_UNSUPPORTED, std::string("ptb_context_create: kernels are built for sm_100a only, device is ") + prop.name);
    CU(cudaSetDevice(cuda_device));
    ptb_context* c = new ptb_context();
    c->device = cuda_device;
    e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { delete c; return cuda_fail(e, "cudaStreamCreate"); }
    e = cudaMalloc((void**)&c->counter, 64);
    if (e != cudaSuccess) { cudaStreamDestroy(c->stream); delete c; return cuda_fail(e, "cudaMalloc(counter)"); }
    *out = c;
    return PTB_OK;
}
void ptb_context_destroy(ptb_context* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaStreamDestroy(ctx->stream);
    cudaFree(ctx->counter);
    delete ctx;
}
int ptb_context_sync(ptb_context* ctx) {
    if (!ctx) return fail(PTB_E_INVALID, "ptb_context_sync: NULL context");
    CU(cudaStreamSynchronize(ctx->stream));
    return PTB_OK;
}
void* ptb_context_stream(ptb_context* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

// ---- scene ----
int ptb_scene_create(ptb_context* ctx, ptb_scene** out) {
    if (!ctx || !out) return fail(PTB_E_INVALID, "ptb_scene_create: NULL argument");
    ptb_scene* s = new ptb_scene();
    s->ctx = ctx;
    s->upgrade = nullptr;
    std::memset(&s->ds, 0, sizeof s->ds);
    std::memset(&s->wide_info, 0, sizeof s->wide_info);
    s->ds.wide_tlas_root = 0xFFFFFFFFu;
    s->ds.globals.SkyboxTextureIndex = 0xFFFFFFFFu;
    s->raycast_variant = PTB_TRACE_VARIANT_DEFAULT;
    s->builder = PTB_BUILDER_SAH_HOST;
    if (const char* e = std::getenv("PTB_BUILDER")) s->builder = std::atoi(e) == 1 ? PTB_BUILDER_LBVH_GPU : std::atoi(e) == 2 ? PTB_BUILDER_PLOC_GPU : std::atoi(e) == 3 ? PTB_BUILDER_AUTO : PTB_BUILDER_SAH_HOST;
    s->material_features = 0;
    *out = s;
    return PTB_OK;
}

int ptb_scene_update(ptb_scene* s, const ptb_scene_view* v, uint32_t dirty) {
    if (!s || !v) return fail(PTB_E_INVALID, "ptb_scene_update: NULL argument");
    cudaStream_t st = s->ctx->stream;
    CU(cudaSetDevice(s->ctx->device));
    CU(cudaStreamSynchronize(st));      // the reference does vkDeviceWaitIdle here (scene.cpp:1704)
    if (dirty & PTB_SCENE_DIRTY_TEXTURES) {
        CU(s->textures.upload(v->Textures, v->TextureCount, st));
        for (auto& l : s->layers) l.release();
        s->layers.assign(v->AtlasLayerCount, DevBuf<float>());
        std::vector<const float*> table(v->AtlasLayerCount);
        const size_t texels = (size_t)v->AtlasWidth * v->AtlasHeight * 4;
        for (uint32_t i = 0; i < v->AtlasLayerCount; i++) {
            CU(s->layers[i].upload(v->AtlasLayers[i], texels, st));
            table[i] = s->layers[i].p;
        }
        CU(s->layer_table.upload(table.data(), table.size(), st));
        CU(cudaStreamSynchronize(st));   // `table` is a stack object
        s->ds.textures = s->textures.p; s->ds.texture_count = v->TextureCount;
        s->ds.atlas_layers = s->layer_table.p; s->ds.atlas_width = v->AtlasWidth; s->ds.atlas_height = v->AtlasHeight;
    }
    if (dirty & PTB_SCENE_DIRTY_MATERIALS) {
        CU(s->materials.upload(v->MaterialData, v->MaterialWordCount, st));
        s->ds.material_data = s->materials.p; s->ds.material_word_count = v->MaterialWordCount;
        s->material_features = 0;
        for (uint32_t slot = 0; slot + 1 <= v->MaterialWordCount / 32u; slot++) {
            const uint32_t type = v->MaterialData[32u * slot];          // word 0 = material type (scene.glsl.inc:687-762)
            s->material_features |= type == 0u ? PTB_FEAT_DIFFUSE : type == 1u ? PTB_FEAT_METAL : type == 2u ? PTB_FEAT_TRANSLUCENT : 0u;
        }
    }
    // a background tree build reads the device copies of shapes / faces / mesh nodes
    std::unique_lock<std::mutex> geometry_lock;
    if (s->upgrade && (dirty & (PTB_SCENE_DIRTY_SHAPES | PTB_SCENE_DIRTY_MESHES))) {
        geometry_lock = std::unique_lock<std::mutex>(s->upgrade->mu);
        upgrade_invalidate_locked(s->upgrade);
    }
    if (dirty & PTB_SCENE_DIRTY_SHAPES) {
        CU(s->shapes.upload(v->Shapes, v->ShapeCount, st));
        CU(s->shape_nodes.upload(v->ShapeNodes, v->ShapeNodeCount, st));
        s->ds.shapes = s->shapes.p; s->ds.shape_count = v->ShapeCount;
        s->ds.shape_nodes = s->shape_nodes.p; s->ds.shape_node_count = v->ShapeNodeCount;
    }
    if (dirty & PTB_SCENE_DIRTY_MESHES) {
        CU(s->faces.upload(v->MeshFaces, v->MeshFaceCount, st));
        CU(s->vertices.upload(v->MeshVertices, v->MeshVertexCount, st));
        CU(s->mesh_nodes.upload(v->MeshNodes, v->MeshNodeCount, st));
        s->ds.mesh_faces = s->faces.p; s->ds.mesh_face_count = v->MeshFaceCount;
        s->ds.mesh_vertices = s->vertices.p;
        s->ds.mesh_nodes = s->mesh_nodes.p; s->ds.mesh_node_count = v->MeshNodeCount;
        { int rc = ptb_scene_build_face_attrs(s); if (rc) return rc; }
    }
    if (geometry_lock.owns_lock()) {
        CU(cudaStreamSynchronize(st));
        geometry_lock.unlock();
    }
    if (dirty & PTB_SCENE_DIRTY_CAMERAS) {
        CU(s->cameras.upload(v->Cameras, v->CameraCount, st));
        s->ds.cameras = s->cameras.p; s->ds.camera_count = v->CameraCount;
    }
    if ((dirty & (PTB_SCENE_DIRTY_GLOBALS | PTB_SCENE_DIRTY_SHAPES | PTB_SCENE_DIRTY_SKYBOX_TEXTURE)) && v->Globals) s->ds.globals = *v->Globals;
    // what scatter can meet in this scene: selects the k_scatter instantiation (ptb_kernels.cu)
    s->ds.features = s->material_features | (s->ds.texture_count ? PTB_FEAT_TEXTURES : 0u)
                   | (s->ds.globals.SkyboxTextureIndex != 0xFFFFFFFFu ? PTB_FEAT_SKY_TEXTURE : 0u)
                   | (s->ds.globals.SceneScatterRate > 0.0f ? PTB_FEAT_SCENE_SCATTER : 0u);
    if (dirty & (PTB_SCENE_DIRTY_SHAPES | PTB_SCENE_DIRTY_MESHES)) {
        // scene flattening: packed two-level BVH2 scene -> one compressed BVH8
        std::string err;
        int rc = PTB_E_UNSUPPORTED;
        bool built = false;
        const bool gpu_builder = s->builder == PTB_BUILDER_LBVH_GPU || s->builder == PTB_BUILDER_PLOC_GPU || s->builder == PTB_BUILDER_AUTO;
        if (gpu_builder) {
            // needs the device copies of the shapes, faces and (for the background host build) mesh nodes; an update that
            // did not carry them finds them resident from the update that did - re-sent only when they are missing
            {
                std::unique_lock<std::mutex> l;
                if (s->upgrade) l = std::unique_lock<std::mutex>(s->upgrade->mu);
                if (!(dirty & PTB_SCENE_DIRTY_SHAPES) && (s->shapes.n != v->ShapeCount || !s->shapes.p)) CU(s->shapes.upload(v->Shapes, v->ShapeCount, st));
                if (!(dirty & PTB_SCENE_DIRTY_MESHES) && (s->faces.n != v->MeshFaceCount || !s->faces.p)) CU(s->faces.upload(v->MeshFaces, v->MeshFaceCount, st));
                if (!(dirty & PTB_SCENE_DIRTY_MESHES) && (s->mesh_nodes.n != v->MeshNodeCount || !s->mesh_nodes.p)) CU(s->mesh_nodes.upload(v->MeshNodes, v->MeshNodeCount, st));
                CU(cudaStreamSynchronize(st));
            }
            GpuWideBuild gb;
            rc = build_wide_bvh_gpu(v, s->shapes.p, s->faces.p, st, s->builder == PTB_BUILDER_LBVH_GPU ? 0 : 1, &gb, &err);
            if (rc == PTB_OK) {
                s->wide_nodes.release(); s->wide_prims.release();
                s->wide_nodes.p = gb.nodes; s->wide_nodes.n = (size_t)gb.node_count * 20;
                s->wide_prims.p = gb.prims; s->wide_prims.n = (size_t)gb.prim_count * 12;
                CU(s->wide_planes.upload(gb.planes.data(), gb.planes.size(), st));
                CU(cudaStreamSynchronize(st));
                s->ds.wide_plane_count = (uint32_t)(gb.planes.size() / 2);
                s->ds.wide_tlas_root = gb.node_count ? 0u : 0xFFFFFFFFu;
                s->wide_info.node_count = gb.node_count; s->wide_info.prim_count = gb.prim_count;
                s->wide_info.plane_count = s->ds.wide_plane_count;
                s->wide_info.depth = gb.depth; s->wide_info.bvh2_depth = gb.bvh2_depth;
                s->wide_info.sah_cost = gb.sah_cost; s->wide_info.build_seconds = gb.build_seconds;
                s->wide_info.builder = s->builder == PTB_BUILDER_LBVH_GPU ? PTB_BUILDER_LBVH_GPU : PTB_BUILDER_PLOC_GPU;
                s->wide_info.device_build_ms = gb.device_ms;
                built = true;
                if (s->builder == PTB_BUILDER_AUTO && gb.node_count) upgrade_request(s);
            }
        }
        if (!built) {
            WideBuild wb;
            rc = build_wide_bvh(v, &wb, &err);
            if (rc == PTB_OK) {
                CU(s->wide_nodes.upload(wb.nodes.data(), wb.nodes.size(), st));
                CU(s->wide_prims.upload(wb.prims.data(), wb.prims.size(), st));
                CU(s->wide_planes.upload(wb.planes.data(), wb.planes.size(), st));
                CU(cudaStreamSynchronize(st));
                s->ds.wide_plane_count = (uint32_t)(wb.planes.size() / 2);
                s->ds.wide_tlas_root = wb.node_count ? 0u : 0xFFFFFFFFu;
                s->wide_info.node_count = wb.node_count; s->wide_info.prim_count = wb.prim_count;
                s->wide_info.plane_count = (uint32_t)(wb.planes.size() / 2);
                s->wide_info.depth = wb.depth; s->wide_info.bvh2_depth = wb.bvh2_depth;
                s->wide_info.sah_cost = wb.sah_cost; s->wide_info.build_seconds = wb.build_seconds;
                s->wide_info.builder = PTB_BUILDER_SAH_HOST; s->wide_info.device_build_ms = 0.0f;
            }
        }
        s->wide_info.status = rc;
        if (rc == PTB_OK) {
            s->ds.wide_nodes = (const uint4*)s->wide_nodes.p;
            s->ds.wide_tris = (const float4*)s->wide_prims.p;
            s->ds.wide_planes = s->wide_planes.p;
        } else {
            // The reference-layout traversal stays usable; the wide backend reports the reason.
            s->ds.wide_nodes = nullptr; s->ds.wide_tlas_root = 0xFFFFFFFFu; s->ds.wide_plane_count = 0;
            g_error = "wide BVH build failed: " + err;
        }
    }
    CU(cudaStreamSynchronize(st));
    return PTB_OK;
}

int ptb_scene_build_face_attrs(ptb_scene* s) {
    s->ds.face_attrs = nullptr;
    const size_t n = s->faces.n;
    static const bool off = std::getenv("PTB_NO_FACE_ATTRS") != nullptr;      // A/B switch
    if (n == 0 || !s->faces.p || !s->vertices.p || off) { s->face_attrs.release(); return PTB_OK; }
    if (s->face_attrs.n != 2 * n || !s->face_attrs.p) {
        if (s->face_attrs.alloc(2 * n) != cudaSuccess) { cudaGetLastError(); s->face_attrs.release(); return PTB_OK; }     // indexed path stays
    }
    CU(launch_face_attrs(s->faces.p, s->vertices.p, s->face_attrs.p, (unsigned)n, s->ctx->stream));
    s->ds.face_attrs = s->face_attrs.p;
    return PTB_OK;
}

void ptb_scene_destroy(ptb_scene* s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->stream);
    upgrade_shutdown(s);
    s->textures.release(); s->materials.release(); s->shapes.release(); s->shape_nodes.release();
    s->faces.release(); s->vertices.release(); s->mesh_nodes.release(); s->cameras.release(); s->face_attrs.release();
    for (auto& l : s->layers) l.release();
    s->layer_table.release(); s->wide_nodes.release(); s->wide_prims.release(); s->wide_planes.release();
    delete s;
}

int ptb_scene_set_builder(ptb_scene* s, uint32_t builder) {
    if (!s) return fail(PTB_E_INVALID, "ptb_scene_set_builder: NULL scene");
    if (builder > PTB_BUILDER_AUTO) return fail(PTB_E_INVALID, "ptb_scene_set_builder: unknown builder");
    s->builder = builder;
    return PTB_OK;
}
int ptb_scene_wide_download(ptb_scene* s, uint32_t* host_nodes, size_t node_words, float* host_prims, size_t prim_floats) {
    if (!s) return fail(PTB_E_INVALID, "ptb_scene_wide_download: NULL scene");
    { int rc = upgrade_poll(s); if (rc) return rc; }
    if (s->wide_info.status != PTB_OK) return fail(PTB_E_UNSUPPORTED, "ptb_scene_wide_download: no wide BVH");
    if (node_words != (size_t)s->wide_info.node_count * 20 || prim_floats != (size_t)s->wide_info.prim_count * 12)
        return fail(PTB_E_INVALID, "ptb_scene_wide_download: sizes do not match ptb_scene_wide_info");
    CU(cudaSetDevice(s->ctx->device));
    CU(cudaStreamSynchronize(s->ctx->stream));
    if (node_words) CU(cudaMemcpy(host_nodes, s->wide_nodes.p, node_words * 4, cudaMemcpyDeviceToHost));
    if (prim_floats) CU(cudaMemcpy(host_prims, s->wide_prims.p, prim_floats * 4, cudaMemcpyDeviceToHost));
    return PTB_OK;
}
int ptb_scene_wide_info(ptb_scene* s, ptb_wide_info* out) {
    if (!s || !out) return fail(PTB_E_INVALID, "ptb_scene_wide_info: NULL argument");
    { int rc = upgrade_poll(s); if (rc) return rc; }
    *out = s->wide_info;
    return PTB_OK;
}

// ---- film ----
int ptb_film_create(ptb_context* ctx, uint32_t width, uint32_t height, void* external, ptb_film** out) {
    if (!ctx || !out || width == 0 || height == 0) return fail(PTB_E_INVALID, "ptb_film_create: bad argument");
    CU(cudaSetDevice(ctx->device));
    ptb_film* f = new ptb_film();
    f->ctx = ctx; f->width = width; f->height = height; f->external = external != nullptr;
    f->accum = (float4*)external; f->srgb = nullptr; f->linear = nullptr;
    const size_t n = (size_t)width * height;
    if (!external) {
        cudaError_t e = cudaMalloc((void**)&f->accum, n * sizeof(float4));
        if (e != cudaSuccess) { delete f; return cuda_fail(e, "cudaMalloc(film)"); }
        cudaMemsetAsync(f->accum, 0, n * sizeof(float4), ctx->stream);
    }
    *out = f;
    return PTB_OK;
}
void ptb_film_destroy(ptb_film* f) {
    if (!f) return;
    cudaSetDevice(f->ctx->device);
    cudaStreamSynchronize(f->ctx->stream);
    if (!f->external && f->accum) cudaFree(f->accum);
    if (f->srgb) cudaFree(f->srgb);
    if (f->linear) cudaFree(f->linear);
    delete f;
}
void* ptb_film_device_accum(ptb_film* f) { return f ? (void*)f->accum : nullptr; }
int ptb_film_download_accum(ptb_film* f, float* host, size_t count) {
    if (!f || !host || count != (size_t)f->width * f->height * 4) return fail(PTB_E_INVALID, "ptb_film_download_accum: bad argument");
    CU(cudaMemcpyAsync(host, f->accum, count * sizeof(float), cudaMemcpyDeviceToHost, f->ctx->stream));
    CU(cudaStreamSynchronize(f->ctx->stream));
    return PTB_OK;
}
int ptb_film_upload_accum(ptb_film* f, const float* host, size_t count) {
    if (!f || !host || count != (size_t)f->width * f->height * 4) return fail(PTB_E_INVALID, "ptb_film_upload_accum: bad argument");
    CU(cudaMemcpyAsync(f->accum, host, count * sizeof(float), cudaMemcpyHostToDevice, f->ctx->stream));
    CU(cudaStreamSynchronize(f->ctx->stream));
    return PTB_OK;
}
int ptb_film_resolve_device(ptb_film* f, const ptb_resolve_params* p, void* device_srgb8) {
    if (!f || !p || !device_srgb8) return fail(PTB_E_INVALID, "ptb_film_resolve_device: NULL argument");
    if (p->ToneMappingMode > PTB_TONE_MAPPING_MODE_ACES) return fail(PTB_E_INVALID, "ptb_film_resolve: unknown tone mapping mode");
    CU(launch_resolve(f->accum, f->width, f->height, *p, (uchar4*)device_srgb8, nullptr, f->ctx->stream));
    return PTB_OK;
}
int ptb_film_resolve(ptb_film* f, const ptb_resolve_params* p, uint8_t* host_srgb8, float* host_linear) {
    if (!f || !p) return fail(PTB_E_INVALID, "ptb_film_resolve: NULL argument");
    if (p->ToneMappingMode > PTB_TONE_MAPPING_MODE_ACES) return fail(PTB_E_INVALID, "ptb_film_resolve: unknown tone mapping mode");
    CU(cudaSetDevice(f->ctx->device));
    const size_t n = (size_t)f->width * f->height;
    if (host_srgb8 && !f->srgb) CU(cudaMalloc((void**)&f->srgb, n * sizeof(uchar4)));
    if (host_linear && !f->linear) CU(cudaMalloc((void**)&f->linear, n * 3 * sizeof(float)));
    CU(launch_resolve(f->accum, f->width, f->height, *p, host_srgb8 ? f->srgb : nullptr, host_linear ? f->linear : nullptr, f->ctx->stream));
    if (host_srgb8) CU(cudaMemcpyAsync(host_srgb8, f->srgb, n * 4, cudaMemcpyDeviceToHost, f->ctx->stream));
    if (host_linear) CU(cudaMemcpyAsync(host_linear, f->linear, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, f->ctx->stream));
    CU(cudaStreamSynchronize(f->ctx->stream));
    return PTB_OK;
}

// ---- renderer ----
int ptb_renderer_create(ptb_context* ctx, ptb_scene* scene, ptb_film* film, ptb_renderer** out) {
    if (!ctx || !scene || !film || !out) return fail(PTB_E_INVALID, "ptb_renderer_create: NULL argument");
    CU(cudaSetDevice(ctx->device));
    ptb_renderer* r = new ptb_renderer();
    r->ctx = ctx; r->scene = scene; r->film = film;
    std::memset(&r->params, 0, sizeof r->params);
    r->params.RenderFlags = PTB_RENDER_FLAG_ACCUMULATE | PTB_RENDER_FLAG_SAMPLE_JITTER;   // application.cpp:106
    r->params.Traversal = PTB_TRAVERSAL_WIDE;
    r->params.TieMode = PTB_TIE_CANONICAL;
    r->params.TraceVariant = PTB_TRACE_VARIANT_DEFAULT;
    r->frame_index = 0; r->rays_traced = 0; r->kernel_launches = 0; r->queue[0] = r->queue[1] = nullptr;
    r->aux = nullptr; r->ev_fork = r->ev_join = nullptr;
    RenderConsts& pc = r->pc;
    std::memset(&pc, 0, sizeof pc);
    set_film_geometry(pc, film->width, film->height);
    const uint32_t tiles = pc.tiles_x * ((film->height + 15) / 16);
    pc.slot_count = tiles * 256u;
    const size_t S = pc.slot_count;
    const size_t bytes = S * (7 * 16) + (size_t)(tiles + 16) * sizeof(unsigned);
    cudaError_t e = cudaMalloc(&r->state_block, bytes);
    if (e != cudaSuccess) { delete r; return cuda_fail(e, "cudaMalloc(wavefront state)"); }
    cudaMemsetAsync(r->state_block, 0, bytes, ctx->stream);
    char* b = (char*)r->state_block;
    r->st.ray = (float4*)b; b += S * 16;
    r->st.hit0 = (uint4*)b; b += S * 16;
    r->st.hit1 = (uint4*)b; b += S * 16;
    r->st.path0 = (uint4*)b; b += S * 16;
    r->st.path1 = (float4*)b; b += S * 16;
    r->st.path2 = (float4*)b; b += S * 16;
    r->st.path3 = (float4*)b; b += S * 16;
    r->st.tile_paths = (unsigned*)b; b += (size_t)tiles * sizeof(unsigned);
    r->st.ray_counter = (unsigned*)b;
    *out = r;
    return PTB_OK;
}
void ptb_renderer_destroy(ptb_renderer* r) {
    if (!r) return;
    cudaSetDevice(r->ctx->device);
    cudaStreamSynchronize(r->ctx->stream);
    cudaFree(r->state_block);
    cudaFree(r->queue[0]); cudaFree(r->queue[1]);
    if (r->aux) cudaStreamDestroy(r->aux);
    if (r->ev_fork) cudaEventDestroy(r->ev_fork);
    if (r->ev_join) cudaEventDestroy(r->ev_join);
    delete r;
}
int ptb_renderer_set_params(ptb_renderer* r, const ptb_render_params* p) {
    if (!r || !p) return fail(PTB_E_INVALID, "ptb_renderer_set_params: NULL argument");
    if (p->Integrator > PTB_INTEGRATOR_MEGAKERNEL || p->Traversal > PTB_TRAVERSAL_WIDE || p->TieMode > PTB_TIE_CANONICAL || p->TraceVariant > PTB_TRACE_VARIANT_MAX || p->ScatterMode > PTB_SCATTER_DIRECT_STAGED || p->Pipelines > 2u)
        return fail(PTB_E_INVALID, "ptb_renderer_set_params: unknown enum value");
    if (p->Traversal == PTB_TRAVERSAL_WIDE && p->TieMode != PTB_TIE_CANONICAL)
        return fail(PTB_E_INVALID, "ptb_renderer_set_params: the wide traversal is order-free and needs PTB_TIE_CANONICAL");
    r->params = *p;
    return PTB_OK;
}
int ptb_renderer_set_frame_index(ptb_renderer* r, uint32_t f) {
    if (!r) return fail(PTB_E_INVALID, "NULL renderer");
    r->frame_index = f;
    return PTB_OK;
}

static int prepare(ptb_renderer* r) {
    { int rc = upgrade_poll(r->scene); if (rc) return rc; }
    const DeviceScene& ds = r->scene->ds;
    if (r->params.CameraIndex >= ds.camera_count) return fail(PTB_E_INVALID, "renderer: CameraIndex out of range");
    if (r->params.Traversal == PTB_TRAVERSAL_WIDE && ds.shape_count > 0 && r->scene->wide_info.status != PTB_OK)
        return fail(PTB_E_UNSUPPORTED, "renderer: wide BVH unavailable for this scene (" + g_error + ")");
    r->pc.camera_index = r->params.CameraIndex;
    r->pc.render_flags = r->params.RenderFlags;
    r->pc.path_length_limit = r->params.PathLengthLimit;
    r->pc.termination_probability = r->params.PathTerminationProbability;
    return PTB_OK;
}

int ptb_renderer_reset(ptb_renderer* r) {
    if (!r) return fail(PTB_E_INVALID, "NULL renderer");
    int rc = prepare(r);
    if (rc) return rc;
    CU(cudaSetDevice(r->ctx->device));
    r->pc.random_seed = r->frame_index;      // basic.cpp:285-304: seed = FrameIndex, Restart = 1
    CU(launch_restart(r->scene->ds, r->pc, r->st, r->film->accum, r->ctx->stream));
    r->kernel_launches += 1;
    return PTB_OK;
}

static int run_rounds(ptb_renderer* r, uint32_t rounds, uint32_t seed_stride) {
    const DeviceScene& ds = r->scene->ds;
    cudaStream_t st = r->ctx->stream;
    const uint64_t live = (uint64_t)r->film->width * r->film->height;
    if (r->params.Integrator == PTB_INTEGRATOR_MEGAKERNEL) {
        CU(launch_megakernel(ds, r->pc, r->st, r->film->accum, r->params.Traversal, r->params.TieMode, rounds, seed_stride, st));
        r->kernel_launches += 1;
    } else {
        const unsigned pipes = pipeline_count(r);
        RenderConsts pc[2] = {r->pc, r->pc};
        WavefrontState wst[2] = {r->st, r->st};
        cudaStream_t str[2] = {st, st};
        if (pipes == 2u) {
            if (!r->aux) {
                CU(cudaStreamCreateWithFlags(&r->aux, cudaStreamNonBlocking));
                CU(cudaEventCreateWithFlags(&r->ev_fork, cudaEventDisableTiming));
                CU(cudaEventCreateWithFlags(&r->ev_join, cudaEventDisableTiming));
            }
            const uint32_t half = (r->pc.slot_count / 512u) * 256u;
            pc[0].slot_count = half;
            pc[1].slot_base = half; pc[1].slot_count = r->pc.slot_count - half;
            wst[1].ray_counter = r->st.ray_counter + 1;      // and its scatter tile counter at [+2] = base + 3
            str[1] = r->aux;
            CU(cudaEventRecord(r->ev_fork, st));
            CU(cudaStreamWaitEvent(r->aux, r->ev_fork, 0));
        }
        for (uint32_t k = 0; k < rounds; k++) {
            for (unsigned p = 0; p < pipes; p++) CU(launch_trace(ds, pc[p], wst[p], r->params.Traversal, r->params.TieMode, r->params.TraceVariant, str[p]));
            r->kernel_launches += pipes;
            for (unsigned p = 0; p < pipes; p++) {
                int rc = scatter_step(r, pc[p], wst[p], p, str[p]);
                if (rc) return rc;
                pc[p].random_seed += seed_stride;
            }
        }
        if (pipes == 2u) {
            CU(cudaEventRecord(r->ev_join, r->aux));
            CU(cudaStreamWaitEvent(st, r->ev_join, 0));
        }
    }
    r->rays_traced += live * rounds;
    return PTB_OK;
}

int ptb_renderer_run(ptb_renderer* r, uint32_t rounds) {
    if (!r) return fail(PTB_E_INVALID, "NULL renderer");
    int rc = prepare(r);
    if (rc) return rc;
    CU(cudaSetDevice(r->ctx->device));
    r->frame_index += 1;                      // basic.cpp:315
    r->pc.random_seed = r->frame_index;
    return run_rounds(r, rounds, 0u);         // all rounds of one call share the seed (basic.cpp:323-331)
}

int ptb_renderer_run_frames(ptb_renderer* r, uint32_t frames) {
    if (!r) return fail(PTB_E_INVALID, "NULL renderer");
    if (frames == 0) return PTB_OK;
    int rc = prepare(r);
    if (rc) return rc;
    CU(cudaSetDevice(r->ctx->device));
    r->pc.random_seed = r->frame_index + 1;   // frame k of this call runs with seed FrameIndex + k
    rc = run_rounds(r, frames, 1u);
    r->frame_index += frames;
    return rc;
}

int ptb_renderer_profile_frames(ptb_renderer* r, uint32_t frames, ptb_kernel_times* out) {
    if (!r || !out) return fail(PTB_E_INVALID, "ptb_renderer_profile_frames: NULL argument");
    std::memset(out, 0, sizeof *out);
    if (frames == 0) return PTB_OK;
    if (frames > 4096) return fail(PTB_E_INVALID, "ptb_renderer_profile_frames: at most 4096 frames per call");
    int rc = prepare(r);
    if (rc) return rc;
    CU(cudaSetDevice(r->ctx->device));
    cudaStream_t st = r->ctx->stream;
    const DeviceScene& ds = r->scene->ds;
    const bool mega = r->params.Integrator == PTB_INTEGRATOR_MEGAKERNEL;
    const uint32_t per = mega ? 2u : 3u;
    std::vector<cudaEvent_t> ev((size_t)frames * per);
    for (auto& e : ev) CU(cudaEventCreate(&e));
    RenderConsts pc = r->pc;
    pc.random_seed = r->frame_index + 1;
    for (uint32_t k = 0; k < frames; k++) {
        cudaEvent_t* e = &ev[(size_t)k * per];
        CU(cudaEventRecord(e[0], st));
        if (mega) {
            CU(launch_megakernel(ds, pc, r->st, r->film->accum, r->params.Traversal, r->params.TieMode, 1u, 1u, st));
            CU(cudaEventRecord(e[1], st));
        } else {
            CU(launch_trace(ds, pc, r->st, r->params.Traversal, r->params.TieMode, r->params.TraceVariant, st));
            CU(cudaEventRecord(e[1], st));
            int rc2 = scatter_step(r, pc, r->st, 0u, st);
            if (rc2) return rc2;
            CU(cudaEventRecord(e[2], st));
        }
        pc.random_seed += 1;
    }
    CU(cudaStreamSynchronize(st));
    for (uint32_t k = 0; k < frames; k++) {
        cudaEvent_t* e = &ev[(size_t)k * per];
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, e[0], e[1]));
        if (mega) { out->mega_ms += ms; out->mega_launches++; }
        else {
            out->trace_ms += ms; out->trace_launches++;
            CU(cudaEventElapsedTime(&ms, e[1], e[2]));
            out->scatter_ms += ms; out->scatter_launches++;
        }
    }
    for (auto& e : ev) cudaEventDestroy(e);
    r->frame_index += frames;
    r->rays_traced += (uint64_t)r->film->width * r->film->height * frames;
    r->kernel_launches += (uint64_t)frames * 1u;      // trace or megakernel; scatter_step counts its own
    return PTB_OK;
}

int ptb_renderer_trace_stats(ptb_renderer* r, ptb_trace_stats* out) {
    if (!r || !out) return fail(PTB_E_INVALID, "ptb_renderer_trace_stats: NULL argument");
    std::memset(out, 0, sizeof *out);
    int rc = prepare(r);
    if (rc) return rc;
    if (r->params.Traversal != PTB_TRAVERSAL_WIDE || r->scene->ds.wide_tlas_root == 0xFFFFFFFFu)
        return fail(PTB_E_UNSUPPORTED, "ptb_renderer_trace_stats: counts the wide-BVH trace kernel only");
    CU(cudaSetDevice(r->ctx->device));
    cudaStream_t st = r->ctx->stream;
    unsigned long long* d = nullptr;
    CU(cudaMalloc((void**)&d, 8 * sizeof(unsigned long long)));
    unsigned long long h[8] = {};
    cudaError_t e = cudaMemsetAsync(d, 0, 8 * sizeof(unsigned long long), st);
    // the trace step is idempotent on the wavefront state (it rewrites the same hit records); scatter re-arms the ray counter
    if (e == cudaSuccess) e = cudaMemsetAsync(r->st.ray_counter, 0, sizeof(unsigned), st);
    if (e == cudaSuccess) e = launch_trace_stats(r->scene->ds, r->pc, r->st, d, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(r->st.ray_counter, 0, sizeof(unsigned), st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(h, d, sizeof h, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d);
    if (e != cudaSuccess) return cuda_fail(e, "ptb_renderer_trace_stats");
    out->node_steps = h[0]; out->prim_tests = h[1]; out->warp_iterations = h[2]; out->flushes = h[3]; out->rays = h[4];
    out->node_bytes = 80; out->prim_bytes = 48; out->state_bytes = 48;
    return PTB_OK;
}

int ptb_renderer_sync(ptb_renderer* r) {
    if (!r) return fail(PTB_E_INVALID, "NULL renderer");
    CU(cudaStreamSynchronize(r->ctx->stream));
    return PTB_OK;
}

int ptb_renderer_stats(ptb_renderer* r, ptb_stats* out) {
    if (!r || !out) return fail(PTB_E_INVALID, "ptb_renderer_stats: NULL argument");
    const uint32_t tiles = r->pc.slot_count / 256u;
    std::vector<unsigned> host(tiles);
    CU(cudaMemcpyAsync(host.data(), r->st.tile_paths, tiles * sizeof(unsigned), cudaMemcpyDeviceToHost, r->ctx->stream));
    CU(cudaStreamSynchronize(r->ctx->stream));
    uint64_t paths = 0;
    for (unsigned c : host) paths += c;
    out->rays_traced = r->rays_traced;
    out->paths_finished = paths;
    out->kernel_launches = r->kernel_launches;
    out->frame_index = r->frame_index;
    return PTB_OK;
}

uint32_t ptb_renderer_slot_count(ptb_renderer* r) { return r ? r->pc.slot_count : 0; }

// Named SoA arrays of the reference's state (basic.glsl.inc:23-59), gathered from the 16-byte records.
int ptb_renderer_download_state(ptb_renderer* r, const char* name, void* host, size_t bytes) {
    if (!r || !name || !host) return fail(PTB_E_INVALID, "ptb_renderer_download_state: NULL argument");
    const size_t S = r->pc.slot_count;
    if (bytes != S * 4) return fail(PTB_E_INVALID, "ptb_renderer_download_state: expected slot_count*4 bytes");
    struct Field { const char* name; int record; int word; };
    static const Field fields[] = {
        {"OriginX", 0, 0}, {"OriginY", 0, 1}, {"OriginZ", 0, 2}, {"PackedVelocity", 0, 3},
        {"Time", 1, 0}, {"ShapeAndMaterialIndex", 1, 1}, {"PrimitiveIndex", 1, 2}, {"ShapeType", 1, 3},
        {"PrimitiveCoordinateX", 2, 0}, {"PrimitiveCoordinateY", 2, 1}, {"PrimitiveCoordinateZ", 2, 2},
        {"ImagePosition", 3, 0}, {"NormalizedLambda0", 3, 1}, {"ActiveShapeIndex01", 3, 2}, {"ActiveShapeIndex23", 3, 3},
        {"Throughput0", 4, 0}, {"Throughput1", 4, 1}, {"Throughput2", 4, 2}, {"Throughput3", 4, 3},
        // the reference stores Probability.w in array 2 and .z in array 3 (basic.glsl.inc:176-177,208-209)
        {"Probability0", 5, 0}, {"Probability1", 5, 1}, {"Probability2", 5, 3}, {"Probability3", 5, 2},
        {"SampleR", 6, 0}, {"SampleG", 6, 1}, {"SampleB", 6, 2}, {"PathLength", 6, 3},
    };
    for (const Field& f : fields) {
        if (std::strcmp(f.name, name)) continue;
        std::vector<uint32_t> rec(S * 4);
        const char* base = (const char*)r->state_block + (size_t)f.record * S * 16;
        CU(cudaMemcpyAsync(rec.data(), base, S * 16, cudaMemcpyDeviceToHost, r->ctx->stream));
        CU(cudaStreamSynchronize(r->ctx->stream));
        uint32_t* o = (uint32_t*)host;
        for (size_t i = 0; i < S; i++) o[i] = rec[4 * i + f.word];
        return PTB_OK;
    }
    // derived from the stored hit on demand (the scatter kernel derives them the same way)
    static const char* derived[] = {"PackedNormal", "PackedTangentX", "TextureU", "TextureV"};
    for (int k = 0; k < 4; k++) {
        if (std::strcmp(derived[k], name)) continue;
        uint4* tmp = nullptr;
        CU(cudaMalloc((void**)&tmp, S * 16));
        cudaError_t e = cudaMemsetAsync(tmp, 0, S * 16, r->ctx->stream);
        if (e == cudaSuccess) e = launch_hit_attributes(r->scene->ds, r->pc, r->st, tmp, r->ctx->stream);
        std::vector<uint32_t> rec(S * 4);
        if (e == cudaSuccess) e = cudaMemcpyAsync(rec.data(), tmp, S * 16, cudaMemcpyDeviceToHost, r->ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(r->ctx->stream);
        cudaFree(tmp);
        if (e != cudaSuccess) return cuda_fail(e, "ptb_renderer_download_state");
        uint32_t* o = (uint32_t*)host;
        for (size_t i = 0; i < S; i++) o[i] = rec[4 * i + k];
        return PTB_OK;
    }
    if (!std::strcmp(name, "Duration")) {
        float* o = (float*)host;
        for (size_t i = 0; i < S; i++) o[i] = PTB_HIT_TIME_LIMIT;
        return PTB_OK;
    }
    return fail(PTB_E_INVALID, std::string("ptb_renderer_download_state: unknown array ") + name);
}

// ---- ray cast harness ----
int ptb_raycast(ptb_scene* s, const ptb_ray* host_rays, size_t n, uint32_t traversal, uint32_t tie, ptb_hit* host_hits) {
    if (!s || (n && (!host_rays || !host_hits))) return fail(PTB_E_INVALID, "ptb_raycast: NULL argument");
    { int rc = upgrade_poll(s); if (rc) return rc; }
    if (traversal > PTB_TRAVERSAL_WIDE || tie > PTB_TIE_CANONICAL) return fail(PTB_E_INVALID, "ptb_raycast: unknown enum value");
    if (traversal == PTB_TRAVERSAL_WIDE && tie != PTB_TIE_CANONICAL) return fail(PTB_E_INVALID, "ptb_raycast: the wide traversal needs PTB_TIE_CANONICAL");
    if (traversal == PTB_TRAVERSAL_WIDE && s->ds.shape_count > 0 && s->wide_info.status != PTB_OK) return fail(PTB_E_UNSUPPORTED, "ptb_raycast: wide BVH unavailable");
    if (n == 0) return PTB_OK;
    CU(cudaSetDevice(s->ctx->device));
    cudaStream_t st = s->ctx->stream;
    ptb_ray* d_rays = nullptr; ptb_hit* d_hits = nullptr;
    CU(cudaMalloc((void**)&d_rays, n * sizeof(ptb_ray)));
    cudaError_t e = cudaMalloc((void**)&d_hits, n * sizeof(ptb_hit));
    if (e != cudaSuccess) { cudaFree(d_rays); return cuda_fail(e, "cudaMalloc(hits)"); }
    e = cudaMemcpyAsync(d_rays, host_rays, n * sizeof(ptb_ray), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = launch_raycast(s->ds, d_rays, n, traversal, tie, s->raycast_variant, s->ctx->counter, d_hits, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(host_hits, d_hits, n * sizeof(ptb_hit), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(d_rays); cudaFree(d_hits);
    if (e != cudaSuccess) return cuda_fail(e, "ptb_raycast");
    return PTB_OK;
}

int ptb_raycast_device(ptb_scene* s, const ptb_ray* d_rays, size_t n, uint32_t traversal, uint32_t tie, ptb_hit* d_hits) {
    if (!s || (n && (!d_rays || !d_hits))) return fail(PTB_E_INVALID, "ptb_raycast_device: NULL argument");
    { int rc = upgrade_poll(s); if (rc) return rc; }
    if (traversal > PTB_TRAVERSAL_WIDE || tie > PTB_TIE_CANONICAL) return fail(PTB_E_INVALID, "ptb_raycast_device: unknown enum value");
    if (traversal == PTB_TRAVERSAL_WIDE && tie != PTB_TIE_CANONICAL) return fail(PTB_E_INVALID, "ptb_raycast_device: the wide traversal needs PTB_TIE_CANONICAL");
    if (traversal == PTB_TRAVERSAL_WIDE && s->ds.shape_count > 0 && s->wide_info.status != PTB_OK) return fail(PTB_E_UNSUPPORTED, "ptb_raycast_device: wide BVH unavailable");
    CU(cudaSetDevice(s->ctx->device));
    CU(launch_raycast(s->ds, d_rays, n, traversal, tie, s->raycast_variant, s->ctx->counter, d_hits, s->ctx->stream));
    return PTB_OK;
}

int ptb_preview_render(ptb_scene* s, const ptb_preview_params* p, float* host_rgba, uint32_t* host_shapes, uint32_t* hit_shape) {
    if (!s || !p || !host_rgba) return fail(PTB_E_INVALID, "ptb_preview_render: NULL argument");
    { int rc = upgrade_poll(s); if (rc) return rc; }
    if (p->RenderMode > PTB_PREVIEW_MODE_SCENE_COMPLEXITY || p->Traversal > PTB_TRAVERSAL_WIDE) return fail(PTB_E_INVALID, "ptb_preview_render: unknown mode");
    if (p->Traversal == PTB_TRAVERSAL_WIDE && s->wide_info.status != PTB_OK) return fail(PTB_E_UNSUPPORTED, "ptb_preview_render: wide BVH unavailable");
    const size_t n = (size_t)p->RenderSizeX * p->RenderSizeY;
    if (n == 0 || n > 0x7FFFFFFFull) return fail(PTB_E_INVALID, "ptb_preview_render: bad render size");
    CU(cudaSetDevice(s->ctx->device));
    cudaStream_t st = s->ctx->stream;
    PreviewConsts c;
    std::memcpy(c.to, p->CameraTransform.To, sizeof c.to);
    c.mode = p->RenderMode; c.brightness = p->Brightness; c.selected_shape = p->SelectedShapeIndex;
    c.width = p->RenderSizeX; c.height = p->RenderSizeY; c.mouse_x = p->MouseX; c.mouse_y = p->MouseY;
    DevBuf<float4> rgba;
    DevBuf<uint32_t> shapes, pick;
    CU(rgba.alloc(n));
    CU(shapes.alloc(n));
    const uint32_t none = 0xFFFFFFFFu;
    CU(pick.upload(&none, 1, st));
    cudaError_t e = launch_preview(s->ds, c, p->Traversal, PTB_TIE_CANONICAL, rgba.p, shapes.p, pick.p, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(host_rgba, rgba.p, n * sizeof(float4), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && host_shapes) e = cudaMemcpyAsync(host_shapes, shapes.p, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && hit_shape) e = cudaMemcpyAsync(hit_shape, pick.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    rgba.release(); shapes.release(); pick.release();
    if (e != cudaSuccess) return cuda_fail(e, "ptb_preview_render");
    return PTB_OK;
}

}  // extern "C"
