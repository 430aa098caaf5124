// TEST INFRASTRUCTURE ONLY — not part of the product, never loaded by libptb.so.
//
// Compiles the per-thread device functions of path_tracer_b200/csrc/*.cuh with
// g++ (they are __host__ __device__) so that the kernel logic can be debugged
// against the oracle in containers without a GPU.  The product has no CPU
// execution path; this file exists so that a bug found on the GPU box can be
// reproduced under gdb/asan here.  Transcendentals come from glibc instead of
// libdevice, so results match the GPU only to rounding.
#include <cuda_runtime.h>
#include <thread>
#include <vector>
#include <atomic>
#include "ptb_wide.cuh"
#include "ptb_scatter.cuh"
#include "ptb_resolve.cuh"
#include "ptb_wide_build.h"

using namespace ptb;

namespace {
template <class F>
void parallel_for(uint64_t n, uint64_t chunk, F&& fn) {
    int threads = (int)std::thread::hardware_concurrency();
    if (threads <= 1 || n <= chunk) { fn(0, n); return; }
    std::atomic<uint64_t> next(0);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++) pool.emplace_back([&]() { for (;;) { uint64_t b = next.fetch_add(chunk); if (b >= n) break; fn(b, std::min(n, b + chunk)); } });
    for (auto& th : pool) th.join();
}

struct Emu {
    DeviceScene ds;
    WideBuild wb;
    std::vector<PathState> path; std::vector<Ray> ray; std::vector<PackedHit> hit; std::vector<u32> prim;
    std::vector<float4> accum;
    RenderConsts pc;
    u32 frame_index = 0;
    uint64_t paths = 0;
};

HitId hit_mode(const DeviceScene& s, const Ray& r, int mode) {
    if (mode == 0) return closest_hit_bvh2<false>(s, r);
    if (mode == 1) return closest_hit_bvh2<true>(s, r);
    return closest_hit_wide(s, r);
}
}  // namespace

extern "C" {

void* emu_create(const ptb_scene_view* v, u32 w, u32 h) {
    Emu* e = new Emu();
    memset(&e->ds, 0, sizeof e->ds);
    e->ds.textures = v->Textures; e->ds.material_data = v->MaterialData; e->ds.shapes = v->Shapes; e->ds.shape_nodes = v->ShapeNodes;
    e->ds.mesh_faces = v->MeshFaces; e->ds.mesh_vertices = v->MeshVertices; e->ds.mesh_nodes = v->MeshNodes; e->ds.cameras = v->Cameras;
    e->ds.atlas_layers = v->AtlasLayers; e->ds.atlas_width = v->AtlasWidth; e->ds.atlas_height = v->AtlasHeight;
    e->ds.shape_count = v->ShapeCount; e->ds.camera_count = v->CameraCount; e->ds.globals = *v->Globals;
    std::string err;
    int rc = build_wide_bvh(v, &e->wb, &err);
    if (rc != 0) { fprintf(stderr, "emu: wide build failed: %s\n", err.c_str()); delete e; return nullptr; }
    e->ds.wide_nodes = (const uint4*)e->wb.nodes.data(); e->ds.wide_tris = (const float4*)e->wb.prims.data();
    e->ds.wide_planes = e->wb.planes.data(); e->ds.wide_plane_count = (u32)(e->wb.planes.size() / 2);
    e->ds.wide_tlas_root = e->wb.node_count ? 0u : 0xFFFFFFFFu;
    e->ds.features = PTB_FEAT_ALL;
    memset(&e->pc, 0, sizeof e->pc);
    set_film_geometry(e->pc, w, h); e->pc.slot_count = w * h;
    e->pc.render_flags = 3;
    size_t n = (size_t)w * h;
    e->path.resize(n); e->ray.resize(n); e->hit.resize(n); e->prim.assign(n, 0xFFFFFFFFu); e->accum.assign(n, make_float4(0, 0, 0, 0));
    return e;
}
void emu_destroy(void* p) { delete (Emu*)p; }
void emu_udivmod(u32 n, u32 d, u32* out2) { RenderConsts pc{}; set_film_geometry(pc, d, 16); udivmod(n, d, pc.width_magic, out2[0], out2[1]); }
void emu_unpack_snorm2x16(u32 p, float* out2) { f2 v = unpack_snorm2x16(p); out2[0] = v.x; out2[1] = v.y; }
void emu_wide_info(void* p, u32* out) { Emu* e = (Emu*)p; out[0] = e->wb.node_count; out[1] = e->wb.prim_count; out[2] = e->wb.depth; out[3] = e->wb.bvh2_depth; out[4] = (u32)(e->wb.build_seconds * 1000); }

void emu_wide_download(void* p, u32* nodes, float* prims) {
    Emu* e = (Emu*)p;
    memcpy(nodes, e->wb.nodes.data(), (size_t)e->wb.node_count * 80);
    memcpy(prims, e->wb.prims.data(), (size_t)e->wb.prim_count * 48);
}

void emu_raycast(void* p, const ptb_ray* rays, uint64_t n, int mode, ptb_hit* out) {
    Emu* e = (Emu*)p;
    parallel_for(n, 1024, [&](uint64_t b, uint64_t en) {
        for (uint64_t i = b; i < en; i++) {
            Ray r; r.origin = mk3(rays[i].ox, rays[i].oy, rays[i].oz); r.velocity = mk3(rays[i].dx, rays[i].dy, rays[i].dz); r.duration = rays[i].duration;
            HitId h = hit_mode(e->ds, r, mode);
            ptb_hit o; memset(&o, 0, sizeof o);
            o.t = h.t; o.shape = h.shape; o.prim = 0xFFFFFFFFu;
            if (h.shape != PTB_SHAPE_NONE) {
                Surface sf = finish_hit(e->ds, h);
                o.prim = h.prim; o.material = sf.material; o.shape_type = h.shape_type;
                o.nx = sf.normal.x; o.ny = sf.normal.y; o.nz = sf.normal.z; o.tx = sf.tangent_x.x; o.ty = sf.tangent_x.y; o.tz = sf.tangent_x.z;
                o.u = sf.uv.x; o.v = sf.uv.y; o.bx = h.coords.x; o.by = h.coords.y; o.bz = h.coords.z;
            }
            out[i] = o;
        }
    });
}

// Traversal statistics of the wide BVH on caller rays: node steps and primitive tests per ray
// (builder tuning aid: lets BVH quality be compared on the CPU).
void emu_wide_stats(void* p, const ptb_ray* rays, uint64_t n, uint64_t* out3) {
    Emu* e = (Emu*)p;
    std::atomic<uint64_t> nodes(0), prims(0), maxsp(0);
    parallel_for(n, 1024, [&](uint64_t b, uint64_t en) {
        uint64_t ln = 0, lp = 0, ms = 0;
        for (uint64_t i = b; i < en; i++) {
            Ray r; r.origin = mk3(rays[i].ox, rays[i].oy, rays[i].oz); r.velocity = mk3(rays[i].dx, rays[i].dy, rays[i].dz); r.duration = rays[i].duration;
            HitId h = wide_begin(e->ds, r);
            if (e->ds.wide_tlas_root == 0xFFFFFFFFu) continue;
            const WideCtx c = wide_ctx(r);
            LocalStack stack; stack.sp = 0;
            uint2 ng = make_uint2(0u, 0x80000000u);
            u32 cur = PTB_SHAPE_NONE; Ray obj = r;
            while (true) {
                PrimGroup pg;
                wide_node_step(e->ds, c, h.t, ng, pg, stack); ln++;
                if ((uint64_t)stack.sp > ms) ms = stack.sp;
                while (pg.bits) { wide_prim_step(e->ds, r, pg, cur, obj, h); lp++; }
                if (ng.y <= 0x00FFFFFFu) { if (stack.empty()) break; ng = stack.pop(); }
            }
        }
        nodes += ln; prims += lp;
        uint64_t m = maxsp.load(); while (ms > m && !maxsp.compare_exchange_weak(m, ms)) {}
    });
    out3[0] = nodes; out3[1] = prims; out3[2] = maxsp;
}

void emu_set(void* p, u32 camera, u32 flags, u32 limit, float term, u32 frame_index) {
    Emu* e = (Emu*)p;
    e->pc.camera_index = camera; e->pc.render_flags = flags; e->pc.path_length_limit = limit; e->pc.termination_probability = term; e->frame_index = frame_index;
}
void emu_reset(void* p) {
    Emu* e = (Emu*)p;
    e->pc.random_seed = e->frame_index;
    const u32 W = e->pc.width;
    parallel_for(e->path.size(), 4096, [&](uint64_t b, uint64_t en) {
        for (uint64_t i = b; i < en; i++) {
            u32 x = (u32)(i % W), y = (u32)(i / W);
            Rng g; g.state = y * 65537u + x + e->pc.random_seed * 277803737u;
            generate_new_path(e->ds, e->pc, g, x, y, e->path[i], e->ray[i]);
            e->ray[i].velocity = unpack_unit_vector(pack_unit_vector(e->ray[i].velocity));
            e->accum[i] = make_float4(0, 0, 0, 0);
        }
    });
    e->paths = 0;
}
void emu_run(void* p, u32 rounds, int mode, int per_round_seed) {
    Emu* e = (Emu*)p;
    const u32 W = e->pc.width;
    if (!per_round_seed) { e->frame_index += 1; e->pc.random_seed = e->frame_index; }
    for (u32 k = 0; k < rounds; k++) {
        if (per_round_seed) { e->frame_index += 1; e->pc.random_seed = e->frame_index; }
        std::atomic<uint64_t> fin(0);
        parallel_for(e->path.size(), 1024, [&](uint64_t b, uint64_t en) {
            uint64_t f = 0;
            for (uint64_t i = b; i < en; i++) {
                u32 x = (u32)(i % W), y = (u32)(i / W);
                HitId h = hit_mode(e->ds, e->ray[i], mode);
                PackedHit ph = packed_miss();
                e->prim[i] = 0xFFFFFFFFu;
                if (h.shape != PTB_SHAPE_NONE) { Surface sf = finish_hit(e->ds, h); ph = pack_hit(h, sf); e->prim[i] = h.prim; }
                e->hit[i] = ph;
                SurfaceHit sh = unpack_hit(ph, e->ray[i]);
                if (scatter_invocation(e->ds, e->pc, x, y, e->path[i], e->ray[i], sh, e->accum.data())) f++;
                e->ray[i].velocity = unpack_unit_vector(pack_unit_vector(e->ray[i].velocity));
                u32 w01, w23; pack_active(e->path[i].active, w01, w23); unpack_active(w01, w23, e->path[i].active);
            }
            fin += f;
        });
        e->paths += fin.load();
    }
}
const float* emu_accum(void* p) { return (const float*)((Emu*)p)->accum.data(); }
uint64_t emu_paths(void* p) { return ((Emu*)p)->paths; }
void emu_resolve(void* p, float brightness, u32 mode, float white, float* out_linear, unsigned char* out_srgb8) {
    Emu* e = (Emu*)p;
    for (size_t i = 0; i < e->accum.size(); i++) {
        f3 c = resolve_pixel(e->accum[i], brightness, mode, white);
        out_linear[3 * i] = c.x; out_linear[3 * i + 1] = c.y; out_linear[3 * i + 2] = c.z;
        out_srgb8[4 * i] = (unsigned char)linear_to_srgb8(c.x); out_srgb8[4 * i + 1] = (unsigned char)linear_to_srgb8(c.y);
        out_srgb8[4 * i + 2] = (unsigned char)linear_to_srgb8(c.z); out_srgb8[4 * i + 3] = 255;
    }
}
}
