// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// extern "C" surface over the shader restatement, bound by tests/ (ctypes),
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
// Nothing under path_tracer_b200/ links or loads this library.
//
// Parity: trace, scatter, resolve and the preview visualiser are pinned bit for bit against the
// reference's own GLSL sources compiled as C++ (oracle/ref -> oracle/_ref/libref.so,
// tests/test_reference_shaders.py, tests/golden/ref_*.npz).
//
// Build: see oracle/Makefile  (g++ -O2 -ffp-contract=off, no -ffast-math).
#include <atomic>
#include <thread>
#include <vector>
#include <algorithm>
#include "sh_integrator.h"
#include "sh_preview.h"

using namespace orc;

namespace {

template <class F>
void parallel_for(uint64_t n, uint64_t chunk, int threads, F&& fn) {
    if (threads <= 1 || n <= chunk) { fn(0, n, 0); return; }
    std::atomic<uint64_t> next(0);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; t++) {
        pool.emplace_back([&, t]() {
            for (;;) {
                uint64_t b = next.fetch_add(chunk);
                if (b >= n) break;
                fn(b, std::min(n, b + chunk), t);
            }
        });
    }
    for (auto& th : pool) th.join();
}

struct renderer {
    wavefront_state W;
    uint FrameIndex = 0;   // basic.hpp:19
};

}  // namespace

extern "C" {

// One ray in, one hit out (flat PODs shared with tests/ via ctypes).
struct orc_ray { float ox, oy, oz, dx, dy, dz, duration; };
struct orc_hit {
    float t;
    uint shape, prim, material, shape_type;
    float nx, ny, nz, tx, ty, tz, u, v;
    float bx, by, bz;   // PrimitiveCoordinates
    uint scene_complexity, mesh_complexity, node_fetches, face_tests, shape_visits;
};

int orc_sizeof_scene_view() { return (int)sizeof(scene_view); }

// K1/K5 on caller-supplied rays (SURVEY.md §8 "ray cast" parity level).
void orc_raycast(const scene_view* S, const orc_ray* rays, uint64_t n, int tie, orc_hit* out, int threads) {
    parallel_for(n, 4096, threads, [&](uint64_t b, uint64_t e, int) {
        for (uint64_t i = b; i < e; i++) {
            ray R;
            R.Origin = V3(rays[i].ox, rays[i].oy, rays[i].oz);
            R.Velocity = V3(rays[i].dx, rays[i].dy, rays[i].dz);
            R.Duration = rays[i].duration;
            hit H = Trace(*S, R, (tie_mode)tie);
            orc_hit& o = out[i];
            memset(&o, 0, sizeof(o));
            o.t = H.Time; o.shape = H.ShapeIndex; o.prim = H.PrimitiveIndex;
            o.scene_complexity = H.SceneComplexity; o.mesh_complexity = H.MeshComplexity;
            o.node_fetches = H.NodeFetches; o.face_tests = H.FaceTests; o.shape_visits = H.ShapeVisits;
            if (H.ShapeIndex != SHAPE_INDEX_NONE) {
                o.material = H.MaterialIndex; o.shape_type = H.ShapeType;
                o.nx = H.Normal.x; o.ny = H.Normal.y; o.nz = H.Normal.z;
                o.tx = H.TangentX.x; o.ty = H.TangentX.y; o.tz = H.TangentX.z;
                o.u = H.UV.x; o.v = H.UV.y;
                o.bx = H.PrimitiveCoordinates.x; o.by = H.PrimitiveCoordinates.y; o.bz = H.PrimitiveCoordinates.z;
            } else {
                o.prim = 0xFFFFFFFFu;
            }
        }
    });
}

// Brute force over every face of every mesh shape + analytic shapes, canonical
// ties.  Independent of both BVH levels: used to validate traversal itself.
void orc_raycast_bruteforce(const scene_view* S, const orc_ray* rays, uint64_t n, orc_hit* out, int threads) {
    parallel_for(n, 64, threads, [&](uint64_t b, uint64_t e, int) {
        for (uint64_t i = b; i < e; i++) {
            ray R;
            R.Origin = V3(rays[i].ox, rays[i].oy, rays[i].oz);
            R.Velocity = V3(rays[i].dx, rays[i].dy, rays[i].dz);
            R.Duration = rays[i].duration;
            hit H; memset(&H, 0, sizeof(H));
            H.ShapeIndex = SHAPE_INDEX_NONE; H.PrimitiveIndex = 0xFFFFFFFFu; H.Time = R.Duration;
            for (uint s = 0; s < S->ShapeCount; s++) {
                const packed_shape& Sh = S->Shapes[s];
                if (Sh.Type == SHAPE_TYPE_MESH_INSTANCE) {
                    ray O = InverseTransformRay(R, Sh.Transform);
                    // Faces of a mesh: the union of the leaf ranges below its root.
                    std::vector<uint> stack{Sh.MeshRootNodeIndex};
                    while (!stack.empty()) {
                        packed_mesh_node N = S->MeshNodes[stack.back()]; stack.pop_back();
                        if (N.FaceEndIndex > 0) {
                            for (uint f = N.FaceBeginOrNodeIndex; f < N.FaceEndIndex; f++) IntersectMeshFace(*S, O, f, s, H, TIE_CANONICAL);
                        } else { stack.push_back(N.FaceBeginOrNodeIndex); stack.push_back(N.FaceBeginOrNodeIndex + 1); }
                    }
                } else {
                    IntersectShape(*S, R, s, H, TIE_CANONICAL);
                }
            }
            orc_hit& o = out[i]; memset(&o, 0, sizeof(o));
            o.t = H.Time; o.shape = H.ShapeIndex; o.prim = H.PrimitiveIndex;
        }
    });
}

// ---- wavefront renderer (CreateBasicRenderer/Reset/Run, basic.cpp:166-332) ----
void* orc_renderer_create(uint w, uint h) {
    renderer* r = new renderer();
    r->W.Resize(w, h);
    return r;
}
void orc_renderer_destroy(void* p) { delete (renderer*)p; }
void orc_renderer_set_frame_index(void* p, uint f) { ((renderer*)p)->FrameIndex = f; }
uint orc_renderer_frame_index(void* p) { return ((renderer*)p)->FrameIndex; }

struct orc_render_params {
    uint camera_index, render_flags, path_length_limit;
    float path_termination_probability;
    int tie_mode;
    int threads;
    // bounded-sample support for CPU baselines: only tiles with
    // (tile_index % tile_stride) == 0 are processed (1 = all).
    uint tile_stride;
};

static void dispatch_scatter(renderer* r, const scene_view* S, const push_constants& PC, const orc_render_params& P) {
    wavefront_state& W = r->W;
    uint tiles = W.TilesX * W.TilesY;
    std::vector<uint64_t> finished(std::max(1, P.threads), 0);
    parallel_for(tiles, 4, P.threads, [&](uint64_t b, uint64_t e, int t) {
        for (uint64_t tile = b; tile < e; tile++) {
            if (P.tile_stride > 1 && (tile % P.tile_stride) != 0) continue;
            uint tx = (uint)(tile % W.TilesX), ty = (uint)(tile / W.TilesX);
            for (uint ly = 0; ly < 16; ly++) for (uint lx = 0; lx < 16; lx++) {
                uint X = tx * 16 + lx, Y = ty * 16 + ly;
                if (X >= W.Width || Y >= W.Height) continue;
                size_t o = ((size_t)Y * W.Width + X) * 4 + 3;
                float a0 = W.Accum[o];
                ScatterMain(*S, PC, W, X, Y);
                if (!PC.Restart && W.Accum[o] != a0) finished[t]++;
            }
        }
    });
    for (uint64_t f : finished) W.PathsFinished += f;
}

static void dispatch_trace(renderer* r, const scene_view* S, const orc_render_params& P) {
    wavefront_state& W = r->W;
    uint tiles = W.TilesX * W.TilesY;
    int nt = std::max(1, P.threads);
    std::vector<uint64_t> rays(nt, 0), nodes(nt, 0), faces(nt, 0), shapes(nt, 0), mesh(nt, 0);
    parallel_for(tiles, 4, P.threads, [&](uint64_t b, uint64_t e, int t) {
        for (uint64_t tile = b; tile < e; tile++) {
            if (P.tile_stride > 1 && (tile % P.tile_stride) != 0) continue;
            uint tx = (uint)(tile % W.TilesX), ty = (uint)(tile / W.TilesX);
            for (uint l = 0; l < 256; l++) {
                uint X = tx * 16 + (l & 15), Y = ty * 16 + (l >> 4);
                if (X >= W.Width || Y >= W.Height) continue;
                hit H;
                TraceMain(*S, W, (uint)tile * 256 + l, (tie_mode)P.tie_mode, &H);
                rays[t]++; nodes[t] += H.NodeFetches; faces[t] += H.FaceTests; shapes[t] += H.ShapeVisits;
                if (H.ShapeIndex != SHAPE_INDEX_NONE && H.ShapeType == SHAPE_TYPE_MESH_INSTANCE) mesh[t]++;
            }
        }
    });
    for (int t = 0; t < nt; t++) {
        W.RaysTraced += rays[t]; W.NodeFetches += nodes[t]; W.FaceTests += faces[t]; W.ShapeVisits += shapes[t]; W.MeshHits += mesh[t];
    }
}

// ResetBasicRenderer (basic.cpp:285-304): scatter dispatch with Restart=1, seed = FrameIndex.
void orc_renderer_reset(void* p, const scene_view* S, const orc_render_params* P) {
    renderer* r = (renderer*)p;
    push_constants PC{P->camera_index, P->render_flags, P->path_length_limit, P->path_termination_probability, r->FrameIndex, 1u};
    dispatch_scatter(r, S, PC, *P);
}

// RunBasicRenderer (basic.cpp:306-332): FrameIndex += 1; Rounds x {trace, scatter}, same seed.
void orc_renderer_run(void* p, const scene_view* S, const orc_render_params* P, uint rounds) {
    renderer* r = (renderer*)p;
    r->FrameIndex += 1;
    push_constants PC{P->camera_index, P->render_flags, P->path_length_limit, P->path_termination_probability, r->FrameIndex, 0u};
    for (uint k = 0; k < rounds; k++) {
        dispatch_trace(r, S, *P);
        dispatch_scatter(r, S, PC, *P);
    }
}

// Only the trace half of a round (for ray-cast parity on in-flight rays).
void orc_renderer_trace_only(void* p, const scene_view* S, const orc_render_params* P) { dispatch_trace((renderer*)p, S, *P); }

uint orc_renderer_slot_count(void* p) { return ((renderer*)p)->W.SlotCount; }
float* orc_renderer_accum(void* p) { return ((renderer*)p)->W.Accum.data(); }

struct orc_stats { uint64_t rays, paths, node_fetches, face_tests, shape_visits, mesh_hits; };
void orc_renderer_stats(void* p, orc_stats* s) {
    wavefront_state& W = ((renderer*)p)->W;
    s->rays = W.RaysTraced; s->paths = W.PathsFinished; s->node_fetches = W.NodeFetches;
    s->face_tests = W.FaceTests; s->shape_visits = W.ShapeVisits; s->mesh_hits = W.MeshHits;
}

// Named access to the SoA state (tests compare slot state with the CUDA path).
void* orc_renderer_state_array(void* p, const char* name) {
    wavefront_state& W = ((renderer*)p)->W;
#define ARR(N) if (!strcmp(name, #N)) return (void*)W.N.data();
    ARR(OriginX) ARR(OriginY) ARR(OriginZ) ARR(Duration) ARR(PackedVelocity)
    ARR(Time) ARR(TextureU) ARR(TextureV) ARR(ShapeAndMaterialIndex) ARR(PackedNormal) ARR(PackedTangentX)
    ARR(ImagePosition) ARR(NormalizedLambda0) ARR(Throughput0) ARR(Throughput1) ARR(Throughput2) ARR(Throughput3)
    ARR(Probability0) ARR(Probability1) ARR(Probability2) ARR(Probability3) ARR(SampleR) ARR(SampleG) ARR(SampleB)
    ARR(ActiveShapeIndex01) ARR(ActiveShapeIndex23) ARR(PathLength) ARR(PrimitiveIndex)
#undef ARR
    return nullptr;
}

// ---- resolve (RenderSampleBuffer, integrator.cpp:105-159 + resolve.glsl) ----
void orc_resolve(const float* accum, uint w, uint h, float brightness, uint mode, float white, float* out_linear_rgb, uint8_t* out_srgb8) {
    resolve_parameters P{brightness, mode, white};
    for (size_t i = 0; i < (size_t)w * h; i++) {
        vec3 C = ResolvePixel(V4(accum[4 * i], accum[4 * i + 1], accum[4 * i + 2], accum[4 * i + 3]), P);
        if (out_linear_rgb) { out_linear_rgb[3 * i] = C.x; out_linear_rgb[3 * i + 1] = C.y; out_linear_rgb[3 * i + 2] = C.z; }
        if (out_srgb8) {
            out_srgb8[4 * i] = LinearToSRGB8(C.x); out_srgb8[4 * i + 1] = LinearToSRGB8(C.y);
            out_srgb8[4 * i + 2] = LinearToSRGB8(C.z); out_srgb8[4 * i + 3] = 255;
        }
    }
}

// ---- known-answer-test helpers ----
void orc_kat_rng(uint seed, uint n, uint* out_u32, float* out_f32) {
    rng G; G.State = seed;
    for (uint i = 0; i < n; i++) {
        rng H = G;
        uint u = G.Next();
        if (out_u32) out_u32[i] = u;
        if (out_f32) out_f32[i] = H.Uniform();
    }
}
uint orc_kat_pack_unit_vector(float x, float y, float z) { return PackUnitVector(V3(x, y, z)); }
void orc_kat_unpack_unit_vector(uint p, float* out3) { vec3 v = UnpackUnitVector(p); out3[0] = v.x; out3[1] = v.y; out3[2] = v.z; }
uint orc_kat_pack_half2(float x, float y) { return packHalf2x16(V2(x, y)); }
void orc_kat_unpack_half2(uint p, float* out2) { vec2 v = unpackHalf2x16(p); out2[0] = v.x; out2[1] = v.y; }
void orc_kat_observer(float lambda, float* out3) { vec3 v = SampleStandardObserver(lambda); out3[0] = v.x; out3[1] = v.y; out3[2] = v.z; }
float orc_kat_parametric_spectrum(float b0, float b1, float b2, float lambda) { return SampleParametricSpectrum(V3(b0, b1, b2), lambda); }
void orc_kat_observe_d65(float b0, float b1, float b2, float* out3) { vec3 v = ObserveParametricSpectrumUnderD65(V3(b0, b1, b2)); out3[0] = v.x; out3[1] = v.y; out3[2] = v.z; }
void orc_kat_tonemap(float r, float g, float b, uint mode, float white, float* out3) {
    vec3 C = V3(r, g, b);
    if (mode == TONE_MAPPING_MODE_CLAMP) C = V3(gl_clamp(C.x, 0, 1), gl_clamp(C.y, 0, 1), gl_clamp(C.z, 0, 1));
    if (mode == TONE_MAPPING_MODE_REINHARD) C = ToneMapReinhard(C, white);
    if (mode == TONE_MAPPING_MODE_HABLE) C = ToneMapHable(C);
    if (mode == TONE_MAPPING_MODE_ACES) C = ToneMapACES(C);
    out3[0] = C.x; out3[1] = C.y; out3[2] = C.z;
}
float orc_kat_box(const orc_ray* r, float reach, const float* mn, const float* mx) {
    ray R; R.Origin = V3(r->ox, r->oy, r->oz); R.Velocity = V3(r->dx, r->dy, r->dz); R.Duration = r->duration;
    return IntersectBoundingBox(R, reach, V3(mn[0], mn[1], mn[2]), V3(mx[0], mx[1], mx[2]));
}
void orc_kat_camera_ray(const packed_camera* cam, float nx, float ny, uint seed, orc_ray* out) {
    rng G; G.State = seed;
    ray R = GenerateCameraRay(*cam, V2(nx, ny), G);
    out->ox = R.Origin.x; out->oy = R.Origin.y; out->oz = R.Origin.z;
    out->dx = R.Velocity.x; out->dy = R.Velocity.y; out->dz = R.Velocity.z; out->duration = R.Duration;
}
uint8_t orc_kat_srgb8(float l) { return LinearToSRGB8(l); }

// Preview visualiser (preview_render.glsl): RGBA float image (alpha 1), shape index per pixel, pick result.
void orc_preview(const scene_view* S, const preview_params* P, float* rgba, uint* shapes, uint* hit_shape, int threads) {
    const uint64_t n = (uint64_t)P->RenderSizeX * P->RenderSizeY;
    parallel_for(n, 1024, threads, [&](uint64_t b, uint64_t e, int) {
        for (uint64_t i = b; i < e; i++) {
            const uint x = (uint)(i % P->RenderSizeX), y = (uint)(i / P->RenderSizeX);
            uint shape;
            vec3 c = PreviewPixel(*S, *P, x, y, &shape);
            rgba[4 * i] = c.x; rgba[4 * i + 1] = c.y; rgba[4 * i + 2] = c.z; rgba[4 * i + 3] = 1.0f;
            if (shapes) shapes[i] = shape;
            if (hit_shape && x == P->MouseX && y == P->MouseY) *hit_shape = shape;
        }
    });
}

}  // extern "C"
