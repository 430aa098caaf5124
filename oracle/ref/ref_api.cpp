// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// oracle/_ref/libref.so: the reference's own trace / scatter shader sources (rewritten into
// oracle/_ref/gen/ by transpile.py, read from /root/reference/src where they lie) compiled as C++
// against ref_glsl.h, plus the dispatch loops of src/integrator/basic.cpp:285-332.  This is what the
// restatement in oracle/sh_*.h is checked against, and what tests/golden/ref_*.npz are generated
// from (tests/golden/make_ref_golden.py), so that the CUDA path is compared with outputs of the
// reference's code on the GPU box too, where /root/reference does not exist.
// Single-threaded by design: the shaders' interface blocks are plain globals here.
#include <cstdio>
#include <cstring>

#include "ptb.h"            // ptb_scene_view: the packed arrays, byte-identical to scene.hpp:82-173 (layout only; nothing is linked)
#include "ref_glsl.h"

namespace glsl {
uvec3 gl_GlobalInvocationID, gl_WorkGroupID, gl_LocalInvocationID;
#include "integrator/basic_trace.glsl"
#include "integrator/basic_scatter.glsl"

static_assert(sizeof(packed_texture) == 32 && sizeof(packed_shape) == 144 && sizeof(packed_shape_node) == 32, "std430 sizes");
static_assert(sizeof(packed_mesh_face) == 48 && sizeof(packed_mesh_vertex) == 8 && sizeof(packed_mesh_node) == 32, "std430 sizes");
static_assert(sizeof(packed_camera) == 160 && sizeof(packed_transform) == 128, "std430 sizes");
}  // namespace glsl

using namespace glsl;

extern "C" {

struct ref_ray { float ox, oy, oz, dx, dy, dz, duration; };
struct ref_hit {
    float t;
    uint32_t shape, prim, material, shape_type;
    float nx, ny, nz, tx, ty, tz, u, v;
    float bx, by, bz;
    uint32_t scene_complexity, mesh_complexity;
};
struct ref_params { uint32_t CameraIndex, RenderFlags, PathLengthLimit; float PathTerminationProbability; };

// UpdateVulkanScene (scene.cpp:1698-1960): binds the packed arrays to the shaders' interface blocks.
void ref_bind_scene(const ptb_scene_view* v) {
#include "ref_bind_scene.inc"
}

// Trace() (scene.glsl.inc:522-611) on caller rays.
void ref_raycast(const ref_ray* rays, uint64_t n, ref_hit* out) {
    for (uint64_t i = 0; i < n; i++) {
        ray R;
        R.Origin = vec3(rays[i].ox, rays[i].oy, rays[i].oz);
        R.Velocity = vec3(rays[i].dx, rays[i].dy, rays[i].dz);
        R.Duration = rays[i].duration;
        hit H = Trace(R);
        ref_hit o;
        std::memset(&o, 0, sizeof o);
        o.t = H.Time; o.shape = H.ShapeIndex;
        if (H.ShapeIndex != SHAPE_INDEX_NONE) {
            o.prim = H.PrimitiveIndex; o.material = H.MaterialIndex; o.shape_type = H.ShapeType;
            o.nx = H.Normal.x; o.ny = H.Normal.y; o.nz = H.Normal.z;
            o.tx = H.TangentX.x; o.ty = H.TangentX.y; o.tz = H.TangentX.z;
            o.u = H.UV.x; o.v = H.UV.y;
            o.bx = H.PrimitiveCoordinates.x; o.by = H.PrimitiveCoordinates.y; o.bz = H.PrimitiveCoordinates.z;
        } else {
            o.prim = 0xFFFFFFFFu;
        }
        o.scene_complexity = H.SceneComplexity; o.mesh_complexity = H.MeshComplexity;
        out[i] = o;
    }
}

// CreateSampleBuffer (integrator.cpp:15-98): the accumulator image, owned by the caller (w x h x RGBA32F).
void ref_bind_film(float* rgba, uint32_t w, uint32_t h) {
    SampleAccumulatorImage.rgba = rgba; SampleAccumulatorImage.width = (int)w; SampleAccumulatorImage.height = (int)h;
}

static void push(const ref_params* p, uint32_t seed, uint32_t restart) {
    CameraIndex = p->CameraIndex; RenderFlags = p->RenderFlags; PathLengthLimit = p->PathLengthLimit;
    PathTerminationProbability = p->PathTerminationProbability; RandomSeed = seed; Restart = restart;
}
// A bounded, shardable sample of the dispatch (bench.py's reference arm: one process per host core, each with its
// own copy of this library's globals): only work groups whose linear index g has g % modulus == residue run.
static uint32_t g_modulus = 1, g_residue = 0;
static uint64_t g_rays = 0;
void ref_set_groups(uint32_t modulus, uint32_t residue) { g_modulus = modulus ? modulus : 1; g_residue = residue; }
uint64_t ref_rays_traced() { return g_rays; }
int ref_slot_capacity() { return TRACE_COUNT; }

// vkCmdDispatch of basic_scatter.glsl over 16x16 work groups (basic.cpp:101-118).
static void dispatch_scatter() {
    const uint gx = ((uint)SampleAccumulatorImage.width + 15u) / 16u, gy = ((uint)SampleAccumulatorImage.height + 15u) / 16u;
    for (uint wy = 0; wy < gy; wy++) for (uint wx = 0; wx < gx; wx++) {
        if ((wy * gx + wx) % g_modulus != g_residue) continue;
        for (uint ly = 0; ly < 16; ly++) for (uint lx = 0; lx < 16; lx++) {
            gl_WorkGroupID = uvec3(wx, wy, 0); gl_LocalInvocationID = uvec3(lx, ly, 0);
            gl_GlobalInvocationID = uvec3(wx * 16 + lx, wy * 16 + ly, 0);
            basic_scatter_main();
        }
    }
}
// basic_trace.glsl over the slots that hold a path (the reference dispatches all TRACE_COUNT slots, basic.cpp:32-60;
// slots outside the film hold no path and their results are never read).
static void dispatch_trace() {
    const uint W = (uint)SampleAccumulatorImage.width, H = (uint)SampleAccumulatorImage.height;
    const uint gx = W / 16u, gy = (H + 15u) / 16u;                     // the slot formula of basic_scatter.glsl:325-329
    for (uint g = 0; g < gx * gy; g++) {
        if (g % g_modulus != g_residue) continue;
        const uint wy = g / gx;
        for (uint l = 0; l < 256u; l++) {
            const uint i = g * 256u + l;
            if (i >= (uint)TRACE_COUNT || wy * 16u + l / 16u >= H) continue;
            gl_GlobalInvocationID = uvec3(i, 0, 0); gl_WorkGroupID = uvec3(i / 256u, 0, 0); gl_LocalInvocationID = uvec3(i % 256u, 0, 0);
            basic_trace_main();
            g_rays++;
        }
    }
}
// ResetBasicRenderer / RunBasicRenderer (basic.cpp:285-332); the caller keeps FrameIndex like basic_renderer does.
void ref_reset(const ref_params* p, uint32_t frame_index) { push(p, frame_index, 1u); dispatch_scatter(); }
void ref_run(const ref_params* p, uint32_t frame_index, uint32_t rounds) {
    push(p, frame_index, 0u);
    for (uint32_t k = 0; k < rounds; k++) { dispatch_trace(); dispatch_scatter(); }
}

// Wavefront state arrays by the names of basic.glsl.inc:23-59.
void* ref_state_array(const char* name) {
#define A(B, N) if (!std::strcmp(name, #N)) return (void*)B.N;
    A(TraceBuffer, OriginX) A(TraceBuffer, OriginY) A(TraceBuffer, OriginZ) A(TraceBuffer, PackedVelocity) A(TraceBuffer, Duration)
    A(TraceBuffer, Time) A(TraceBuffer, ShapeAndMaterialIndex) A(TraceBuffer, PackedNormal) A(TraceBuffer, PackedTangentX)
    A(TraceBuffer, TextureU) A(TraceBuffer, TextureV)
    A(PathBuffer, ImagePosition) A(PathBuffer, NormalizedLambda0) A(PathBuffer, Throughput0) A(PathBuffer, Throughput1)
    A(PathBuffer, Throughput2) A(PathBuffer, Throughput3) A(PathBuffer, Probability0) A(PathBuffer, Probability1)
    A(PathBuffer, Probability2) A(PathBuffer, Probability3) A(PathBuffer, SampleR) A(PathBuffer, SampleG) A(PathBuffer, SampleB)
    A(PathBuffer, ActiveShapeIndex01) A(PathBuffer, ActiveShapeIndex23)
#undef A
    return nullptr;
}

// Small pieces for known-answer checks of the shim against the restatement.
uint32_t ref_pack_unit_vector(float x, float y, float z) { return PackUnitVector(vec3(x, y, z)); }
void ref_unpack_unit_vector(uint32_t p, float* o) { vec3 v = UnpackUnitVector(p); o[0] = v.x; o[1] = v.y; o[2] = v.z; }
void ref_observer(float lambda, float* o) { vec3 v = SampleStandardObserver(lambda); o[0] = v.x; o[1] = v.y; o[2] = v.z; }
void ref_rng(uint32_t seed, uint32_t n, float* o) { RandomState = seed; for (uint32_t i = 0; i < n; i++) o[i] = Random0To1(); }

}  // extern "C"
