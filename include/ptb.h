/*
 * ptb.h — C ABI of the B200-native path-tracing integrator.
 *
 * Drop-in boundary for the trace / scatter / resolve dispatch of
 * samukallio/path-tracer (`src/integrator`).  Each entry point names the
 * reference interface it replaces (file:line relative to the reference root).
 * The scene data model is unchanged: `ptb_scene_view` is a set of
 * {pointer,count} pairs over the arrays `PackSceneData()` produces
 * (src/scene/scene.hpp:347-356), byte-identical to src/scene/scene.hpp:82-173.
 *
 * Conventions
 *   - every function returns 0 on success or a negative PTB_E_* code;
 *     ptb_last_error() returns a thread-local description of the last failure;
 *   - handles are opaque and caller-owned; a renderer borrows its scene and
 *     film (they must outlive it), exactly like basic_renderer borrows
 *     vulkan_scene / vulkan_sample_buffer (src/integrator/basic.cpp:177-178);
 *   - work is enqueued on the context's CUDA stream and returns immediately;
 *     ptb_renderer_sync / downloads / host resolves synchronise;
 *   - there is NO CPU fallback: every call fails with PTB_E_CUDA when no
 *     sm_100 device is usable.
 */
#ifndef PTB_H
#define PTB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PTB_OK            0
#define PTB_E_INVALID    -1   /* bad argument / inconsistent scene arrays */
#define PTB_E_CUDA       -2   /* CUDA runtime error (message in ptb_last_error) */
#define PTB_E_NOMEM      -3
#define PTB_E_UNSUPPORTED -4

/* ---- packed scene records: src/scene/scene.hpp:82-173 (std430) ------------ */
typedef struct { float To[16]; float From[16]; } ptb_packed_transform;               /* :82-86  col-major */
typedef struct { float AtlasPlacementMinimum[2]; float AtlasPlacementMaximum[2];
                 uint32_t AtlasImageIndex, Type, Flags, Unused0; } ptb_packed_texture; /* :90-98  */
typedef struct { int32_t Type; uint32_t MaterialIndex, MeshRootNodeIndex, Pad0;
                 ptb_packed_transform Transform; } ptb_packed_shape;                 /* :102-108 */
typedef struct { float Minimum[3]; uint32_t ChildNodeIndices;
                 float Maximum[3]; uint32_t ShapeIndex; } ptb_packed_shape_node;     /* :112-118 */
typedef struct { float Position0[3]; uint32_t VertexIndex0;
                 float Position1[3]; uint32_t VertexIndex1;
                 float Position2[3]; uint32_t VertexIndex2; } ptb_packed_mesh_face;  /* :122-130 */
typedef struct { uint32_t PackedNormal, PackedUV; } ptb_packed_mesh_vertex;          /* :134-138 */
typedef struct { float Minimum[3]; uint32_t FaceBeginOrNodeIndex;
                 float Maximum[3]; uint32_t FaceEndIndex; } ptb_packed_mesh_node;    /* :142-148 */
typedef struct { float SkyboxMeanDirection[3]; float SkyboxConcentration;
                 float SkyboxSamplingProbability, SkyboxBrightness;
                 uint32_t SkyboxTextureIndex, ShapeCount;
                 float SceneScatterRate; uint32_t Pad[3]; } ptb_packed_scene_globals; /* :152-161 */
typedef struct { uint32_t Model; float FocalLength, ApertureRadius, SensorDistance;
                 float SensorSize[2]; uint32_t Pad[2];
                 ptb_packed_transform Transform; } ptb_packed_camera;                /* :165-173 */

/* scene_dirty_flag: src/scene/scene.hpp:323-333 (same bit values) */
#define PTB_SCENE_DIRTY_GLOBALS        (1u << 0)
#define PTB_SCENE_DIRTY_TEXTURES       (1u << 1)
#define PTB_SCENE_DIRTY_MATERIALS      (1u << 2)
#define PTB_SCENE_DIRTY_SHAPES         (1u << 3)
#define PTB_SCENE_DIRTY_MESHES         (1u << 4)
#define PTB_SCENE_DIRTY_CAMERAS        (1u << 5)
#define PTB_SCENE_DIRTY_SKYBOX_TEXTURE (1u << 6)
#define PTB_SCENE_DIRTY_ALL            0xFFFFFFFFu

/* The host arrays of `scene` after PackSceneData (src/scene/scene.hpp:347-356).
 * Every pointer is HOST memory; ptb_scene_update copies what the dirty flags
 * select, so the arrays may be freed on return. */
typedef struct {
    const ptb_packed_texture*     Textures;      uint32_t TextureCount;      uint32_t _r0;
    const uint32_t*               MaterialData;  uint32_t MaterialWordCount; uint32_t _r1;  /* 32 words / slot */
    const ptb_packed_shape*       Shapes;        uint32_t ShapeCount;        uint32_t _r2;
    const ptb_packed_shape_node*  ShapeNodes;    uint32_t ShapeNodeCount;    uint32_t _r3;
    const ptb_packed_mesh_face*   MeshFaces;     uint32_t MeshFaceCount;     uint32_t _r4;
    const ptb_packed_mesh_vertex* MeshVertices;  uint32_t MeshVertexCount;   uint32_t _r5;
    const ptb_packed_mesh_node*   MeshNodes;     uint32_t MeshNodeCount;     uint32_t _r6;
    const ptb_packed_camera*      Cameras;       uint32_t CameraCount;       uint32_t _r7;
    const float* const*           AtlasLayers;   uint32_t AtlasLayerCount;   /* scene::Images, RGBA32F */
    uint32_t                      AtlasWidth;                                /* 4096 in the reference (scene.cpp:1122) */
    uint32_t                      AtlasHeight;   uint32_t _r8;
    const ptb_packed_scene_globals* Globals;
} ptb_scene_view;

/* render_flag: src/integrator/integrator.hpp:6-10 */
#define PTB_RENDER_FLAG_ACCUMULATE    (1u << 0)
#define PTB_RENDER_FLAG_SAMPLE_JITTER (1u << 1)

/* tone_mapping_mode: src/integrator/integrator.hpp:12-19 */
#define PTB_TONE_MAPPING_MODE_CLAMP    0u
#define PTB_TONE_MAPPING_MODE_REINHARD 1u
#define PTB_TONE_MAPPING_MODE_HABLE    2u
#define PTB_TONE_MAPPING_MODE_ACES     3u

/* resolve_parameters: src/integrator/integrator.hpp:43-48 */
typedef struct { float Brightness; uint32_t ToneMappingMode; float ToneMappingWhiteLevel; } ptb_resolve_params;

/* Integrator structure (north_star: both are implemented). */
#define PTB_INTEGRATOR_WAVEFRONT  0u   /* trace kernel + scatter kernel per round, SoA state in HBM */
#define PTB_INTEGRATOR_MEGAKERNEL 1u   /* one persistent kernel, path state in registers across rounds */

/* Traversal backend for the trace step. */
#define PTB_TRAVERSAL_BVH2_REFERENCE 0u /* the packed two-level BVH2 as-is, reference visiting order */
#define PTB_TRAVERSAL_WIDE           1u /* compressed BVH8 re-flattened at ptb_scene_update */

/* Trace-step kernel for PTB_TRAVERSAL_WIDE.  0 = one thread per slot; 1..4 = persistent warps
 * with dynamic ray fetch; 5..MAX = the same with cooperative leaf tests (numbered tuning points,
 * see ptb_kernels.cu).  All variants produce identical hit records. */
#define PTB_TRACE_VARIANT_SIMPLE  0u
#define PTB_TRACE_VARIANT_DEFAULT 14u
#define PTB_TRACE_VARIANT_MAX     19u

/* Scatter-step structure of the wavefront integrator. */
#define PTB_SCATTER_AUTO   0u  /* queues when the scene mixes material types or has media, else direct */
#define PTB_SCATTER_DIRECT 1u  /* one kernel over all slots in slot order */
#define PTB_SCATTER_QUEUES 2u  /* per-round material queues {miss, diffuse, metal, other} + one specialised kernel each */
#define PTB_SCATTER_DIRECT_STAGED 3u  /* DIRECT as resident blocks over 256-slot tiles whose state is staged through shared memory by TMA bulk
                                         copies, two tiles deep (same results; measured 19 % slower than DIRECT on C2, kept for A/B) */

/* Tie rule for equal hit times. */
#define PTB_TIE_FAITHFUL  0u  /* reference rules; needs PTB_TRAVERSAL_BVH2_REFERENCE */
#define PTB_TIE_CANONICAL 1u  /* lowest (t, ShapeIndex, PrimitiveIndex) */

/* The public tunables of basic_renderer (src/integrator/basic.hpp:19-25) that
 * the application pokes before Reset/Run (src/application/application.cpp:104-107),
 * plus the build's own knobs. */
typedef struct {
    uint32_t CameraIndex;
    uint32_t RenderFlags;
    uint32_t PathLengthLimit;            /* 0 = unlimited (reference behaviour: value is never read) */
    float    PathTerminationProbability;
    uint32_t Integrator;                 /* PTB_INTEGRATOR_* */
    uint32_t Traversal;                  /* PTB_TRAVERSAL_*  */
    uint32_t TieMode;                    /* PTB_TIE_*        */
    uint32_t TraceVariant;               /* PTB_TRACE_VARIANT_*: kernel shape of the wide-BVH trace step */
    uint32_t ScatterMode;                /* PTB_SCATTER_* */
    uint32_t Pipelines;                  /* 0 = automatic (one), 1 = one stream, 2 = the film's slots as two halves on two
                                            streams (kernel tails overlap the other half's next kernel; same results) */
    uint32_t Reserved[2];                /* must be 0 */
} ptb_render_params;

typedef struct {
    uint64_t rays_traced;      /* trace invocations on live slots  (Mrays/s numerator)  */
    uint64_t paths_finished;   /* accumulator alpha increments     (Mpaths/s numerator) */
    uint64_t kernel_launches;  /* kernels launched by this renderer since creation      */
    uint64_t frame_index;      /* basic_renderer::FrameIndex */
} ptb_stats;

/* Test / preview ray cast record (K1 on caller rays; preview_render.glsl:96-109). */
typedef struct { float ox, oy, oz, dx, dy, dz, duration; } ptb_ray;
typedef struct {
    float    t;
    uint32_t shape, prim, material, shape_type;   /* shape == 0xFFFFFFFF: miss */
    float    nx, ny, nz, tx, ty, tz, u, v;
    float    bx, by, bz;
} ptb_hit;

/* Preview visualiser: the push constants of src/application/preview_render.glsl:48-60. */
#define PTB_PREVIEW_MODE_BASE_COLOR        0u
#define PTB_PREVIEW_MODE_BASE_COLOR_SHADED 1u
#define PTB_PREVIEW_MODE_NORMAL            2u
#define PTB_PREVIEW_MODE_MATERIAL_INDEX    3u
#define PTB_PREVIEW_MODE_PRIMITIVE_INDEX   4u
#define PTB_PREVIEW_MODE_MESH_COMPLEXITY   5u   /* BVH2 traversal: mesh-node iterations; wide: leaf primitives tested */
#define PTB_PREVIEW_MODE_SCENE_COMPLEXITY  6u   /* BVH2 traversal: shape-node iterations; wide: BVH8 nodes visited   */
typedef struct {
    ptb_packed_transform CameraTransform;
    uint32_t RenderMode;
    float    Brightness;
    uint32_t SelectedShapeIndex;
    uint32_t RenderSizeX, RenderSizeY;
    uint32_t MouseX, MouseY;
    uint32_t Traversal;                  /* PTB_TRAVERSAL_* (ties are canonical) */
} ptb_preview_params;

typedef struct ptb_context  ptb_context;
typedef struct ptb_scene    ptb_scene;
typedef struct ptb_film     ptb_film;
typedef struct ptb_renderer ptb_renderer;

const char* ptb_last_error(void);
const char* ptb_version(void);

/* Replaces CreateVulkan/DestroyVulkan as far as the integrator needs a device
 * (src/core/vulkan.cpp:1816): binds one CUDA device (one process per GPU). */
int  ptb_context_create(int cuda_device, ptb_context** out);
void ptb_context_destroy(ptb_context* ctx);
int  ptb_context_sync(ptb_context* ctx);
/* cudaStream_t of the context, for callers that share buffers with it. */
void* ptb_context_stream(ptb_context* ctx);

/* CreateVulkanScene / UpdateVulkanScene / DestroyVulkanScene
 * (src/scene/scene.hpp:440-442, scene.cpp:1643-2006).  Update synchronises the
 * stream first (the reference does vkDeviceWaitIdle, scene.cpp:1704), copies the
 * arrays selected by dirty_flags and re-flattens the wide BVH when meshes change. */
int  ptb_scene_create(ptb_context* ctx, ptb_scene** out);
int  ptb_scene_update(ptb_scene* scene, const ptb_scene_view* view, uint32_t dirty_flags);
void ptb_scene_destroy(ptb_scene* scene);
/* Result of the scene flattening done inside ptb_scene_update: the compressed
 * 8-wide BVH that replaces the reference's nested BVH2 walks on the hot path
 * (scene.glsl.inc:336-399,468-520).  status != 0: only the reference-layout
 * traversal (PTB_TRAVERSAL_BVH2_REFERENCE) is usable for this scene. */
typedef struct {
    int32_t  status;
    uint32_t node_count, prim_count, plane_count, depth, bvh2_depth;
    double   sah_cost, build_seconds;
    uint32_t builder;                    /* PTB_BUILDER_* that produced the tree */
    float    device_build_ms;            /* GPU builder: CUDA-event time of its kernels; 0 for the host builder */
} ptb_wide_info;
int  ptb_scene_wide_info(ptb_scene* scene, ptb_wide_info* out);
/* Which builder flattens the scene at the next ptb_scene_update that touches shapes or meshes.
 * The reference has one builder, the single-threaded binned-SAH recursion BuildMeshNode
 * (src/scene/scene.cpp:435-599), run at import time; here the traversal tree is rebuilt from the
 * packed arrays, so the choice is a property of the device scene:
 *   SAH_HOST  threaded binned SAH + SAH-optimal 8-wide collapse on the host (best traversal speed; default)
 *   LBVH_GPU  Morton-order radix tree + SAH-optimal collapse as CUDA kernels (milliseconds per 10 M triangles;
 *             slowest to traverse).
 *   PLOC_GPU  locally-ordered agglomerative clustering on the Morton order as CUDA kernels (tens of
 *             milliseconds per 10 M triangles; between the two in traversal speed).
 *   AUTO      both: first image after milliseconds, full traversal speed once the host build is done.
 * Hits are bit-identical whichever builder ran: the tree only culls.
 * If the GPU builder cannot represent a scene (tree deeper than the traversal stack) the host builder runs. */
#define PTB_BUILDER_SAH_HOST 0u
#define PTB_BUILDER_LBVH_GPU 1u
#define PTB_BUILDER_PLOC_GPU 2u
#define PTB_BUILDER_AUTO     3u   /* PLOC_GPU at once; a host thread then builds the SAH_HOST tree from the device copies of
                                   * the packed arrays and the next call that traverses the scene swaps it in (one stream
                                   * synchronise).  Results do not change at the swap.  ptb_scene_wide_info reports the
                                   * builder of the tree in use. */
int  ptb_scene_set_builder(ptb_scene* scene, uint32_t builder);
/* Debug/test: copies the flattened tree to the host (20 words per node, 12 floats per leaf primitive; layout in
 * csrc/ptb_wide.cuh) so that a checker can verify it structurally.  Counts come from ptb_scene_wide_info. */
int  ptb_scene_wide_download(ptb_scene* scene, uint32_t* host_nodes, size_t node_words, float* host_prims, size_t prim_floats);

/* Spectral upsampling (src/core/spectrum.hpp:13-21).
 * ptb_spectrum_table_build replaces BuildParametricSpectrumTableForSRGB (src/core/spectrum.cpp:365-411,
 * called from CreateScene when sRGBSpectrumTable.dat is missing, src/scene/scene.cpp:918-927): it fits the
 * float[3][64][64][64][3] table on the GPU (one warp per column) and writes it to host memory in the layout
 * SaveParametricSpectrumTable dumps (spectrum.cpp:413-424), so the file stays interchangeable.
 * ptb_spectrum_upsample is the texel conversion of the atlas packer (src/scene/scene.cpp:1187-1212) over
 * `texels` RGBA32F values: texture_type 0 raw, 1 reflectance with alpha, 2 radiance. */
int  ptb_spectrum_table_build(ptb_context* ctx, float* host_table, size_t count_floats, float* device_ms);
int  ptb_spectrum_upsample(ptb_context* ctx, const float* host_table, size_t count_floats, const float* host_rgba_in,
                           float* host_rgba_out, size_t texels, uint32_t texture_type);

/* CreateSampleBuffer / DestroySampleBuffer (src/integrator/integrator.hpp:51-53):
 * the RGBA32F accumulator (rgb = sum of CIE XYZ, a = sample count).
 * `external_device_rgba` (may be NULL) lets the caller own the device buffer,
 * e.g. a tensor that a collective reduces. */
int  ptb_film_create(ptb_context* ctx, uint32_t width, uint32_t height, void* external_device_rgba, ptb_film** out);
void ptb_film_destroy(ptb_film* film);
void* ptb_film_device_accum(ptb_film* film);
int  ptb_film_download_accum(ptb_film* film, float* host_rgba, size_t count_floats);
int  ptb_film_upload_accum(ptb_film* film, const float* host_rgba, size_t count_floats);
/* RenderSampleBuffer (src/integrator/integrator.hpp:55-60, resolve.glsl:112-130):
 * normalise by alpha, XYZ->sRGB, Brightness, tone-map; then the sRGB OETF the
 * reference gets from its _SRGB swapchain.  Either output may be NULL.
 * host_srgb8 is W*H*4 bytes (RGBA), host_linear_rgb is W*H*3 floats (pre-OETF). */
int  ptb_film_resolve(ptb_film* film, const ptb_resolve_params* params, uint8_t* host_srgb8, float* host_linear_rgb);
/* Same, output left in device memory (W*H*4 bytes), no synchronisation. */
int  ptb_film_resolve_device(ptb_film* film, const ptb_resolve_params* params, void* device_srgb8);

/* CreateBasicRenderer / DestroyBasicRenderer (src/integrator/basic.hpp:28-29). */
int  ptb_renderer_create(ptb_context* ctx, ptb_scene* scene, ptb_film* film, ptb_renderer** out);
void ptb_renderer_destroy(ptb_renderer* r);
/* The caller-poked public fields of basic_renderer (basic.hpp:19-25). */
int  ptb_renderer_set_params(ptb_renderer* r, const ptb_render_params* params);
int  ptb_renderer_set_frame_index(ptb_renderer* r, uint32_t frame_index);
/* ResetBasicRenderer (basic.cpp:285-304): clear the film, start one path per pixel. */
int  ptb_renderer_reset(ptb_renderer* r);
/* RunBasicRenderer (basic.cpp:306-332): FrameIndex += 1, then `rounds` x {trace, scatter}
 * sharing that seed. */
int  ptb_renderer_run(ptb_renderer* r, uint32_t rounds);
/* `frames` consecutive RunBasicRenderer(…, 1) calls — what Update() issues in
 * steady state, one per displayed frame (application.cpp:114) — as one call. */
int  ptb_renderer_run_frames(ptb_renderer* r, uint32_t frames);
int  ptb_renderer_sync(ptb_renderer* r);
/* Measurement aid (no reference counterpart: the reference has no timers):
 * same work as ptb_renderer_run_frames, with CUDA events around every kernel launch
 * on the context stream; returns the summed device time per kernel. */
typedef struct {
    float    trace_ms, scatter_ms, mega_ms;
    uint32_t trace_launches, scatter_launches, mega_launches;
} ptb_kernel_times;
int  ptb_renderer_profile_frames(ptb_renderer* r, uint32_t frames, ptb_kernel_times* out);
/* Measurement aid: ONE extra launch of the default wide-BVH trace kernel, instrumented, over the renderer's current
 * rays (the trace step is idempotent: it rewrites the same hit records).  Counts what the kernel's own data layout
 * makes it fetch: 80-byte BVH8 nodes and 48-byte leaf primitives, plus 48 B of wavefront state per ray (16 B ray in,
 * 2 x 16 B hit out) - the "own-layout algorithmic bytes" of bench.py's roofline - and the lane occupancy of its phases. */
typedef struct {
    uint64_t rays, node_steps, prim_tests, warp_iterations, flushes;
    uint32_t node_bytes, prim_bytes, state_bytes, _pad;
} ptb_trace_stats;
int  ptb_renderer_trace_stats(ptb_renderer* r, ptb_trace_stats* out);
int  ptb_renderer_stats(ptb_renderer* r, ptb_stats* out);
/* Debug/test: download one named SoA state array (names of basic.glsl.inc:23-59). */
int  ptb_renderer_download_state(ptb_renderer* r, const char* name, void* host, size_t bytes);
uint32_t ptb_renderer_slot_count(ptb_renderer* r);

/* Trace() on caller-supplied host rays (K1/K5; the ray-cast parity harness). */
int  ptb_raycast(ptb_scene* scene, const ptb_ray* host_rays, size_t n, uint32_t traversal, uint32_t tie_mode, ptb_hit* host_hits);
/* Same with both buffers already in device memory; enqueued on the context stream, no synchronisation. */
int  ptb_raycast_device(ptb_scene* scene, const ptb_ray* device_rays, size_t n, uint32_t traversal, uint32_t tie_mode, ptb_hit* device_hits);

/* ptb_raycast without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same kernel and bit-identical hits as ptb_raycast; the device staging of the rays and hits and the kernel's ray-fetch
 * counter live in one file-scope static __device__ array (csrc/ptb_noalloc.cu) of PTB_NOALLOC_MAX_RAYS rays = 5.75 MiB,
 * resident on every GPU that loads the library until the process exits.  n > PTB_NOALLOC_MAX_RAYS fails with
 * PTB_E_INVALID.  Calls no cudaMalloc / cudaFree / allocating library routine.  All calls on a device share the array:
 * the counter is cleared on the stream before each launch, the hits of the call's n rays are overwritten; calls that
 * overlap on the GPU (two contexts or two host threads on one device) race - serialise them.  Synchronises. */
#define PTB_NOALLOC_MAX_RAYS 65536
int    ptb_raycast_noalloc(ptb_scene* scene, const ptb_ray* host_rays, size_t n, uint32_t traversal, uint32_t tie_mode, ptb_hit* host_hits);
size_t ptb_raycast_noalloc_max_rays(void);

/* ---- preview visualiser (src/application/preview_render.glsl:96-178; application.cpp preview pass) ----
 * One camera ray per pixel, closest hit, colour by RenderMode; no graphics API: the image is written to
 * host_rgba (RenderSizeX*RenderSizeY*4 floats, alpha 1).  host_shapes (optional) receives Hit.ShapeIndex
 * per pixel, *hit_shape the shape under (MouseX, MouseY) — the reference's pick query (QuerySSBO). */
int  ptb_preview_render(ptb_scene* scene, const ptb_preview_params* params, float* host_rgba, uint32_t* host_shapes, uint32_t* hit_shape);

/* ---- multi-GPU groups: sample-split rendering of one image on several GPUs of one NVLink box -------------------
 * No reference counterpart: the reference binds one VkPhysicalDevice (src/core/vulkan.cpp:1485-1506).  Partial
 * films add (rgb = sum of XYZ samples, a = sample count; resolve divides by a, resolve.glsl:117-118), so every rank
 * renders a disjoint block of RandomSeed values (basic.cpp:315,323) into its own film and ONE exchange step combines
 * them.  A group is a communicator over NCCL (loaded with dlopen at first use: libptb.so has no NCCL dependency);
 * every group call is enqueued on its context's stream and is collective - all ranks make the same call in the same
 * order.  Two ways to form a group:
 *   - one process per GPU (torchrun / mpirun): rank 0 calls ptb_group_unique_id, the launcher hands the 128 bytes to
 *     the other ranks, every rank calls ptb_group_create;
 *   - one process driving n GPUs (bin/ptb_headless --devices): ptb_group_create_local, then ONE HOST THREAD PER
 *     DEVICE issues that device's calls (collectives block until every rank has joined). */
typedef struct { char bytes[128]; } ptb_group_id;
typedef struct ptb_group ptb_group;
int  ptb_group_unique_id(ptb_group_id* out);
int  ptb_group_create(ptb_context* ctx, const ptb_group_id* id, int rank, int world, ptb_group** out);
int  ptb_group_create_local(ptb_context* const* contexts, int count, ptb_group** out_groups /* [count], rank k on contexts[k] */);
void ptb_group_destroy(ptb_group* group);
int  ptb_group_rank(const ptb_group* group);
int  ptb_group_world(const ptb_group* group);
int  ptb_group_barrier(ptb_group* group);                 /* all ranks' streams have reached this point; synchronises */
int  ptb_group_nccl_version(int* version);
/* Seed schedule of the sample split: first FrameIndex of `rank` when every rank renders frames_per_rank frames after
 * its reset (rank r owns base + r * (frames_per_rank + 1) + [0, frames_per_rank]: disjoint and gap-free over ranks). */
uint32_t ptb_group_first_frame(int rank, int world, uint32_t frames_per_rank, uint32_t base);

/* The root's device scene - packed arrays, atlas layers and the flattened BVH8 - sent device-to-device over NVLink
 * (one ncclBroadcast per array).  Ranks other than `root` pass a freshly created, never updated scene: they neither
 * pack nor build.  (UpdateVulkanScene, scene.cpp:1704-1960, is what runs on the root only.) */
int  ptb_scene_replicate(ptb_scene* scene, ptb_group* group, int root);

/* Sum of all ranks' films into `sum` on `root` (root < 0: on every rank).  `sum` may be the rank's own film (in
 * place) or another film of the same size; ranks that receive nothing may pass NULL. */
int  ptb_film_reduce(ptb_film* film, ptb_group* group, int root, ptb_film* sum);
/* The exchange step fused with the resolve that follows it (RenderSampleBuffer on the combined film).  The resolved
 * sRGB8 frame lands in host_srgb8 (W*H*4 bytes) on the root; other ranks pass NULL.  `sum` (root only, optional)
 * also receives the float sum.  Synchronises. */
#define PTB_REDUCE_NCCL        0u  /* ncclReduce of the float film (16 B/pixel) to the root + k_resolve there */
#define PTB_REDUCE_NCCL_SLICED 1u  /* ncclReduceScatter -> each rank resolves its 1/n slice -> ncclAllGather of sRGB8 (4 B/pixel
                                      on the second leg); needs W*H divisible by the group size and no `sum`, else = NCCL */
#define PTB_REDUCE_P2P_FUSED   2u  /* one kernel per rank over NVLink peer memory: reads its slice of every rank's film, adds in
                                      rank order (bit-identical to a sequential sum), tone-maps, writes sRGB8 (and the float sum)
                                      straight into the root's buffers.  Films must be library-owned (no external_device_rgba)
                                      when the ranks are separate processes. */
int  ptb_film_reduce_resolve(ptb_film* film, ptb_group* group, int root, uint32_t mode, const ptb_resolve_params* params,
                             ptb_film* sum, uint8_t* host_srgb8);
typedef struct { float exchange_ms; uint32_t mode; uint64_t bytes_in; } ptb_reduce_times;   /* device time of the last reduce_resolve on this rank */
int  ptb_group_last_reduce(ptb_group* group, ptb_reduce_times* out);

#ifdef __cplusplus
}
#endif
#endif /* PTB_H */
