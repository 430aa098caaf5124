// Host-visible launch interface of the sm_100a kernels (ptb_kernels.cu).
// Called only by the C ABI implementation (ptb_api.cu).
#pragma once
#include <cuda_runtime.h>
#include "ptb_scatter.cuh"

namespace ptb {

// Wavefront state, one record group per path slot (slot = 16x16-tile-major pixel
// order, basic_scatter.glsl:324-328 with tiles_x = ceil(W/16)).  The reference
// keeps 26 separate 4-byte arrays (basic.glsl.inc:23-59); here the same fields
// are grouped into 16-byte records so every access is one coalesced LDG.128 /
// STG.128 per thread.  Ray.Duration is not stored: every stored ray has
// HIT_TIME_LIMIT (basic_scatter.glsl:307, scene.glsl.inc:617).
struct WavefrontState {
    float4* ray;      // origin.xyz, oct-packed direction (bits)
    uint4*  hit0;     // t (bits), shape<<16|material (0xFFFFFFFF = miss), primitive index, shape type
    uint4*  hit1;     // primitive coordinates xyz (barycentrics / object-space point), unused
                      // (normal, tangent and UV are derived from these in the scatter kernel, where all
                      //  lanes of a warp do it together, instead of by whichever lane finishes traversal)
    uint4*  path0;    // pixel index, lambda0 (bits), active shapes 0|1, active shapes 2|3
    float4* path1;    // throughput at the 4 wavelengths
    float4* path2;    // probability at the 4 wavelengths
    float4* path3;    // sample XYZ, path length (bits)
    unsigned int* ray_counter;  // dynamic ray fetch of the persistent trace kernel; zeroed by restart/scatter
    unsigned int* tile_paths;   // per 256-slot tile: paths finished since reset (u32 wraps are summed by the host in u64 deltas)
};

struct LaunchStats { unsigned int launches; };

// 2 x uint4 per face = {n0, uv0, n1, uv1}, {n2, uv2, 0, 0} (DeviceScene::face_attrs)
cudaError_t launch_face_attrs(const ptb_packed_mesh_face* faces, const ptb_packed_mesh_vertex* vertices, uint4* out, unsigned n, cudaStream_t stream);
cudaError_t launch_restart(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum, cudaStream_t stream);
// variant 0: one thread per slot; variant >= 1: persistent warps with dynamic ray fetch (wide traversal only),
// tuning point (refill threshold, postpone threshold) selected by the variant number.
cudaError_t launch_trace(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, unsigned traversal, unsigned tie,
                         unsigned variant, cudaStream_t stream);
// The default cooperative trace kernel with work counters (5 x u64, see k_trace_coop): same hits, not timed.
cudaError_t launch_trace_stats(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, unsigned long long* stats, cudaStream_t stream);
// Debug/test: closest-hit attributes of the stored hits as {packed normal, packed tangent, u, v} per slot.
cudaError_t launch_hit_attributes(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, uint4* out, cudaStream_t stream);
// staged: the TMA-staged persistent kernel (k_scatter_staged, an experiment kept for A/B); else k_scatter
cudaError_t launch_scatter(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum, bool staged, cudaStream_t stream);
// Scatter through per-material queues (classify + one specialised kernel per class); `queue` holds scatter_queue_bytes().
size_t scatter_queue_bytes(unsigned slot_count);
unsigned scatter_queue_launches(const DeviceScene& s);
cudaError_t launch_scatter_queues(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum, unsigned* queue,
                                  cudaStream_t stream);
// `rounds` x {trace, scatter} with the path state in registers; seed = pc.random_seed + k * seed_stride for round k.
cudaError_t launch_megakernel(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum,
                              unsigned traversal, unsigned tie, unsigned rounds, unsigned seed_stride, cudaStream_t stream);
struct PreviewConsts;
// Preview visualiser: RGBA32F image, optional shape index per pixel, shape under the mouse pixel.
cudaError_t launch_preview(const DeviceScene& s, const PreviewConsts& p, unsigned traversal, unsigned tie, float4* rgba, unsigned* shapes,
                           unsigned* hit_shape, cudaStream_t stream);
cudaError_t launch_resolve(const float4* accum, unsigned width, unsigned height, ptb_resolve_params params,
                           uchar4* out_srgb8, float* out_linear_rgb, cudaStream_t stream);
// counter: device u32 (zeroed by the launcher) for the persistent variant; may be null for variant 0.
cudaError_t launch_raycast(const DeviceScene& s, const ptb_ray* rays, size_t n, unsigned traversal, unsigned tie, unsigned variant,
                           unsigned* counter, ptb_hit* out, cudaStream_t stream);

}  // namespace ptb
