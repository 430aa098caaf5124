// sm_100a kernels of the path-tracing integrator and their launchers.
//
//   k_restart   K3  basic_scatter.glsl main, Restart=1  (clear film, start one path per pixel)
//   k_trace     K1  basic_trace.glsl                    (closest hit per slot)
//   k_scatter   K2  basic_scatter.glsl main, Restart=0  (shade, accumulate, regenerate)
//   k_mega      K1+K2 fused over `rounds` with the path state in registers
//   k_resolve   K4  resolve.glsl
//   k_raycast   K1/K5 on caller rays (parity harness, preview_render.glsl:96-109)
//
// Compiled with --fmad=false (see ptb_vec.cuh).
#include "ptb_kernels.h"
#include "ptb_resolve.cuh"
#include "ptb_wide.cuh"

namespace ptb {

enum { MODE_BVH2_FAITHFUL = 0, MODE_BVH2_CANONICAL = 1, MODE_WIDE = 2 };

template <int MODE>
__device__ __forceinline__ HitId closest_hit(const DeviceScene& s, const Ray& r) {
    if (MODE == MODE_BVH2_FAITHFUL) return closest_hit_bvh2<false>(s, r);
    if (MODE == MODE_BVH2_CANONICAL) return closest_hit_bvh2<true>(s, r);
    return closest_hit_wide(s, r);
}

static int mode_of(unsigned traversal, unsigned tie) {
    if (traversal == PTB_TRAVERSAL_WIDE) return MODE_WIDE;
    return tie == PTB_TIE_CANONICAL ? MODE_BVH2_CANONICAL : MODE_BVH2_FAITHFUL;
}

// slot -> pixel (tile-major 16x16, basic_scatter.glsl:324-328)
__device__ __forceinline__ bool slot_pixel(const RenderConsts& pc, u32 slot, u32& x, u32& y) {
    const u32 tile = slot >> 8, l = slot & 255u;
    const u32 tx = tile % pc.tiles_x, ty = tile / pc.tiles_x;
    x = tx * 16u + (l & 15u);
    y = ty * 16u + (l >> 4);
    return slot < pc.slot_count && x < pc.width && y < pc.height;
}

// ---- state records <-> register structs ----
__device__ __forceinline__ Ray load_ray(const WavefrontState& st, u32 slot) {
    const float4 r = st.ray[slot];
    Ray ray;
    ray.origin = mk3(r.x, r.y, r.z);
    ray.velocity = unpack_unit_vector(__float_as_uint(r.w));
    ray.duration = PTB_HIT_TIME_LIMIT;
    return ray;
}
__device__ __forceinline__ void store_ray(const WavefrontState& st, u32 slot, const Ray& ray) {
    st.ray[slot] = make_float4(ray.origin.x, ray.origin.y, ray.origin.z, __uint_as_float(pack_unit_vector(ray.velocity)));
}
__device__ __forceinline__ PathState load_path(const WavefrontState& st, u32 slot) {
    const uint4 a = st.path0[slot];
    const float4 t = st.path1[slot], p = st.path2[slot], c = st.path3[slot];
    PathState path;
    path.pixel = a.x; path.lambda0 = __uint_as_float(a.y);
    unpack_active(a.z, a.w, path.active);
    path.throughput = mk4(t.x, t.y, t.z, t.w);
    path.probability = mk4(p.x, p.y, p.z, p.w);
    path.sample = mk3(c.x, c.y, c.z);
    path.length = __float_as_uint(c.w);
    return path;
}
__device__ __forceinline__ void store_path(const WavefrontState& st, u32 slot, const PathState& path) {
    u32 w01, w23;
    pack_active(path.active, w01, w23);
    st.path0[slot] = make_uint4(path.pixel, __float_as_uint(path.lambda0), w01, w23);
    st.path1[slot] = make_float4(path.throughput.x, path.throughput.y, path.throughput.z, path.throughput.w);
    st.path2[slot] = make_float4(path.probability.x, path.probability.y, path.probability.z, path.probability.w);
    st.path3[slot] = make_float4(path.sample.x, path.sample.y, path.sample.z, __uint_as_float(path.length));
}

// ---- K3: restart ----
__global__ void __launch_bounds__(256) k_restart(const DeviceScene s, const RenderConsts pc, const WavefrontState st, float4* __restrict__ accum) {
    const u32 slot = blockIdx.x * 256u + threadIdx.x;
    u32 x, y;
    if (!slot_pixel(pc, slot, x, y)) return;
    Rng g;
    g.state = y * 65537u + x + pc.random_seed * 277803737u;
    PathState path; Ray ray;
    generate_new_path(s, pc, g, x, y, path, ray);
    store_ray(st, slot, ray);
    store_path(st, slot, path);
    accum[y * pc.width + x] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (threadIdx.x == 0) st.tile_paths[blockIdx.x] = 0u;
}

// ---- K1: trace ----
template <int MODE>
__global__ void __launch_bounds__(128) k_trace(const DeviceScene s, const RenderConsts pc, const WavefrontState st) {
    const u32 slot = blockIdx.x * 128u + threadIdx.x;
    u32 x, y;
    if (!slot_pixel(pc, slot, x, y)) return;
    const Ray ray = load_ray(st, slot);
    const HitId h = closest_hit<MODE>(s, ray);
    if (h.shape == PTB_SHAPE_NONE) {
        st.hit0[slot] = make_uint4(__float_as_uint(PTB_HIT_TIME_LIMIT), 0xFFFFFFFFu, 0u, 0u);
        st.hit1[slot] = make_uint4(0u, 0u, 0xFFFFFFFFu, 0u);
        return;
    }
    const Surface sf = finish_hit(s, h);
    const PackedHit p = pack_hit(h, sf);
    st.hit0[slot] = make_uint4(__float_as_uint(p.t), p.shape_material, p.normal, p.tangent);
    st.hit1[slot] = make_uint4(__float_as_uint(p.u), __float_as_uint(p.v), h.prim, h.shape_type);
}

// ---- K2: scatter ----
__global__ void __launch_bounds__(256) k_scatter(const DeviceScene s, const RenderConsts pc, const WavefrontState st, float4* __restrict__ accum) {
    const u32 slot = blockIdx.x * 256u + threadIdx.x;
    u32 x, y;
    bool finished = false;
    if (slot_pixel(pc, slot, x, y)) {
        PathState path = load_path(st, slot);
        Ray ray = load_ray(st, slot);
        const uint4 h0 = st.hit0[slot];
        PackedHit p;
        p.t = __uint_as_float(h0.x); p.shape_material = h0.y; p.normal = h0.z; p.tangent = h0.w; p.u = p.v = 0.0f;
        if (h0.y != 0xFFFFFFFFu) { const uint4 h1 = st.hit1[slot]; p.u = __uint_as_float(h1.x); p.v = __uint_as_float(h1.y); }
        const SurfaceHit hit = unpack_hit(p, ray);
        finished = scatter_invocation(s, pc, x, y, path, ray, hit, accum);
        store_ray(st, slot, ray);
        store_path(st, slot, path);
    }
    const int c = __syncthreads_count(finished ? 1 : 0);
    if (threadIdx.x == 0 && c) st.tile_paths[blockIdx.x] += (unsigned)c;
}

// ---- megakernel: rounds x {trace, scatter}, state in registers ----
template <int MODE>
__global__ void __launch_bounds__(128) k_mega(const DeviceScene s, RenderConsts pc, const WavefrontState st, float4* __restrict__ accum,
                                               u32 rounds, u32 seed_stride) {
    const u32 slot = blockIdx.x * 128u + threadIdx.x;
    u32 x, y;
    u32 finished = 0;
    if (slot_pixel(pc, slot, x, y)) {
        PathState path = load_path(st, slot);
        Ray ray = load_ray(st, slot);
        HitId h; Surface sf;
        h.shape = PTB_SHAPE_NONE; h.prim = 0xFFFFFFFFu; h.shape_type = 0; h.t = PTB_HIT_TIME_LIMIT;
        for (u32 k = 0; k < rounds; k++) {
            h = closest_hit<MODE>(s, ray);
            PackedHit p = packed_miss();
            if (h.shape != PTB_SHAPE_NONE) { sf = finish_hit(s, h); p = pack_hit(h, sf); }
            const SurfaceHit hit = unpack_hit(p, ray);
            if (scatter_invocation(s, pc, x, y, path, ray, hit, accum)) finished++;
            // the same quantisation the wavefront state applies between rounds
            ray.velocity = unpack_unit_vector(pack_unit_vector(ray.velocity));
            u32 w01, w23;
            pack_active(path.active, w01, w23);
            unpack_active(w01, w23, path.active);
            pc.random_seed += seed_stride;
        }
        store_ray(st, slot, ray);
        store_path(st, slot, path);
    }
    finished = __reduce_add_sync(0xFFFFFFFFu, finished);
    if ((threadIdx.x & 31u) == 0 && finished) atomicAdd(&st.tile_paths[slot >> 8], finished);
}

// ---- K4: resolve ----
__global__ void __launch_bounds__(256) k_resolve(const float4* __restrict__ accum, u32 count, ptb_resolve_params params,
                                                  uchar4* __restrict__ out_srgb8, float* __restrict__ out_linear) {
    const u32 i = blockIdx.x * 256u + threadIdx.x;
    if (i >= count) return;
    const f3 c = resolve_pixel(accum[i], params.Brightness, params.ToneMappingMode, params.ToneMappingWhiteLevel);
    if (out_srgb8) out_srgb8[i] = make_uchar4((unsigned char)linear_to_srgb8(c.x), (unsigned char)linear_to_srgb8(c.y), (unsigned char)linear_to_srgb8(c.z), 255);
    if (out_linear) { out_linear[3 * (size_t)i] = c.x; out_linear[3 * (size_t)i + 1] = c.y; out_linear[3 * (size_t)i + 2] = c.z; }
}

// ---- ray cast on caller rays ----
template <int MODE>
__global__ void __launch_bounds__(128) k_raycast(const DeviceScene s, const ptb_ray* __restrict__ rays, size_t n, ptb_hit* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * 128u + threadIdx.x;
    if (i >= n) return;
    const ptb_ray in = rays[i];
    Ray r;
    r.origin = mk3(in.ox, in.oy, in.oz); r.velocity = mk3(in.dx, in.dy, in.dz); r.duration = in.duration;
    const HitId h = closest_hit<MODE>(s, r);
    ptb_hit o;
    memset(&o, 0, sizeof o);
    o.t = h.t; o.shape = h.shape; o.prim = 0xFFFFFFFFu;
    if (h.shape != PTB_SHAPE_NONE) {
        const Surface sf = finish_hit(s, h);
        o.prim = h.prim; o.material = sf.material; o.shape_type = h.shape_type;
        o.nx = sf.normal.x; o.ny = sf.normal.y; o.nz = sf.normal.z;
        o.tx = sf.tangent_x.x; o.ty = sf.tangent_x.y; o.tz = sf.tangent_x.z;
        o.u = sf.uv.x; o.v = sf.uv.y;
        o.bx = h.coords.x; o.by = h.coords.y; o.bz = h.coords.z;
    }
    out[i] = o;
}

// ---- launchers ----
cudaError_t launch_restart(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum, cudaStream_t stream) {
    k_restart<<<pc.slot_count / 256u, 256, 0, stream>>>(s, pc, st, accum);
    return cudaGetLastError();
}

cudaError_t launch_trace(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, unsigned traversal, unsigned tie, cudaStream_t stream) {
    const unsigned grid = pc.slot_count / 128u;
    switch (mode_of(traversal, tie)) {
        case MODE_BVH2_FAITHFUL: k_trace<MODE_BVH2_FAITHFUL><<<grid, 128, 0, stream>>>(s, pc, st); break;
        case MODE_BVH2_CANONICAL: k_trace<MODE_BVH2_CANONICAL><<<grid, 128, 0, stream>>>(s, pc, st); break;
        default: k_trace<MODE_WIDE><<<grid, 128, 0, stream>>>(s, pc, st); break;
    }
    return cudaGetLastError();
}

cudaError_t launch_scatter(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum, cudaStream_t stream) {
    k_scatter<<<pc.slot_count / 256u, 256, 0, stream>>>(s, pc, st, accum);
    return cudaGetLastError();
}

cudaError_t launch_megakernel(const DeviceScene& s, const RenderConsts& pc, const WavefrontState& st, float4* accum,
                              unsigned traversal, unsigned tie, unsigned rounds, unsigned seed_stride, cudaStream_t stream) {
    const unsigned grid = pc.slot_count / 128u;
    switch (mode_of(traversal, tie)) {
        case MODE_BVH2_FAITHFUL: k_mega<MODE_BVH2_FAITHFUL><<<grid, 128, 0, stream>>>(s, pc, st, accum, rounds, seed_stride); break;
        case MODE_BVH2_CANONICAL: k_mega<MODE_BVH2_CANONICAL><<<grid, 128, 0, stream>>>(s, pc, st, accum, rounds, seed_stride); break;
        default: k_mega<MODE_WIDE><<<grid, 128, 0, stream>>>(s, pc, st, accum, rounds, seed_stride); break;
    }
    return cudaGetLastError();
}

cudaError_t launch_resolve(const float4* accum, unsigned width, unsigned height, ptb_resolve_params params,
                           uchar4* out_srgb8, float* out_linear_rgb, cudaStream_t stream) {
    const unsigned count = width * height;
    k_resolve<<<(count + 255u) / 256u, 256, 0, stream>>>(accum, count, params, out_srgb8, out_linear_rgb);
    return cudaGetLastError();
}

cudaError_t launch_raycast(const DeviceScene& s, const ptb_ray* rays, size_t n, unsigned traversal, unsigned tie, ptb_hit* out, cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    const unsigned grid = (unsigned)((n + 127) / 128);
    switch (mode_of(traversal, tie)) {
        case MODE_BVH2_FAITHFUL: k_raycast<MODE_BVH2_FAITHFUL><<<grid, 128, 0, stream>>>(s, rays, n, out); break;
        case MODE_BVH2_CANONICAL: k_raycast<MODE_BVH2_CANONICAL><<<grid, 128, 0, stream>>>(s, rays, n, out); break;
        default: k_raycast<MODE_WIDE><<<grid, 128, 0, stream>>>(s, rays, n, out); break;
    }
    return cudaGetLastError();
}

}  // namespace ptb
