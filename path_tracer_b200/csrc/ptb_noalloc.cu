// ptb_raycast_noalloc: ptb_raycast for callers that must not allocate (code that runs where cudaMalloc is forbidden:
// tight serving loops, harnesses that ban it, regions under CUDA graph capture).  No counterpart in the
// reference.
//
// ptb_raycast stages the caller's host rays and hits through two cudaMalloc'd device buffers per call and uses the
// context's counter for the persistent kernel's dynamic ray fetch.  Here all three live in ONE file-scope
// `static __device__` array of a fixed size: PTB_NOALLOC_MAX_RAYS (65 536) rays per call, 256 + 65 536 x (28 + 64)
// bytes = 5.75 MiB, held by every GPU that loads libptb.so until the process exits.  The entry point calls no
// cudaMalloc / cudaFree / allocating library routine; the kernel is the same k_trace_coop<RaycastJob> (or k_raycast)
// ptb_raycast launches, so the hits are bit-identical.
//
// The array is SHARED by every call on a device and keeps the last call's contents.  What must start at zero - the
// ray-fetch counter - is cleared on the stream before each launch (launch_raycast does it); the hit records are fully
// overwritten for the n rays of the call and nothing beyond n is read back.  Two calls that overlap on the GPU (two
// contexts = two streams on one device, or two host threads) RACE on the array: serialise them.
//
// In host code `g_noalloc_scratch` is a host-side stand-in for the symbol, not a device address; the device address
// is looked up once per device with cudaGetSymbolAddress while that device is current, and that pointer is what the
// kernel gets.
#include <mutex>

#include "ptb_internal.h"

using namespace ptb;

#define PTB_NOALLOC_COUNTER_BYTES 256
#define PTB_NOALLOC_SCRATCH_BYTES (PTB_NOALLOC_COUNTER_BYTES + (size_t)PTB_NOALLOC_MAX_RAYS * (sizeof(ptb_ray) + sizeof(ptb_hit)))
#define PTB_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_noalloc_scratch[PTB_NOALLOC_SCRATCH_BYTES];

namespace {
std::mutex g_mu;
unsigned char* g_device_ptr[PTB_NOALLOC_MAX_DEVICES];      // per device: address of g_noalloc_scratch there

// Set-up, once per device, with that device current.
int scratch_of_current_device(int device, unsigned char** out) {
    if (device < 0 || device >= PTB_NOALLOC_MAX_DEVICES) return api_fail(PTB_E_UNSUPPORTED, "ptb_raycast_noalloc: device index beyond 63");
    std::lock_guard<std::mutex> l(g_mu);
    if (!g_device_ptr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return api_cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_scratch)");
        g_device_ptr[device] = (unsigned char*)p;
    }
    *out = g_device_ptr[device];
    return PTB_OK;
}
}  // namespace

extern "C" {

size_t ptb_raycast_noalloc_max_rays(void) { return PTB_NOALLOC_MAX_RAYS; }

int ptb_raycast_noalloc(ptb_scene* s, const ptb_ray* host_rays, size_t n, uint32_t traversal, uint32_t tie, ptb_hit* host_hits) {
    if (!s || (n && (!host_rays || !host_hits))) return api_fail(PTB_E_INVALID, "ptb_raycast_noalloc: NULL argument");
    if (n > (size_t)PTB_NOALLOC_MAX_RAYS)
        return api_fail(PTB_E_INVALID, "ptb_raycast_noalloc: " + std::to_string(n) + " rays exceed the static scratch of " + std::to_string(PTB_NOALLOC_MAX_RAYS) +
                                       " rays per call (split the batch, or use ptb_raycast)");
    if (traversal > PTB_TRAVERSAL_WIDE || tie > PTB_TIE_CANONICAL) return api_fail(PTB_E_INVALID, "ptb_raycast_noalloc: unknown enum value");
    if (traversal == PTB_TRAVERSAL_WIDE && tie != PTB_TIE_CANONICAL) return api_fail(PTB_E_INVALID, "ptb_raycast_noalloc: the wide traversal needs PTB_TIE_CANONICAL");
    if (traversal == PTB_TRAVERSAL_WIDE && s->ds.shape_count > 0 && s->wide_info.status != PTB_OK) return api_fail(PTB_E_UNSUPPORTED, "ptb_raycast_noalloc: wide BVH unavailable");
    if (n == 0) return PTB_OK;
    // (A host-built tree parked by PTB_BUILDER_AUTO is NOT swapped in here - the swap frees the old tree.  The tree in use
    //  gives the same hits.)
    CU(cudaSetDevice(s->ctx->device));
    unsigned char* base = nullptr;
    int rc = scratch_of_current_device(s->ctx->device, &base);
    if (rc) return rc;
    unsigned* counter = (unsigned*)base;
    ptb_ray* d_rays = (ptb_ray*)(base + PTB_NOALLOC_COUNTER_BYTES);
    ptb_hit* d_hits = (ptb_hit*)(base + PTB_NOALLOC_COUNTER_BYTES + (size_t)PTB_NOALLOC_MAX_RAYS * sizeof(ptb_ray));
    cudaStream_t st = s->ctx->stream;
    CU(cudaMemcpyAsync(d_rays, host_rays, n * sizeof(ptb_ray), cudaMemcpyHostToDevice, st));
    CU(launch_raycast(s->ds, d_rays, n, traversal, tie, s->raycast_variant, counter, d_hits, st));      // zeroes the counter on the stream first
    CU(cudaMemcpyAsync(host_hits, d_hits, n * sizeof(ptb_hit), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    return PTB_OK;
}

}  // extern "C"
