// Multi-GPU rendering behind the C ABI (include/ptb.h, "groups"): sample-split rendering of one
// image on several B200s of one NVLink / NVSwitch box, scene replicated per GPU.
//
// The reference is single-device (it picks one VkPhysicalDevice, src/core/vulkan.cpp:1485-1506), so
// this is the build's own component.  What makes it possible is the film's algebra: the accumulator
// is additive (rgb = sum of XYZ samples, a = sample count) and resolve divides by a
// (resolve.glsl:117-118), so partial films of disjoint seed blocks simply add.
//
//   ptb_group_*            communicators: one process per GPU (unique id handed over by the host's
//                          launcher) or one process driving several GPUs, one host thread each
//   ptb_scene_replicate    the FLATTENED scene (packed arrays + BVH8 + atlas) built once on the root
//                          and sent device-to-device, instead of N host builds
//   ptb_film_reduce        sum of the float films (ncclReduce / ncclAllReduce)
//   ptb_film_reduce_resolve   the exchange step fused with the resolve that follows it, three ways:
//       NCCL          ncclReduce of 16 B/pixel to the root, k_resolve there
//       NCCL_SLICED   ncclReduceScatter -> every rank resolves its 1/n slice -> ncclAllGather of sRGB8
//       P2P_FUSED     ONE kernel per rank over NVLink peer memory: loads its slice of every rank's
//                     film (16-byte coalesced peer loads), sums in rank order (a fixed order: the result is
//                     bit-identical to a sequential host sum, which NCCL's ring / tree order is not),
//                     tone-maps and stores sRGB8 - and optionally the float sum - into the root's buffers
//
// NCCL is loaded at first use (dlopen of libnccl.so.2, the copy already in the process if there is
// one), so libptb.so itself has no NCCL dependency and single-GPU hosts never touch it.
#include <dlfcn.h>
#include <nccl.h>
#include <unistd.h>

#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "ptb_internal.h"
#include "ptb_resolve.cuh"

using namespace ptb;

namespace {

struct Nccl {
    void* handle = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommInitAll) CommInitAll = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclReduce) Reduce = nullptr;
    decltype(&ncclReduceScatter) ReduceScatter = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclBroadcast) Broadcast = nullptr;
    std::string error;
};

Nccl* nccl() {
    static Nccl n;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) { n.handle = dlopen(nm, RTLD_NOW | RTLD_NOLOAD); if (n.handle) break; }     // the copy torch already mapped, if any
        if (!n.handle && std::getenv("PTB_NCCL_LIB")) n.handle = dlopen(std::getenv("PTB_NCCL_LIB"), RTLD_NOW | RTLD_GLOBAL);
        for (const char* nm : names) { if (!n.handle) n.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); }
        if (!n.handle) { n.error = std::string("cannot load libnccl.so.2: ") + dlerror(); return; }
#define SYM(name) n.name = (decltype(n.name))dlsym(n.handle, "nccl" #name); if (!n.name) { n.error = "libnccl has no symbol nccl" #name; return; }
        SYM(GetUniqueId) SYM(CommInitRank) SYM(CommInitAll) SYM(CommDestroy) SYM(GetErrorString) SYM(GetVersion)
        SYM(AllReduce) SYM(Reduce) SYM(ReduceScatter) SYM(AllGather) SYM(Broadcast)
#undef SYM
    });
    return n.error.empty() ? &n : nullptr;
}

int nccl_fail(ncclResult_t r, const char* what) {
    Nccl* n = nccl();
    return api_fail(PTB_E_CUDA, std::string(what) + ": NCCL " + (n ? n->GetErrorString(r) : "unavailable"));
}
#define NC(x) do { ncclResult_t r_ = (x); if (r_ != ncclSuccess) return nccl_fail(r_, #x); } while (0)

}  // namespace

#define PTB_MAX_GROUP 16

// What a rank tells its peers about one film (exchanged once per (film, group) with an all-gather).
struct PeerRecord {
    long long pid;
    int device, external;
    void* accum; void* srgb;
    cudaIpcMemHandle_t accum_handle, srgb_handle;
    unsigned width, height;
    unsigned pad[2];
};

struct FilmPeers {                     // per (film, role, group): every rank's film as this rank can address it
    ptb_film* film = nullptr;
    int role = 0;                       // 0: the films being summed; 1: the destination of the float sum (only the root's entry is used)
    float4* accum[PTB_MAX_GROUP] = {};
    uchar4* srgb[PTB_MAX_GROUP] = {};
    std::vector<void*> opened;          // cudaIpcOpenMemHandle results to close
};

struct ptb_group {
    ptb_context* ctx;
    ncclComm_t comm;
    int rank, world;
    float* d_token;                    // stream-ordered barrier: a 1-float all-reduce
    unsigned char* d_xchg;             // world x sizeof(PeerRecord) staging for the peer-pointer exchange
    float* slice_sum; uchar4* slice_srgb; size_t slice_pixels;      // NCCL_SLICED scratch
    std::vector<FilmPeers*> films;
    ptb_reduce_times last;
};

struct PeerFilms { const float4* accum[PTB_MAX_GROUP]; };

// P2P_FUSED: pixels [begin, end) of the image - this rank's slice.  Every load is a coalesced 16-byte access, local for
// one of the `world` films and over NVLink for the others; all of them are issued before the first add.
template <int WORLD>
__global__ void __launch_bounds__(256) k_reduce_resolve_p2p(const PeerFilms peers, int world, u32 begin, u32 end, ptb_resolve_params params,
                                                             uchar4* __restrict__ out_srgb, float4* __restrict__ out_sum) {
    const u32 i = begin + blockIdx.x * 256u + threadIdx.x;
    if (i >= end) return;
    float4 v[WORLD];
#pragma unroll
    for (int k = 0; k < WORLD; k++) if (k < world) v[k] = peers.accum[k][i];
    float4 a = v[0];
#pragma unroll
    for (int k = 1; k < WORLD; k++) if (k < world) { a.x += v[k].x; a.y += v[k].y; a.z += v[k].z; a.w += v[k].w; }      // rank order: a fixed association
    if (out_sum) out_sum[i] = a;
    const f3 c = resolve_pixel(a, params.Brightness, params.ToneMappingMode, params.ToneMappingWhiteLevel);
    out_srgb[i] = make_uchar4((unsigned char)linear_to_srgb8(c.x), (unsigned char)linear_to_srgb8(c.y), (unsigned char)linear_to_srgb8(c.z), 255);
}

static int group_barrier(ptb_group* g) {
    Nccl* n = nccl();
    NC(n->AllReduce(g->d_token, g->d_token, 1, ncclFloat32, ncclSum, g->comm, g->ctx->stream));
    return PTB_OK;
}

static int group_finish(ptb_group* g, ptb_context* ctx, ncclComm_t comm, int rank, int world, ptb_group** out) {
    g->ctx = ctx; g->comm = comm; g->rank = rank; g->world = world;
    g->d_token = nullptr; g->d_xchg = nullptr; g->slice_sum = nullptr; g->slice_srgb = nullptr; g->slice_pixels = 0;
    std::memset(&g->last, 0, sizeof g->last);
    CU(cudaSetDevice(ctx->device));
    CU(cudaMalloc((void**)&g->d_token, 256));
    CU(cudaMemsetAsync(g->d_token, 0, 256, ctx->stream));
    CU(cudaMalloc((void**)&g->d_xchg, (size_t)world * sizeof(PeerRecord)));
    *out = g;
    return PTB_OK;
}

// Every rank's film as this rank can address it (same process: the pointer itself after cudaDeviceEnablePeerAccess;
// another process: cudaIpcOpenMemHandle).  Collective: all ranks call it for their film at the same point.
static FilmPeers* film_peers_lookup(ptb_group* g, ptb_film* film, int role) {
    for (FilmPeers* p : g->films) if (p->film == film && p->role == role) return p;
    return nullptr;
}
static int film_peers_exchange(ptb_group* g, ptb_film* film, int role, FilmPeers** out) {
    for (size_t i = 0; i < g->films.size(); i++) if (g->films[i]->film == film && g->films[i]->role == role) {      // replaced
        for (void* q : g->films[i]->opened) cudaIpcCloseMemHandle(q);
        delete g->films[i];
        g->films.erase(g->films.begin() + (long)i);
        break;
    }
    Nccl* n = nccl();
    cudaStream_t st = g->ctx->stream;
    const size_t pixels = (size_t)film->width * film->height;
    if (!film->srgb) CU(cudaMalloc((void**)&film->srgb, pixels * sizeof(uchar4)));
    PeerRecord mine;
    std::memset(&mine, 0, sizeof mine);
    mine.pid = (long long)getpid(); mine.device = g->ctx->device; mine.external = film->external ? 1 : 0;
    mine.accum = film->accum; mine.srgb = film->srgb; mine.width = film->width; mine.height = film->height;
    if (!film->external) {
        CU(cudaIpcGetMemHandle(&mine.accum_handle, film->accum));
        CU(cudaIpcGetMemHandle(&mine.srgb_handle, film->srgb));
    }
    CU(cudaMemcpyAsync(g->d_xchg + (size_t)g->rank * sizeof(PeerRecord), &mine, sizeof mine, cudaMemcpyHostToDevice, st));
    NC(n->AllGather(g->d_xchg + (size_t)g->rank * sizeof(PeerRecord), g->d_xchg, sizeof(PeerRecord), ncclChar, g->comm, st));
    std::vector<PeerRecord> all(g->world);
    CU(cudaMemcpyAsync(all.data(), g->d_xchg, (size_t)g->world * sizeof(PeerRecord), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    FilmPeers* p = new FilmPeers();
    p->film = film; p->role = role;
    std::string why;
    for (int k = 0; k < g->world && why.empty(); k++) {
        const PeerRecord& r = all[k];
        if (r.width != film->width || r.height != film->height) { why = "films of different sizes"; break; }
        if (k == g->rank) { p->accum[k] = film->accum; p->srgb[k] = film->srgb; continue; }
        if (r.pid == mine.pid) {
            int can = 0;
            cudaDeviceCanAccessPeer(&can, g->ctx->device, r.device);
            if (!can) { why = "no peer access between devices " + std::to_string(g->ctx->device) + " and " + std::to_string(r.device); break; }
            cudaError_t e = cudaDeviceEnablePeerAccess(r.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { why = std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e); break; }
            cudaGetLastError();
            p->accum[k] = (float4*)r.accum; p->srgb[k] = (uchar4*)r.srgb;
        } else {
            if (r.external) { why = "a peer film lives in caller-owned memory (external_device_rgba): not shareable between processes"; break; }
            void* a = nullptr; void* s = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&a, r.accum_handle, cudaIpcMemLazyEnablePeerAccess);
            if (e == cudaSuccess) { p->opened.push_back(a); e = cudaIpcOpenMemHandle(&s, r.srgb_handle, cudaIpcMemLazyEnablePeerAccess); }
            if (e != cudaSuccess) { cudaGetLastError(); why = std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e); break; }
            p->opened.push_back(s);
            p->accum[k] = (float4*)a; p->srgb[k] = (uchar4*)s;
        }
    }
    // every rank must reach the same verdict, or the ranks would take different code paths and hang
    float ok = why.empty() ? 0.0f : 1.0f, total = 0.0f;
    CU(cudaMemcpyAsync(g->d_token + 1, &ok, sizeof ok, cudaMemcpyHostToDevice, st));
    NC(n->AllReduce(g->d_token + 1, g->d_token + 1, 1, ncclFloat32, ncclSum, g->comm, st));
    CU(cudaMemcpyAsync(&total, g->d_token + 1, sizeof total, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (total != 0.0f) {
        for (void* q : p->opened) cudaIpcCloseMemHandle(q);
        delete p;
        return api_fail(PTB_E_UNSUPPORTED, "peer-memory film exchange unavailable: " + (why.empty() ? std::string("refused by another rank") : why));
    }
    g->films.push_back(p);
    *out = p;
    return PTB_OK;
}

// ---- scene replication ----
// The root's device scene - packed arrays, atlas layers, BVH8 - sent device-to-device; the other ranks never pack, never
// build.  Header first (counts, globals, features, tree statistics), then one ncclBroadcast per array.
struct ReplicaHeader {
    uint64_t textures, materials, shapes, shape_nodes, faces, vertices, mesh_nodes, cameras, layers, layer_floats;
    uint64_t wide_nodes, wide_prims, wide_planes;
    DeviceScene ds;
    ptb_wide_info info;
    uint32_t material_features, builder;
};

template <class T>
static int bcast_buf(Nccl* n, ptb_group* g, int root, DevBuf<T>& b, size_t count) {
    if (g->rank != root) CU(b.alloc(count));
    if (count) NC(n->Broadcast(b.p, b.p, count * sizeof(T), ncclChar, root, g->comm, g->ctx->stream));
    return PTB_OK;
}

extern "C" {

int ptb_group_unique_id(ptb_group_id* out) {
    if (!out) return api_fail(PTB_E_INVALID, "ptb_group_unique_id: NULL argument");
    Nccl* n = nccl();
    if (!n) return api_fail(PTB_E_UNSUPPORTED, "ptb_group_unique_id: NCCL unavailable");
    static_assert(sizeof(ptb_group_id) == sizeof(ncclUniqueId), "ptb_group_id is an ncclUniqueId");
    ncclUniqueId id;
    NC(n->GetUniqueId(&id));
    std::memcpy(out, &id, sizeof id);
    return PTB_OK;
}

int ptb_group_create(ptb_context* ctx, const ptb_group_id* id, int rank, int world, ptb_group** out) {
    if (!ctx || !id || !out || world < 1 || world > PTB_MAX_GROUP || rank < 0 || rank >= world) return api_fail(PTB_E_INVALID, "ptb_group_create: bad argument");
    *out = nullptr;
    Nccl* n = nccl();
    if (!n) return api_fail(PTB_E_UNSUPPORTED, "ptb_group_create: NCCL unavailable");
    CU(cudaSetDevice(ctx->device));
    ncclUniqueId nid;
    std::memcpy(&nid, id, sizeof nid);
    ncclComm_t comm = nullptr;
    NC(n->CommInitRank(&comm, world, nid, rank));
    ptb_group* g = new ptb_group();
    int rc = group_finish(g, ctx, comm, rank, world, out);
    if (rc) { n->CommDestroy(comm); delete g; }
    return rc;
}

int ptb_group_create_local(ptb_context* const* ctxs, int count, ptb_group** out) {
    if (!ctxs || !out || count < 1 || count > PTB_MAX_GROUP) return api_fail(PTB_E_INVALID, "ptb_group_create_local: bad argument");
    Nccl* n = nccl();
    if (!n) return api_fail(PTB_E_UNSUPPORTED, "ptb_group_create_local: NCCL unavailable");
    std::vector<int> devices(count);
    for (int k = 0; k < count; k++) { if (!ctxs[k]) return api_fail(PTB_E_INVALID, "ptb_group_create_local: NULL context"); devices[k] = ctxs[k]->device; out[k] = nullptr; }
    for (int a = 0; a < count; a++) for (int b = a + 1; b < count; b++) if (devices[a] == devices[b]) return api_fail(PTB_E_INVALID, "ptb_group_create_local: one context per device");
    std::vector<ncclComm_t> comms(count);
    NC(n->CommInitAll(comms.data(), count, devices.data()));
    for (int k = 0; k < count; k++) {
        ptb_group* g = new ptb_group();
        int rc = group_finish(g, ctxs[k], comms[k], k, count, &out[k]);
        if (rc) return rc;
    }
    return PTB_OK;
}

void ptb_group_destroy(ptb_group* g) {
    if (!g) return;
    cudaSetDevice(g->ctx->device);
    cudaStreamSynchronize(g->ctx->stream);
    for (FilmPeers* p : g->films) { for (void* q : p->opened) cudaIpcCloseMemHandle(q); delete p; }
    if (Nccl* n = nccl()) n->CommDestroy(g->comm);
    cudaFree(g->d_token); cudaFree(g->d_xchg); cudaFree(g->slice_sum); cudaFree(g->slice_srgb);
    delete g;
}

int ptb_group_rank(const ptb_group* g) { return g ? g->rank : -1; }
int ptb_group_world(const ptb_group* g) { return g ? g->world : 0; }

int ptb_group_barrier(ptb_group* g) {
    if (!g) return api_fail(PTB_E_INVALID, "ptb_group_barrier: NULL group");
    CU(cudaSetDevice(g->ctx->device));
    int rc = group_barrier(g);
    if (rc) return rc;
    CU(cudaStreamSynchronize(g->ctx->stream));
    return PTB_OK;
}

uint32_t ptb_group_first_frame(int rank, int world, uint32_t frames_per_rank, uint32_t base) {
    (void)world;
    return base + (uint32_t)rank * (frames_per_rank + 1u);      // rank r owns seeds base + r (F + 1) + [0, F]: reset uses the first
}

int ptb_group_nccl_version(int* version) {
    Nccl* n = nccl();
    if (!n || !version) return api_fail(PTB_E_UNSUPPORTED, "ptb_group_nccl_version: NCCL unavailable");
    NC(n->GetVersion(version));
    return PTB_OK;
}

// ---- film exchange ----
int ptb_film_reduce(ptb_film* film, ptb_group* g, int root, ptb_film* sum) {
    if (!film || !g || root >= g->world) return api_fail(PTB_E_INVALID, "ptb_film_reduce: bad argument");
    const bool receives = root < 0 || root == g->rank;
    if (receives && !sum) return api_fail(PTB_E_INVALID, "ptb_film_reduce: the receiving rank needs a destination film (it may be the source film)");
    if (receives && (sum->width != film->width || sum->height != film->height || sum->ctx != film->ctx)) return api_fail(PTB_E_INVALID, "ptb_film_reduce: destination film does not match");
    Nccl* n = nccl();
    CU(cudaSetDevice(g->ctx->device));
    const size_t count = (size_t)film->width * film->height * 4;
    if (root < 0) NC(n->AllReduce(film->accum, sum->accum, count, ncclFloat32, ncclSum, g->comm, g->ctx->stream));
    else NC(n->Reduce(film->accum, receives ? sum->accum : nullptr, count, ncclFloat32, ncclSum, root, g->comm, g->ctx->stream));
    return PTB_OK;
}

int ptb_film_reduce_resolve(ptb_film* film, ptb_group* g, int root, uint32_t mode, const ptb_resolve_params* params, ptb_film* sum, uint8_t* host_srgb8) {
    if (!film || !g || !params || root < 0 || root >= g->world) return api_fail(PTB_E_INVALID, "ptb_film_reduce_resolve: bad argument");
    if (mode > PTB_REDUCE_P2P_FUSED) return api_fail(PTB_E_INVALID, "ptb_film_reduce_resolve: unknown mode");
    if (params->ToneMappingMode > PTB_TONE_MAPPING_MODE_ACES) return api_fail(PTB_E_INVALID, "ptb_film_reduce_resolve: unknown tone mapping mode");
    const bool is_root = root == g->rank;
    if (sum && is_root && (sum->width != film->width || sum->height != film->height || sum->ctx != film->ctx)) return api_fail(PTB_E_INVALID, "ptb_film_reduce_resolve: destination film does not match");
    Nccl* n = nccl();
    CU(cudaSetDevice(g->ctx->device));
    cudaStream_t st = g->ctx->stream;
    const size_t pixels = (size_t)film->width * film->height;
    const int W = g->world;
    if (mode == PTB_REDUCE_NCCL_SLICED && (pixels % (size_t)W != 0 || sum)) mode = PTB_REDUCE_NCCL;     // slices must be equal; no float sum on this path
    if (!film->srgb) CU(cudaMalloc((void**)&film->srgb, pixels * sizeof(uchar4)));
    FilmPeers* peers = nullptr; FilmPeers* sum_peers = nullptr;
    bool sum_in_place = false;
    if (mode == PTB_REDUCE_P2P_FUSED) {
        // Peer pointers are exchanged once per film and cached.  Whether an exchange is due is decided TOGETHER (one
        // 4-float all-reduce), so that no rank ever waits in a collective the others skipped.  Non-roots stand in for the
        // root's `sum` film with their own film (only the root's entry of that table is ever dereferenced).
        ptb_film* sum_key = is_root ? sum : film;
        peers = film_peers_lookup(g, film, 0);
        if (sum_key) sum_peers = film_peers_lookup(g, sum_key, 1);
        float flags[4] = {(is_root && sum && sum != film) ? 1.0f : 0.0f, peers ? 0.0f : 1.0f, sum_peers ? 0.0f : 1.0f, (is_root && sum == film) ? 1.0f : 0.0f};
        CU(cudaMemcpyAsync(g->d_token + 4, flags, sizeof flags, cudaMemcpyHostToDevice, st));
        NC(n->AllReduce(g->d_token + 4, g->d_token + 4, 4, ncclFloat32, ncclSum, g->comm, st));
        CU(cudaMemcpyAsync(flags, g->d_token + 4, sizeof flags, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        int rc;
        if (flags[1] != 0.0f && (rc = film_peers_exchange(g, film, 0, &peers))) return rc;
        if (flags[0] != 0.0f) { if (flags[2] != 0.0f && (rc = film_peers_exchange(g, sum_key ? sum_key : film, 1, &sum_peers))) return rc; }
        else sum_peers = nullptr;
        sum_in_place = flags[3] != 0.0f;
    }
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
    int rc = PTB_OK;
    uchar4* result = film->srgb;          // root: where the resolved frame ends up
    do {
        if (mode == PTB_REDUCE_P2P_FUSED) {
            if ((rc = group_barrier(g))) break;                       // every rank's render has finished (stream order + the collective)
            cudaEventRecord(e0, st);
            const u32 begin = (u32)(pixels * (size_t)g->rank / W), end = (u32)(pixels * (size_t)(g->rank + 1) / W);
            PeerFilms pf;
            for (int k = 0; k < PTB_MAX_GROUP; k++) pf.accum[k] = peers->accum[k < W ? k : 0];
            float4* out_sum = sum_peers ? sum_peers->accum[root] : sum_in_place ? peers->accum[root] : nullptr;
            const unsigned grid = (end - begin + 255u) / 256u;
            if (grid) {
                if (W <= 2) k_reduce_resolve_p2p<2><<<grid, 256, 0, st>>>(pf, W, begin, end, *params, peers->srgb[root], out_sum);
                else if (W <= 4) k_reduce_resolve_p2p<4><<<grid, 256, 0, st>>>(pf, W, begin, end, *params, peers->srgb[root], out_sum);
                else if (W <= 8) k_reduce_resolve_p2p<8><<<grid, 256, 0, st>>>(pf, W, begin, end, *params, peers->srgb[root], out_sum);
                else k_reduce_resolve_p2p<PTB_MAX_GROUP><<<grid, 256, 0, st>>>(pf, W, begin, end, *params, peers->srgb[root], out_sum);
            }
            cudaError_t e = cudaGetLastError();
            if (e != cudaSuccess) { rc = api_cuda_fail(e, "k_reduce_resolve_p2p"); break; }
            cudaEventRecord(e1, st);
            if ((rc = group_barrier(g))) break;                       // every slice has landed in the root's buffers
        } else if (mode == PTB_REDUCE_NCCL_SLICED) {
            const size_t slice = pixels / W;
            if (g->slice_pixels < slice) {
                cudaFree(g->slice_sum); cudaFree(g->slice_srgb); g->slice_sum = nullptr; g->slice_srgb = nullptr; g->slice_pixels = 0;
                if (cudaMalloc((void**)&g->slice_sum, slice * 16) != cudaSuccess || cudaMalloc((void**)&g->slice_srgb, slice * 4) != cudaSuccess) { rc = api_fail(PTB_E_NOMEM, "ptb_film_reduce_resolve: slice scratch"); break; }
                g->slice_pixels = slice;
            }
            cudaEventRecord(e0, st);
            ncclResult_t r = n->ReduceScatter(film->accum, g->slice_sum, slice * 4, ncclFloat32, ncclSum, g->comm, st);
            if (r != ncclSuccess) { rc = nccl_fail(r, "ncclReduceScatter"); break; }
            cudaError_t e = launch_resolve((const float4*)g->slice_sum, (unsigned)slice, 1u, *params, g->slice_srgb, nullptr, st);
            if (e != cudaSuccess) { rc = api_cuda_fail(e, "k_resolve(slice)"); break; }
            r = n->AllGather(g->slice_srgb, film->srgb, slice * 4, ncclChar, g->comm, st);
            if (r != ncclSuccess) { rc = nccl_fail(r, "ncclAllGather"); break; }
            cudaEventRecord(e1, st);
        } else {
            float4* dst = nullptr;
            if (is_root) {
                if (sum) dst = sum->accum;
                else {
                    if (g->slice_pixels < pixels) {
                        cudaFree(g->slice_sum); cudaFree(g->slice_srgb); g->slice_sum = nullptr; g->slice_srgb = nullptr; g->slice_pixels = 0;
                        if (cudaMalloc((void**)&g->slice_sum, pixels * 16) != cudaSuccess || cudaMalloc((void**)&g->slice_srgb, pixels * 4) != cudaSuccess) { rc = api_fail(PTB_E_NOMEM, "ptb_film_reduce_resolve: sum scratch"); }
                        else g->slice_pixels = pixels;
                    }
                    dst = (float4*)g->slice_sum;
                }
            }
            // (an allocation failure on the root must not leave the other ranks waiting in the collective: it still joins)
            cudaEventRecord(e0, st);
            ncclResult_t r = n->Reduce(film->accum, dst ? (void*)dst : (void*)film->accum, pixels * 4, ncclFloat32, ncclSum, root, g->comm, st);
            if (r != ncclSuccess) { rc = nccl_fail(r, "ncclReduce"); break; }
            if (rc) break;
            if (is_root) {
                cudaError_t e = launch_resolve(dst, film->width, film->height, *params, film->srgb, nullptr, st);
                if (e != cudaSuccess) { rc = api_cuda_fail(e, "k_resolve"); break; }
            }
            cudaEventRecord(e1, st);
        }
        if (is_root && host_srgb8) {
            cudaError_t e = cudaMemcpyAsync(host_srgb8, result, pixels * 4, cudaMemcpyDeviceToHost, st);
            if (e != cudaSuccess) { rc = api_cuda_fail(e, "cudaMemcpyAsync(frame)"); break; }
        }
        cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { rc = api_cuda_fail(e, "ptb_film_reduce_resolve"); break; }
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, e0, e1) == cudaSuccess) { g->last.exchange_ms = ms; g->last.mode = mode; g->last.bytes_in = (uint64_t)pixels * 16; }
    } while (0);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return rc;
}

int ptb_group_last_reduce(ptb_group* g, ptb_reduce_times* out) {
    if (!g || !out) return api_fail(PTB_E_INVALID, "ptb_group_last_reduce: NULL argument");
    *out = g->last;
    return PTB_OK;
}

int ptb_scene_replicate(ptb_scene* s, ptb_group* g, int root) {
    if (!s || !g || root < 0 || root >= g->world) return api_fail(PTB_E_INVALID, "ptb_scene_replicate: bad argument");
    if (s->ctx != g->ctx) return api_fail(PTB_E_INVALID, "ptb_scene_replicate: scene and group belong to different contexts");
    Nccl* n = nccl();
    CU(cudaSetDevice(g->ctx->device));
    cudaStream_t st = g->ctx->stream;
    CU(cudaStreamSynchronize(st));
    const bool is_root = g->rank == root;
    ReplicaHeader h;
    std::memset(&h, 0, sizeof h);
    if (is_root) {
        if (s->upgrade) return api_fail(PTB_E_UNSUPPORTED, "ptb_scene_replicate: the root scene has a background tree build (PTB_BUILDER_AUTO); use another builder");
        h.textures = s->textures.n; h.materials = s->materials.n; h.shapes = s->shapes.n; h.shape_nodes = s->shape_nodes.n;
        h.faces = s->faces.n; h.vertices = s->vertices.n; h.mesh_nodes = s->mesh_nodes.n; h.cameras = s->cameras.n;
        h.layers = s->layers.size(); h.layer_floats = s->layers.empty() ? 0 : s->layers[0].n;
        h.wide_nodes = s->wide_nodes.n; h.wide_prims = s->wide_prims.n; h.wide_planes = s->wide_planes.n;
        h.ds = s->ds; h.info = s->wide_info; h.material_features = s->material_features; h.builder = s->builder;
    }
    DevBuf<unsigned char> dh;
    CU(dh.alloc(sizeof h));
    if (is_root) CU(cudaMemcpyAsync(dh.p, &h, sizeof h, cudaMemcpyHostToDevice, st));
    NC(n->Broadcast(dh.p, dh.p, sizeof h, ncclChar, root, g->comm, st));
    CU(cudaMemcpyAsync(&h, dh.p, sizeof h, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    dh.release();
    int rc;
    if ((rc = bcast_buf(n, g, root, s->textures, h.textures))) return rc;
    if ((rc = bcast_buf(n, g, root, s->materials, h.materials))) return rc;
    if ((rc = bcast_buf(n, g, root, s->shapes, h.shapes))) return rc;
    if ((rc = bcast_buf(n, g, root, s->shape_nodes, h.shape_nodes))) return rc;
    if ((rc = bcast_buf(n, g, root, s->faces, h.faces))) return rc;
    if ((rc = bcast_buf(n, g, root, s->vertices, h.vertices))) return rc;
    if ((rc = bcast_buf(n, g, root, s->mesh_nodes, h.mesh_nodes))) return rc;
    if ((rc = bcast_buf(n, g, root, s->cameras, h.cameras))) return rc;
    if ((rc = bcast_buf(n, g, root, s->wide_nodes, h.wide_nodes))) return rc;
    if ((rc = bcast_buf(n, g, root, s->wide_prims, h.wide_prims))) return rc;
    if ((rc = bcast_buf(n, g, root, s->wide_planes, h.wide_planes))) return rc;
    if (!is_root) { for (auto& l : s->layers) l.release(); s->layers.assign((size_t)h.layers, DevBuf<float>()); }
    for (size_t i = 0; i < h.layers; i++) if ((rc = bcast_buf(n, g, root, s->layers[i], h.layer_floats))) return rc;
    if (!is_root) {
        std::vector<const float*> table(h.layers);
        for (size_t i = 0; i < h.layers; i++) table[i] = s->layers[i].p;
        CU(s->layer_table.upload(table.data(), table.size(), st));
        CU(cudaStreamSynchronize(st));
        DeviceScene& d = s->ds;
        d = h.ds;                                  // counts, globals, features, tree root; then this device's own pointers
        d.textures = s->textures.p; d.material_data = s->materials.p; d.shapes = s->shapes.p; d.shape_nodes = s->shape_nodes.p;
        d.mesh_faces = s->faces.p; d.mesh_vertices = s->vertices.p; d.mesh_nodes = s->mesh_nodes.p; d.cameras = s->cameras.p;
        d.atlas_layers = s->layer_table.p;
        d.wide_nodes = (const uint4*)s->wide_nodes.p; d.wide_tris = (const float4*)s->wide_prims.p; d.wide_planes = s->wide_planes.p;
        s->wide_info = h.info; s->material_features = h.material_features; s->builder = h.builder;
        if ((rc = ptb_scene_build_face_attrs(s))) return rc;      // gathered here from the arrays that just arrived (not broadcast)
    }
    CU(cudaStreamSynchronize(st));
    return PTB_OK;
}

}  // extern "C"
