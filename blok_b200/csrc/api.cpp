// C-ABI implementation (include/blokb200.h): context, scene upload, frame constants, dispatch, readback.
// Host C++20; device work lives in kernels.cu. Every CUDA call is checked; nothing here falls back to the CPU.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "../../include/blokb200.h"
#include "build.h"
#include "comm.h"
#include "kernels.h"
#include "post.h"
#include "scene_build.h"

using namespace blokb;

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CK(expr)                                                                                          \
    do {                                                                                                  \
        cudaError_t e_ = (expr);                                                                          \
        if (e_ != cudaSuccess) return fail(BLOKB_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
    } while (0)

constexpr int kSlots = 5;
constexpr uint32_t kRefillDefault = 0;  // bounce pass: lanes refilled below this many walking lanes (0 = off); measured: DESIGN.md 5
constexpr size_t kSlicePad = 64;  // >= the largest communicator size blokb_reduce_accum slices for
struct blokb_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    // world
    bool haveWorld = false;
    DeviceWorld dw{};
    std::vector<void*> worldAllocs;
    blokb_world_stats stats{};
    // sizes of the device-resident compact arrays (blokb_world_export) and, for device-built worlds, the SubChunkGpu records
    size_t nNodesCompact = 0, nBlocks = 0, nVoxels = 0;
    const blokb_subchunk* dSubs = nullptr;
    size_t nSubsDev = 0;
    // frame
    bool haveFrame = false;
    FrameDev frame{};
    blokb_buffers buf{};
    // denoiser / post-process images (allocated by the first blokb_denoise / blokb_postprocess at this size)
    blokb_frame_ubo ubo{};
    struct Post {
        bool alloc = false;
        float* histColor[2] = {nullptr, nullptr};     // gbuffer.historyColor[2]      rgba32f
        float* histMoments[2] = {nullptr, nullptr};   // gbuffer.historyMoments[2]    rg32f
        uint16_t* histLen[2] = {nullptr, nullptr};    // gbuffer.historyLength[2]     r16f
        float* prevWorldPos = nullptr;                // previousWorldPosition()  — the other half of the double-buffered
        uint16_t* prevNormalRough = nullptr;          // previousNormalRoughness()  G-buffer geometry; roles flip in end_frame
        float* variance = nullptr;                    // gbuffer.variance             r32f
        float* ping = nullptr;                        // gbuffer.filterPing / filterPong rgba32f
        float* pong = nullptr;
        float* taaHist[2] = {nullptr, nullptr};       // PostProcessBuffers::taaHistory[2] rgba32f
        float* taaOut = nullptr;                      // taaOutput (only when tonemapping is off)
        uint8_t* tonemapOut = nullptr;                // tonemapOutput rgba8
        uint8_t* sharpenOut = nullptr;                // sharpenOutput rgba8
        int i = 0, j = 0;                             // DenoiserGBuffer::historyIndex, PostProcessBuffers::historyIndex
        const float* denoised = nullptr;              // Denoiser::getOutputImage() of the current frame
        const void* final = nullptr;                  // PostProcess::getOutputImage() of the current frame
        size_t finalBytesPerPixel = 0;
    } post;
    // peer mappings of the exchanged images (blokb_set_peers): slot 0 hit, 1 shadow, 2 accum, 3 packed hit, 4 colour
    struct Peers {
        void* mapped[kMaxPeers] = {};   // cudaIpcOpenMemHandle results, nullptr at `self`
        uint32_t n = 0, self = 0;
    } peers[kSlots];
    // externally owned symmetric memory standing in for an image (blokb_bind_image); own = the context's allocation, kept aside
    struct Bound {
        void* own = nullptr;
        void* mc = nullptr;
        bool on = false;
        bool readOnlyMc = false;   // blokb_bind_accum_for_reduce: dispatches store into the local copy only; mc serves multimem.ld_reduce
    } bound[kSlots];
    // cross-GPU reduction (blokb_set_comm / blokb_comm_init / blokb_reduce_accum)
    NcclComm comm = nullptr;
    bool commOwned = false;
    int commRank = 0, commSize = 1;
    float* dSlice = nullptr;      // reduce-scatter destination: this rank's slice of the summed accumulation image
    size_t sliceCap = 0;          // pixels
    float* dFence = nullptr;      // 4-byte all-reduce operand (fence around the multimem reduction)
    // blokb_bind_fence: counter word pair in symmetric memory, ranks of the frame, fences enqueued so far
    uint32_t* fenceLocal = nullptr;
    void* fenceMc = nullptr;
    uint32_t fenceRanks = 0, fenceEpoch = 0;
    // dWave / dReorder are one scratch per context: a dispatch that uses one on another stream than the previous user first waits for that
    // user's event, so path dispatches of one context on different streams serialise on the scratch instead of racing on it
    struct ScratchGuard { cudaStream_t last = nullptr; cudaEvent_t ev = nullptr; bool used = false; } waveGuard, reorderGuard;
    // Persisting-L2 window over nodes[]: set-aside bytes granted by the device, 0 = off
    size_t l2SetAside = 0, l2MaxWindow = 0;
    uint32_t* fenceFlagHost = nullptr;   // mapped page-locked word: set by a fence kernel that timed out, read before the next fence is enqueued
    uint32_t* fenceFlagDev = nullptr;
    // editable worlds (blokb_upload_dense keeps the voxel list on the device; blokb_update_dense replaces dirty chunks in it)
    build::VoxelList edit;
    std::vector<blokb_material> editMaterials;
    bool editable = false;
    // scratch for host-ray calls
    float* dRays = nullptr;
    blokb_hit* dHits = nullptr;
    size_t rayCap = 0;
    void* dReorder = nullptr;              // scratch of blokb_trace_rays_reordered (keys, order, cub), grow-only
    size_t reorderCap = 0;
    unsigned long long* dStats = nullptr;  // 2 counters when stats enabled
    // wavefront path tracing (blokb_set_path_options): slot arrays of one batch, grow-only
    blokb_path_options pathOpt{BLOKB_PATH_WAVEFRONT, 8, 8, 6, 2, kRefillDefault};
    void* dWave = nullptr;
    size_t waveCap = 0;
    uint32_t waveBatch = 0;   // batch the current scratch was sized for (<= pathOpt.batchSamples)
    size_t waveThreads = 0;   // ... and the launch size (threads of the pixel grid)
    bool waveLimited = false; // the device could not hold the configured batch: waveBatch is all it gets at this size
    bool statsOn = false;
    uint64_t launches = 0;
    // Dispatches enqueued on caller-supplied streams: one event per foreign stream, recorded after every dispatch on it. Calls that
    // run on the context's own stream (readback, write_image, stats_read, uploads, reallocation) first make that stream wait
    // for these events, so "dispatch on my stream, then blokb_readback" is ordered without the caller synchronising.
    struct Foreign { cudaStream_t s; cudaEvent_t ev; bool dirty; };
    std::vector<Foreign> foreign;
};

static void l2_window(blokb_ctx* c, cudaStream_t q);
// After a dispatch on stream q: remember it if q is not the context's own stream.
static int dispatched(blokb_ctx* c, cudaStream_t q) {
    if (q == c->stream) return 0;
    for (auto& f : c->foreign)
        if (f.s == q) { f.dirty = true; CK(cudaEventRecord(f.ev, q)); return 0; }
    cudaEvent_t ev = nullptr;
    CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    c->foreign.push_back({q, ev, true});
    CK(cudaEventRecord(ev, q));
    l2_window(c, q);   // from this stream's next dispatch on
    return 0;
}
// Access-policy window of a stream: loads from the node pool are marked persisting in L2 (up to the set-aside the device granted),
// so the gigabytes of slot arrays a wavefront batch streams past do not evict the tree the walks descend through. Best effort.
static void l2_window(blokb_ctx* c, cudaStream_t q) {
    if (!c->l2SetAside || !c->haveWorld || !c->dw.nodes) return;
    cudaStreamAttrValue v{};
    const size_t bytes = std::min(size_t(c->dw.nNodes) * sizeof(Node), c->l2MaxWindow);
    v.accessPolicyWindow.base_ptr = const_cast<uint4*>(c->dw.nodes);
    v.accessPolicyWindow.num_bytes = bytes;
    v.accessPolicyWindow.hitRatio = bytes ? std::min(1.0f, float(double(c->l2SetAside) / double(bytes))) : 0.0f;
    v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    cudaStreamSetAttribute(q, cudaStreamAttributeAccessPolicyWindow, &v);
    cudaGetLastError();
}
static void world_windows(blokb_ctx* c) {   // a new world: the windows of every stream this context knows move to the new node pool
    l2_window(c, c->stream);
    for (auto& f : c->foreign) l2_window(c, f.s);
}
// Before work on the context's own stream that reads or replaces what dispatches wrote.
static int join_dispatches(blokb_ctx* c) {
    for (auto& f : c->foreign)
        if (f.dirty) { CK(cudaStreamWaitEvent(c->stream, f.ev, 0)); f.dirty = false; }
    return 0;
}
// Everything this context enqueued anywhere has finished (before buffers are freed or replaced).
static int idle(blokb_ctx* c) {
    if (int rc = join_dispatches(c)) return rc;
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

static void free_world(blokb_ctx* c) {
    for (void* p : c->worldAllocs) cudaFree(p);
    c->worldAllocs.clear();
    c->haveWorld = false;
    c->nNodesCompact = c->nBlocks = c->nVoxels = c->nSubsDev = 0;
    c->dSubs = nullptr;
    c->edit.release();   // a replaced world is editable again only through blokb_upload_dense
    c->editMaterials.clear();
    c->editable = false;
}
static int peer_slot(int image) {
    return image == BLOKB_IMG_HIT ? 0 : image == BLOKB_IMG_SHADOW ? 1 : image == BLOKB_IMG_ACCUM ? 2 : image == BLOKB_IMG_HIT_PACKED ? 3 :
           image == BLOKB_IMG_COLOR ? 4 : -1;
}
static void close_peers(blokb_ctx* c, int slot) {
    auto& P = c->peers[slot];
    for (uint32_t k = 0; k < P.n; k++)
        if (P.mapped[k]) cudaIpcCloseMemHandle(P.mapped[k]);
    P = blokb_ctx::Peers{};
}
// own buffer first, then the peers' mappings of the same image
static PeerOut out_for(blokb_ctx* c, int slot, void* own) {
    PeerOut o;
    o.dst[0] = own;
    o.n = 1;
    if (c->bound[slot].on && c->bound[slot].mc && !c->bound[slot].readOnlyMc) {
        o.mc = c->bound[slot].mc;
        return o;
    }
    const auto& P = c->peers[slot];
    for (uint32_t k = 0; k < P.n; k++)
        if (k != P.self) o.dst[o.n++] = P.mapped[k];
    return o;
}
static void** image_slot_ptr(blokb_ctx* c, int slot) {
    return slot == 0 ? reinterpret_cast<void**>(&c->buf.hit) : slot == 1 ? reinterpret_cast<void**>(&c->buf.shadow) :
           slot == 2 ? reinterpret_cast<void**>(&c->buf.accum) : slot == 3 ? reinterpret_cast<void**>(&c->buf.hitPacked) : reinterpret_cast<void**>(&c->buf.color);
}
static void unbind(blokb_ctx* c, int slot) {
    auto& B = c->bound[slot];
    if (B.on) *image_slot_ptr(c, slot) = B.own;
    B = blokb_ctx::Bound{};
}
static PeerOut own_only(void* own) {
    PeerOut o;
    o.dst[0] = own;
    return o;
}

static void free_images(blokb_ctx* c) {
    for (int sl = 0; sl < kSlots; sl++) { close_peers(c, sl); unbind(c, sl); }
    void* ps[] = {c->buf.hit, c->buf.shadow, c->buf.color, c->buf.worldPos, c->buf.normalRough, c->buf.albedoMetal, c->buf.motion, c->buf.accum, c->buf.hitPacked};
    for (void* p : ps) if (p) cudaFree(p);
    c->buf = blokb_buffers{};
    auto& P = c->post;
    void* qs[] = {P.histColor[0], P.histColor[1], P.histMoments[0], P.histMoments[1], P.histLen[0], P.histLen[1], P.prevWorldPos, P.prevNormalRough,
                  P.variance, P.ping, P.pong, P.taaHist[0], P.taaHist[1], P.taaOut, P.tonemapOut, P.sharpenOut};
    for (void* p : qs) if (p) cudaFree(p);
    P = blokb_ctx::Post{};
}

template <class T>
static int upload_vec(blokb_ctx* c, const std::vector<T>& v, const T** out) {
    void* d = nullptr;
    const size_t bytes = std::max<size_t>(v.size() * sizeof(T), 16);
    CK(cudaMalloc(&d, bytes));
    c->worldAllocs.push_back(d);
    if (!v.empty()) CK(cudaMemcpyAsync(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    *out = static_cast<const T*>(d);
    return 0;
}

extern "C" {

const char* blokb_last_error(void) { return g_err.c_str(); }

int blokb_create(int device, blokb_ctx** out) {
    if (!out) return fail(BLOKB_E_ARG, "out is null");
    *out = nullptr;
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail(BLOKB_E_CUDA, "no such CUDA device");
    CK(cudaSetDevice(device));
    blokb_ctx* c = new (std::nothrow) blokb_ctx;
    if (!c) return fail(BLOKB_E_ARG, "out of host memory");
    c->device = device;
    cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { delete c; return fail(BLOKB_E_CUDA, cudaGetErrorString(e)); }
    e = cudaMalloc(&c->dStats, 4 * sizeof(unsigned long long));
    if (e != cudaSuccess) { cudaStreamDestroy(c->stream); delete c; return fail(BLOKB_E_CUDA, cudaGetErrorString(e)); }
    cudaMemset(c->dStats, 0, 4 * sizeof(unsigned long long));
    // BLOKB_L2_PERSIST_MB=<n>: ask the device to set n MB of L2 aside for persisting lines (device-wide limit; clamped to what it allows)
    // and mark the node pool persisting on every stream this context dispatches on. Off unless set: measured, DESIGN.md 5.
    if (const char* env = std::getenv("BLOKB_L2_PERSIST_MB")) {
        const long mb = std::strtol(env, nullptr, 10);
        int maxPersist = 0, maxWindow = 0;
        cudaDeviceGetAttribute(&maxPersist, cudaDevAttrMaxPersistingL2CacheSize, device);
        cudaDeviceGetAttribute(&maxWindow, cudaDevAttrMaxAccessPolicyWindowSize, device);
        if (mb > 0 && maxPersist > 0 && maxWindow > 0) {
            const size_t want = std::min(size_t(mb) << 20, size_t(maxPersist));
            if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) == cudaSuccess) { c->l2SetAside = want; c->l2MaxWindow = size_t(maxWindow); }
            cudaGetLastError();
        }
    }
    *out = c;
    return 0;
}

void blokb_destroy(blokb_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    idle(c);
    for (auto& f : c->foreign) cudaEventDestroy(f.ev);
    if (c->waveGuard.ev) cudaEventDestroy(c->waveGuard.ev);
    if (c->reorderGuard.ev) cudaEventDestroy(c->reorderGuard.ev);
    free_world(c);
    free_images(c);
    if (c->dRays) cudaFree(c->dRays);
    if (c->dHits) cudaFree(c->dHits);
    if (c->dStats) cudaFree(c->dStats);
    if (c->dReorder) cudaFree(c->dReorder);
    if (c->dWave) cudaFree(c->dWave);
    if (c->dSlice) cudaFree(c->dSlice);
    if (c->dFence) cudaFree(c->dFence);
    if (c->fenceFlagHost) cudaFreeHost(c->fenceFlagHost);
    if (c->comm && c->commOwned) { std::string err; if (const NcclApi* N = nccl_api(err)) N->CommDestroy(c->comm); }
    cudaStreamDestroy(c->stream);
    delete c;
}

int blokb_upload_world(blokb_ctx* c, const blokb_svo_node* nodes, size_t nNodes, const blokb_subchunk* subs, size_t nSubs,
                       const blokb_material* materials, size_t nMaterials) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if ((nNodes && !nodes) || (nSubs && !subs) || (nMaterials && !materials)) return fail(BLOKB_E_ARG, "null array with non-zero count");
    CK(cudaSetDevice(c->device));
    const auto t0 = std::chrono::steady_clock::now();
    HostWorld hw;
    std::string err;
    int rc = build_host_world(nodes, nNodes, subs, nSubs, hw, err);
    if (rc) return fail(rc, err);
    const auto t1 = std::chrono::steady_clock::now();
    if (int rc2 = idle(c)) return rc2;  // updateWorld waits idle before replacing buffers (renderer.hpp:55-71)
    free_world(c);
    DeviceWorld& d = c->dw;
    d = DeviceWorld{};
    for (int a = 0; a < 3; a++) { d.origin[a] = hw.origin[a]; d.dim[a] = hw.dim[a]; }
    d.levels = hw.levels;
    d.matWide = hw.matWide ? 1 : 0;
    {
        const Node* p;
        if ((rc = upload_vec(c, hw.nodes, &p))) return rc;
        d.nodes = reinterpret_cast<const uint4*>(p);
    }
    if ((rc = upload_vec(c, hw.blockFirstLeaf, &d.blockFirstLeaf))) return rc;
    if (hw.matWide) { const uint32_t* p; if ((rc = upload_vec(c, hw.mat32, &p))) return rc; d.voxelMat = p; }
    else { const uint16_t* p; if ((rc = upload_vec(c, hw.mat16, &p))) return rc; d.voxelMat = p; }
    {
        // hit.rchit clamps the index to 65535 and reads whatever is there; keep reads in bounds with a default tail entry
        std::vector<blokb_material> m(materials, materials + nMaterials);
        if (m.empty()) { blokb_material def{{0.8f, 0.8f, 0.8f}, (127u << 16), {0, 0, 0}, 0.f}; m.push_back(def); }
        const blokb_material* p;
        if ((rc = upload_vec(c, m, &p))) return rc;
        d.materials = reinterpret_cast<const float4*>(p);
        d.nMaterials = (uint32_t)m.size();
    }
    CK(cudaStreamSynchronize(c->stream));
    const auto t2 = std::chrono::steady_clock::now();
    blokb_world_stats& s = c->stats;
    s = blokb_world_stats{};
    s.nodes = nNodes; s.subChunks = nSubs; s.materials = nMaterials;
    s.bricks = hw.bricks.size(); s.blocks = hw.blockFirstLeaf.size(); s.voxels = hw.voxels;
    s.compactBytes = hw.compactBytes();
    for (int a = 0; a < 3; a++) { s.gridOrigin[a] = hw.origin[a]; s.gridDim[a] = hw.dim[a]; }
    s.levels = hw.levels;
    s.buildMs = std::chrono::duration<float, std::milli>(t1 - t0).count();
    s.uploadMs = std::chrono::duration<float, std::milli>(t2 - t1).count();
    c->nNodesCompact = hw.nodes.size(); d.nNodes = uint32_t(hw.nodes.size()); c->nBlocks = hw.blockFirstLeaf.size(); c->nVoxels = hw.voxels;
    c->haveWorld = true;
    world_windows(c);
    return 0;
}

// Scene in, built on the device: ChunkManager::setVoxelMaterial per voxel + rebuildDirtyChunks +
// packChunksToGpuSvo + uploadSvoBuffers in one call, without the reference-format node arrays ever existing.
static int finish_device_build(blokb_ctx* c, build::DeviceBuild& B, const blokb_material* materials, size_t nMaterials,
                               std::chrono::steady_clock::time_point t0);

int blokb_upload_voxels(blokb_ctx* c, const int32_t* xyz, const uint32_t* materialId, size_t n, const blokb_material* materials, size_t nMaterials) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if ((n && (!xyz || !materialId)) || (nMaterials && !materials)) return fail(BLOKB_E_ARG, "null array with non-zero count");
    CK(cudaSetDevice(c->device));
    const auto t0 = std::chrono::steady_clock::now();
    if (int rc2 = idle(c)) return rc2;
    build::DeviceBuild B;
    std::string err;
    int rc = build::from_voxels(xyz, materialId, n, c->stream, B, err);
    if (rc) return fail(rc, err);
    return finish_device_build(c, B, materials, nMaterials, t0);
}

int blokb_upload_dense(blokb_ctx* c, const blokb_dense_chunk* chunks, size_t nChunks, const blokb_material* materials, size_t nMaterials) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if ((nChunks && !chunks) || (nMaterials && !materials)) return fail(BLOKB_E_ARG, "null array with non-zero count");
    CK(cudaSetDevice(c->device));
    const auto t0 = std::chrono::steady_clock::now();
    if (int rc2 = idle(c)) return rc2;
    build::DeviceBuild B;
    build::VoxelList L;
    std::string err;
    int rc = build::from_dense(chunks, nChunks, c->stream, B, err, &L);
    if (rc) return fail(rc, err);
    std::vector<blokb_material> mats(materials, materials + nMaterials);
    rc = finish_device_build(c, B, materials, nMaterials, t0);   // frees the previous world (and its voxel list)
    if (rc) { L.release(); return rc; }
    c->edit = std::move(L);
    L = build::VoxelList{};
    c->editMaterials = std::move(mats);
    c->editable = true;
    return 0;
}

int blokb_update_dense(blokb_ctx* c, const blokb_dense_chunk* chunks, size_t nChunks) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if (nChunks && !chunks) return fail(BLOKB_E_ARG, "null array with non-zero count");
    if (!c->haveWorld || !c->editable) return fail(BLOKB_E_STATE, "the current world was not uploaded with blokb_upload_dense");
    if (nChunks == 0) return 0;
    CK(cudaSetDevice(c->device));
    const auto t0 = std::chrono::steady_clock::now();
    if (int rc2 = idle(c)) return rc2;   // updateWorld waits idle before replacing buffers (renderer.hpp:55-71)
    std::string err;
    int rc = build::replace_chunks(c->edit, chunks, nChunks, c->stream, err);   // on failure the list is unchanged
    if (rc) return fail(rc, err);
    build::DeviceBuild B;
    rc = build::from_list(c->edit, c->stream, B, err);
    if (rc) return fail(rc, err);
    // finish_device_build replaces the world and with it the editable state: carry the list and the materials across
    build::VoxelList L = std::move(c->edit);
    c->edit = build::VoxelList{};
    std::vector<blokb_material> mats = std::move(c->editMaterials);
    rc = finish_device_build(c, B, mats.data(), mats.size(), t0);
    c->edit = std::move(L);
    L = build::VoxelList{};
    c->editMaterials = std::move(mats);
    c->editable = rc == 0;
    if (rc) { c->edit.release(); c->editMaterials.clear(); }
    return rc;
}

static int finish_device_build(blokb_ctx* c, build::DeviceBuild& B, const blokb_material* materials, size_t nMaterials,
                               std::chrono::steady_clock::time_point t0) {
    std::string err;
    int rc = 0;
    HostWorld hw;
    hw.subs = B.recs;
    uint64_t brickStart = 0;
    rc = build_top(B.subs.data(), B.subs.size(), B.nBricks, hw, brickStart, err);
    if (rc) { build::release(B); return fail(rc, err); }
    const auto t1 = std::chrono::steady_clock::now();
    free_world(c);
    DeviceWorld& d = c->dw;
    d = DeviceWorld{};
    for (int a = 0; a < 3; a++) { d.origin[a] = hw.origin[a]; d.dim[a] = hw.dim[a]; }
    d.levels = hw.levels;
    d.matWide = B.matWide ? 1 : 0;
    // take ownership of the device arrays first so that any failure below frees them with the world
    for (void* p : {(void*)B.bricks, (void*)B.blockFirstLeaf, B.voxelMat, (void*)B.dSubs}) if (p) c->worldAllocs.push_back(p);
    Node* dBricks = B.bricks;
    d.blockFirstLeaf = B.blockFirstLeaf; d.voxelMat = B.voxelMat; c->dSubs = B.dSubs; c->nSubsDev = B.nSubs;
    B.bricks = nullptr; B.blockFirstLeaf = nullptr; B.voxelMat = nullptr; B.dSubs = nullptr;
    if (!d.blockFirstLeaf) { const std::vector<uint32_t> e; if ((rc = upload_vec(c, e, &d.blockFirstLeaf))) return rc; }
    if (!d.voxelMat) { const std::vector<uint16_t> e; const uint16_t* p; if ((rc = upload_vec(c, e, &p))) return rc; d.voxelMat = p; }
    {
        // nodes[] = host-built top (root ... sub-chunk nodes) followed by the device-built brick nodes
        void* dn = nullptr;
        const size_t total = size_t(brickStart + B.nBricks);
        if (hw.nodes.size() != brickStart) return fail(BLOKB_E_CUDA, "internal: top-of-tree size mismatch");
        CK(cudaMalloc(&dn, std::max<size_t>(total * sizeof(Node), 16)));
        c->worldAllocs.push_back(dn);
        CK(cudaMemcpyAsync(dn, hw.nodes.data(), hw.nodes.size() * sizeof(Node), cudaMemcpyHostToDevice, c->stream));
        if (B.nBricks) CK(cudaMemcpyAsync(static_cast<Node*>(dn) + brickStart, dBricks, B.nBricks * sizeof(Node), cudaMemcpyDeviceToDevice, c->stream));
        d.nodes = reinterpret_cast<const uint4*>(dn);
        c->nNodesCompact = total; d.nNodes = uint32_t(total);
    }
    {
        std::vector<blokb_material> m(materials, materials + nMaterials);
        if (m.empty()) { blokb_material def{{0.8f, 0.8f, 0.8f}, (127u << 16), {0, 0, 0}, 0.f}; m.push_back(def); }
        const blokb_material* p;
        if ((rc = upload_vec(c, m, &p))) return rc;
        d.materials = reinterpret_cast<const float4*>(p);
        d.nMaterials = (uint32_t)m.size();
    }
    CK(cudaStreamSynchronize(c->stream));
    const auto t2 = std::chrono::steady_clock::now();
    blokb_world_stats& s = c->stats;
    s = blokb_world_stats{};
    s.nodes = B.refNodes; s.subChunks = B.nSubs; s.materials = nMaterials;
    s.bricks = B.nBricks; s.blocks = B.nBlocks; s.voxels = B.nVox;
    s.compactBytes = c->nNodesCompact * sizeof(Node) + B.nBlocks * 4 + B.nVox * (B.matWide ? 4 : 2);
    for (int a = 0; a < 3; a++) { s.gridOrigin[a] = hw.origin[a]; s.gridDim[a] = hw.dim[a]; }
    s.levels = hw.levels;
    s.buildMs = std::chrono::duration<float, std::milli>(t1 - t0).count();  // H2D of the voxel list + device build + top of tree
    s.uploadMs = std::chrono::duration<float, std::milli>(t2 - t1).count();
    c->nBlocks = B.nBlocks; c->nVoxels = B.nVox;
    c->haveWorld = true;
    world_windows(c);
    return 0;
}

int blokb_upload_vox(blokb_ctx* c, const char* path, uint32_t modelIndex, const float* offset3) {
    if (!c || !path) return fail(BLOKB_E_ARG, "null argument");
    blokb_vox* v = nullptr;
    if (int rc = blokb_vox_load(path, modelIndex, offset3, &v)) return fail(rc, blokb_vox_last_error());
    uint64_t n = 0, nm = 0;
    blokb_vox_counts(v, &n, &nm, nullptr, nullptr);
    if (n == 0) {  // loadAndImportVox: `return count > 0` (vox_loader.cpp:457) — the app reports the load as failed
        blokb_vox_free(v);
        return fail(BLOKB_E_ARG, std::string(path) + ": model " + std::to_string(modelIndex) + " imports no voxels");
    }
    const int rc = blokb_upload_voxels(c, blokb_vox_xyz(v), blokb_vox_material_ids(v), (size_t)n, blokb_vox_materials(v), (size_t)nm);
    blokb_vox_free(v);
    return rc;
}

// Copies one of the device-resident world arrays to the host (parity checks, tools). bytes must match exactly; pass
// host == NULL to query the size through *bytesOut.
int blokb_world_export(blokb_ctx* c, int which, void* host, size_t bytes, size_t* bytesOut) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if (!c->haveWorld) return fail(BLOKB_E_STATE, "no world uploaded");
    const void* src = nullptr;
    size_t sz = 0;
    switch (which) {
        case BLOKB_WORLD_NODES: src = c->dw.nodes; sz = c->nNodesCompact * sizeof(Node); break;
        case BLOKB_WORLD_BLOCKFIRSTLEAF: src = c->dw.blockFirstLeaf; sz = c->nBlocks * 4; break;
        case BLOKB_WORLD_VOXELMAT: src = c->dw.voxelMat; sz = c->nVoxels * (c->dw.matWide ? 4 : 2); break;
        case BLOKB_WORLD_SUBCHUNKS: src = c->dSubs; sz = c->nSubsDev * sizeof(blokb_subchunk); break;
        default: return fail(BLOKB_E_ARG, "unknown world array id");
    }
    if (bytesOut) *bytesOut = sz;
    if (!host) return 0;
    if (which == BLOKB_WORLD_SUBCHUNKS && !src && sz == 0 && c->stats.subChunks) return fail(BLOKB_E_STATE, "sub-chunk records are kept on the device only for worlds built by blokb_upload_voxels");
    if (bytes != sz) return fail(BLOKB_E_ARG, "host buffer size mismatch: expected " + std::to_string(sz));
    CK(cudaSetDevice(c->device));
    if (sz) CK(cudaMemcpyAsync(host, src, sz, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

// ---- scene dump: the device-resident compact world, as is, to a file and back ---------------
namespace {
struct DumpHeader {
    char magic[8];          // "BLOKBWD1"
    uint32_t version;       // 1
    uint32_t headerBytes;   // sizeof(DumpHeader)
    int32_t origin[3], dim[3], levels, matWide;
    uint64_t nNodes, nBlocks, nVoxels, nMaterials, nSubs;  // element counts of the five payload arrays, in this order
    blokb_world_stats stats;
    uint64_t payloadBytes;  // sum of the arrays' byte sizes, each padded to 16
    uint64_t fnv1a;         // FNV-1a 64 of the payload
};
inline size_t pad16(size_t n) { return (n + 15) & ~size_t(15); }
uint64_t fnv1a64(const unsigned char* p, size_t n) {
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 1099511628211ull; }
    return h;
}
}  // namespace

// A dump is trusted only as far as its checksum (accidental damage); before its arrays go to the device every index the
// traversal follows is range-checked here, so that a stale, truncated-and-rechecksummed or crafted file is refused with
// BLOKB_E_ARG instead of making traverse() read nodes[base + rank], voxelMat[..] or blockFirstLeaf[aux + rank] out of bounds.
// Layout being checked (scene.h): breadth-first levels L, L-1, ..., 2 (sub-chunks), 1 (bricks); children of a node are
// nodes[base .. base + popcount(mask)) inside the next level's range; a brick's base / aux index voxelMat / blockFirstLeaf.
static std::string validate_dump(const DumpHeader& h, const Node* nodes) {
    for (int a = 0; a < 3; a++) {
        if (h.dim[a] <= 0 || (h.dim[a] & 15) || (h.origin[a] & 15)) return "grid origin / extent is not a positive multiple of 16";
        if (int64_t(h.dim[a]) > (int64_t(1) << (2 * h.levels))) return "grid extent exceeds 4^levels";
        const int64_t lo = h.origin[a], hi = int64_t(h.origin[a]) + (int64_t(1) << (2 * h.levels));
        if (lo <= -(int64_t(1) << 23) || hi >= (int64_t(1) << 24)) return "grid coordinates outside the float-exact range";
    }
    auto block_count = [](uint64_t m) {
        int n = 0;
        for (uint32_t b = 0; b < 8; b++) n += (m & (0x0000000000330033ull << (2u * (b & 1u) + 8u * ((b >> 1) & 1u) + 32u * ((b >> 2) & 1u)))) ? 1 : 0;
        return n;
    };
    uint64_t start = 0, end = 1;   // node range of the level being checked
    for (int level = h.levels; level >= 1; level--) {
        if (end > h.nNodes) return "node levels run past the node array";
        uint64_t kids = 0;
        for (uint64_t i = start; i < end; i++) kids += uint64_t(__builtin_popcountll(nodes[i].mask));
        if (level == 1) {
            for (uint64_t i = start; i < end; i++) {
                const Node& n = nodes[i];
                if (uint64_t(n.base) + uint64_t(__builtin_popcountll(n.mask)) > h.nVoxels) return "brick " + std::to_string(i) + " indexes past voxelMat";
                if (uint64_t(n.aux) + uint64_t(block_count(n.mask)) > h.nBlocks) return "brick " + std::to_string(i) + " indexes past blockFirstLeaf";
            }
            if (end != h.nNodes) return "node array is longer than its levels";
            break;
        }
        const uint64_t cs = end, ce = end + kids;   // the next level's range
        if (ce > h.nNodes) return "level " + std::to_string(level) + " has more children than nodes follow";
        for (uint64_t i = start; i < end; i++) {
            const Node& n = nodes[i];
            const uint64_t k = uint64_t(__builtin_popcountll(n.mask));
            if (k && (n.base < cs || uint64_t(n.base) + k > ce)) return "node " + std::to_string(i) + " points outside level " + std::to_string(level - 1);
        }
        start = cs; end = ce;
    }
    return "";
}

int blokb_world_save(blokb_ctx* c, const char* path) {
    if (!c || !path) return fail(BLOKB_E_ARG, "null argument");
    if (!c->haveWorld) return fail(BLOKB_E_STATE, "no world uploaded");
    CK(cudaSetDevice(c->device));
    DumpHeader h{};
    std::memcpy(h.magic, "BLOKBWD1", 8);
    h.version = 1; h.headerBytes = sizeof(DumpHeader);
    for (int a = 0; a < 3; a++) { h.origin[a] = c->dw.origin[a]; h.dim[a] = c->dw.dim[a]; }
    h.levels = c->dw.levels; h.matWide = c->dw.matWide;
    h.nNodes = c->nNodesCompact; h.nBlocks = c->nBlocks; h.nVoxels = c->nVoxels; h.nMaterials = c->dw.nMaterials;
    h.nSubs = c->dSubs ? c->nSubsDev : 0;
    h.stats = c->stats;
    const void* src[5] = {c->dw.nodes, c->dw.blockFirstLeaf, c->dw.voxelMat, c->dw.materials, c->dSubs};
    const size_t sz[5] = {size_t(h.nNodes) * sizeof(Node), size_t(h.nBlocks) * 4, size_t(h.nVoxels) * (h.matWide ? 4 : 2),
                          size_t(h.nMaterials) * sizeof(blokb_material), size_t(h.nSubs) * sizeof(blokb_subchunk)};
    size_t total = 0;
    for (size_t b : sz) total += pad16(b);
    std::vector<unsigned char> buf(total, 0);
    size_t off = 0;
    for (int k = 0; k < 5; k++) {
        if (sz[k]) CK(cudaMemcpyAsync(buf.data() + off, src[k], sz[k], cudaMemcpyDeviceToHost, c->stream));
        off += pad16(sz[k]);
    }
    CK(cudaStreamSynchronize(c->stream));
    h.payloadBytes = total;
    h.fnv1a = fnv1a64(buf.data(), total);
    FILE* f = std::fopen(path, "wb");
    if (!f) return fail(BLOKB_E_ARG, std::string("cannot open for writing: ") + path);
    const bool ok = std::fwrite(&h, sizeof(h), 1, f) == 1 && (total == 0 || std::fwrite(buf.data(), 1, total, f) == total);
    if (std::fclose(f) != 0 || !ok) return fail(BLOKB_E_ARG, std::string("write failed: ") + path);
    return 0;
}

int blokb_world_load(blokb_ctx* c, const char* path) {
    if (!c || !path) return fail(BLOKB_E_ARG, "null argument");
    CK(cudaSetDevice(c->device));
    const auto t0 = std::chrono::steady_clock::now();
    FILE* f = std::fopen(path, "rb");
    if (!f) return fail(BLOKB_E_ARG, std::string("cannot open: ") + path);
    DumpHeader h{};
    std::vector<unsigned char> buf;
    std::string why;
    if (std::fread(&h, sizeof(h), 1, f) != 1) why = "truncated header";
    else if (std::memcmp(h.magic, "BLOKBWD1", 8) != 0) why = "not a blok-b200 scene dump";
    else if (h.version != 1 || h.headerBytes != sizeof(DumpHeader)) why = "unsupported dump version";
    else if (h.levels < 3 || h.levels > kMaxLevels || h.nNodes == 0 || h.nNodes > (1ull << 32) || h.nBlocks > (1ull << 32) || h.nVoxels > (1ull << 32) ||
             h.nMaterials == 0 || h.nMaterials > (1ull << 24) || h.nSubs > (1ull << 32) || (h.matWide != 0 && h.matWide != 1)) why = "implausible header";
    size_t sz[5] = {};
    if (why.empty()) {
        sz[0] = size_t(h.nNodes) * sizeof(Node); sz[1] = size_t(h.nBlocks) * 4; sz[2] = size_t(h.nVoxels) * (h.matWide ? 4 : 2);
        sz[3] = size_t(h.nMaterials) * sizeof(blokb_material); sz[4] = size_t(h.nSubs) * sizeof(blokb_subchunk);
        size_t total = 0;
        for (size_t b : sz) total += pad16(b);
        if (total != h.payloadBytes) why = "payload size does not match the header";
        else {
            buf.resize(total);
            if (total && std::fread(buf.data(), 1, total, f) != total) why = "truncated payload";
            else if (std::fgetc(f) != EOF) why = "trailing bytes after the payload";
            else if (fnv1a64(buf.data(), total) != h.fnv1a) why = "checksum mismatch";
            else why = validate_dump(h, reinterpret_cast<const Node*>(buf.data()));
        }
    }
    std::fclose(f);
    if (!why.empty()) return fail(BLOKB_E_ARG, std::string(path) + ": " + why);
    const auto t1 = std::chrono::steady_clock::now();
    if (int rc2 = idle(c)) return rc2;
    free_world(c);
    void* dst[5] = {};
    size_t off = 0;
    for (int k = 0; k < 5; k++) {
        if (k == 4 && sz[k] == 0) break;
        CK(cudaMalloc(&dst[k], std::max<size_t>(sz[k], 16)));
        c->worldAllocs.push_back(dst[k]);
        if (sz[k]) CK(cudaMemcpyAsync(dst[k], buf.data() + off, sz[k], cudaMemcpyHostToDevice, c->stream));
        off += pad16(sz[k]);
    }
    CK(cudaStreamSynchronize(c->stream));
    DeviceWorld& d = c->dw;
    d = DeviceWorld{};
    for (int a = 0; a < 3; a++) { d.origin[a] = h.origin[a]; d.dim[a] = h.dim[a]; }
    d.levels = h.levels; d.matWide = h.matWide;
    d.nodes = static_cast<const uint4*>(dst[0]); d.blockFirstLeaf = static_cast<const uint32_t*>(dst[1]); d.voxelMat = dst[2];
    d.materials = static_cast<const float4*>(dst[3]); d.nMaterials = uint32_t(h.nMaterials);
    c->nNodesCompact = h.nNodes; d.nNodes = uint32_t(h.nNodes); c->nBlocks = h.nBlocks; c->nVoxels = h.nVoxels;
    c->dSubs = static_cast<const blokb_subchunk*>(dst[4]); c->nSubsDev = h.nSubs;
    c->stats = h.stats;
    const auto t2 = std::chrono::steady_clock::now();
    c->stats.buildMs = std::chrono::duration<float, std::milli>(t1 - t0).count();   // file read + checksum
    c->stats.uploadMs = std::chrono::duration<float, std::milli>(t2 - t1).count();
    c->haveWorld = true;
    world_windows(c);
    return 0;
}

int blokb_world_stats_get(blokb_ctx* c, blokb_world_stats* out) {
    if (!c || !out) return fail(BLOKB_E_ARG, "null argument");
    if (!c->haveWorld) return fail(BLOKB_E_STATE, "no world uploaded");
    *out = c->stats;
    return 0;
}

int blokb_set_frame(blokb_ctx* c, const blokb_frame_ubo* f) {
    if (!c || !f) return fail(BLOKB_E_ARG, "null argument");
    if (f->screenWidth == 0 || f->screenHeight == 0 || f->screenWidth > 65536 || f->screenHeight > 65536) return fail(BLOKB_E_ARG, "bad screen size");
    CK(cudaSetDevice(c->device));
    if (c->buf.width != f->screenWidth || c->buf.height != f->screenHeight) {
        if (int rc = idle(c)) return rc;
        // From here until every image exists the context has NO frame: a failed allocation must not leave the previous frame's
        // size paired with null or partial images (the next dispatch would write through them).
        c->haveFrame = false;
        free_images(c);
        const size_t n = size_t(f->screenWidth) * f->screenHeight;
        void** slots[] = {(void**)&c->buf.hit, (void**)&c->buf.shadow, (void**)&c->buf.color, (void**)&c->buf.worldPos, (void**)&c->buf.normalRough,
                          (void**)&c->buf.albedoMetal, (void**)&c->buf.motion, (void**)&c->buf.accum, (void**)&c->buf.hitPacked};
        // colour and accum carry kSlicePad pixels of slack: blokb_reduce_accum slices them into G equal parts of ceil(n / G) pixels
        const size_t bytes[] = {n * sizeof(blokb_hit), n, (n + kSlicePad) * 16, n * 16, n * 8, n * 4, n * 4, (n + kSlicePad) * 16, n * sizeof(blokb_hit_packed)};
        for (int k = 0; k < 9; k++) {
            const cudaError_t e = cudaMalloc(slots[k], bytes[k]);
            if (e != cudaSuccess) {
                *slots[k] = nullptr;
                free_images(c);
                return fail(BLOKB_E_CUDA, std::string("cudaMalloc of frame image ") + std::to_string(k) + ": " + cudaGetErrorString(e));
            }
        }
        // zero-filled before any dispatch can touch it, whatever stream that dispatch uses
        if (cudaMemsetAsync(c->buf.accum, 0, (n + kSlicePad) * 16, c->stream) != cudaSuccess || cudaStreamSynchronize(c->stream) != cudaSuccess) {
            free_images(c);
            return fail(BLOKB_E_CUDA, "zero-fill of the accumulation image failed");
        }
        c->buf.width = f->screenWidth; c->buf.height = f->screenHeight;
    }
    FrameDev& d = c->frame;
    std::memcpy(d.invView, f->invView, 64);
    std::memcpy(d.invProj, f->invProj, 64);
    std::memcpy(d.prevViewProj, f->prevViewProj, 64);
    std::memcpy(d.camPos, f->camPos, 12);
    d.frameCount = f->frameCount; d.sampleCount = f->sampleCount; d.width = f->screenWidth; d.height = f->screenHeight;
    c->ubo = *f;
    c->haveFrame = true;
    return 0;
}

static int ready(blokb_ctx* c, bool needFrame) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if (!c->haveWorld) return fail(BLOKB_E_STATE, "no world uploaded (blokb_upload_world)");
    if (needFrame && !c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame)");
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return fail(BLOKB_E_CUDA, cudaGetErrorString(e));
    return 0;
}
static cudaStream_t pick(blokb_ctx* c, void* s) { return s ? static_cast<cudaStream_t>(s) : c->stream; }
static unsigned long long* st(blokb_ctx* c) { return c->statsOn ? c->dStats : nullptr; }

int blokb_set_strips(blokb_ctx* c, uint32_t first, uint32_t stride) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if (stride == 0) return fail(BLOKB_E_ARG, "strip stride is 0");
    c->frame.stripFirst = first; c->frame.stripStride = stride;
    return 0;
}
int blokb_ipc_export(blokb_ctx* c, int image, void* handle64) {
    if (!c || !handle64) return fail(BLOKB_E_ARG, "null argument");
    if (!c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame allocates the images)");
    const int slot = peer_slot(image);
    void* p = slot < 0 ? nullptr : *image_slot_ptr(c, slot);
    if (!p) return fail(BLOKB_E_ARG, "image is not exchangeable (hit, shadow, accum, packed hit, colour)");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    CK(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, p));
    std::memcpy(handle64, &h, 64);
    return 0;
}
int blokb_set_peers(blokb_ctx* c, int image, const void* handles, uint32_t n, uint32_t self) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    const int slot = peer_slot(image);
    if (slot < 0) return fail(BLOKB_E_ARG, "image is not exchangeable (hit, shadow, accum, packed hit, colour)");
    CK(cudaSetDevice(c->device));
    if (int rc = idle(c)) return rc;
    close_peers(c, slot);
    if (n == 0) return 0;
    if (!handles) return fail(BLOKB_E_ARG, "handles is null");
    if (n > kMaxPeers) return fail(BLOKB_E_ARG, "more than 8 peers");
    if (self >= n) return fail(BLOKB_E_ARG, "self >= n");
    auto& P = c->peers[slot];
    P.n = n; P.self = self;
    for (uint32_t k = 0; k < n; k++) {
        if (k == self) continue;
        cudaIpcMemHandle_t h;
        std::memcpy(&h, static_cast<const char*>(handles) + 64 * size_t(k), 64);
        cudaError_t e = cudaIpcOpenMemHandle(&P.mapped[k], h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            P.mapped[k] = nullptr;
            close_peers(c, slot);
            return fail(BLOKB_E_CUDA, std::string("cudaIpcOpenMemHandle(peer ") + std::to_string(k) + "): " + cudaGetErrorString(e));
        }
    }
    return 0;
}
int blokb_bind_image(blokb_ctx* c, int image, void* local, void* multicast) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    const int slot = peer_slot(image);
    if (slot < 0) return fail(BLOKB_E_ARG, "image is not exchangeable (hit, shadow, accum, packed hit, colour)");
    if (!c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame allocates the images)");
    if (!local && multicast) return fail(BLOKB_E_ARG, "a multicast address needs the local address of the same memory");
    if (multicast && image == BLOKB_IMG_SHADOW && (c->buf.width & 3u)) return fail(BLOKB_E_ARG, "multicast shadow image needs width % 4 == 0");
    if ((reinterpret_cast<uintptr_t>(local) | reinterpret_cast<uintptr_t>(multicast)) & 15u) return fail(BLOKB_E_ARG, "addresses must be 16-byte aligned");
    CK(cudaSetDevice(c->device));
    if (int rc = idle(c)) return rc;
    close_peers(c, slot);
    unbind(c, slot);
    if (!local) return 0;
    auto& B = c->bound[slot];
    B.own = *image_slot_ptr(c, slot);
    B.mc = multicast;
    B.on = true;
    *image_slot_ptr(c, slot) = local;
    return 0;
}
int blokb_bind_fence(blokb_ctx* c, void* local, void* multicast, uint32_t nranks) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    CK(cudaSetDevice(c->device));
    if (int rc = idle(c)) return rc;
    c->fenceLocal = nullptr; c->fenceMc = nullptr; c->fenceRanks = 0; c->fenceEpoch = 0;   // a new binding starts a new count (the words arrive zero-filled)
    if (c->fenceFlagHost) *reinterpret_cast<volatile uint32_t*>(c->fenceFlagHost) = 0;
    if (nranks == 0) return 0;
    if (!local || !multicast) return fail(BLOKB_E_ARG, "the fence needs the local and the multicast address of the same symmetric word pair");
    if ((reinterpret_cast<uintptr_t>(local) | reinterpret_cast<uintptr_t>(multicast)) & 15u) return fail(BLOKB_E_ARG, "addresses must be 16-byte aligned");
    if (!c->fenceFlagHost) {
        CK(cudaHostAlloc((void**)&c->fenceFlagHost, 64, cudaHostAllocMapped));
        *c->fenceFlagHost = 0;
        CK(cudaHostGetDevicePointer((void**)&c->fenceFlagDev, c->fenceFlagHost, 0));
    }
    c->fenceLocal = static_cast<uint32_t*>(local); c->fenceMc = multicast; c->fenceRanks = nranks;
    return 0;
}
int blokb_fence_status(blokb_ctx* c) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if (!c->fenceRanks) return fail(BLOKB_E_STATE, "no fence bound (blokb_bind_fence)");
    CK(cudaSetDevice(c->device));
    if (int rc = join_dispatches(c)) return rc;
    uint32_t w[2] = {0, 0};
    CK(cudaMemcpyAsync(w, c->fenceLocal, sizeof w, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    if (w[1]) return fail(BLOKB_E_STATE, "a frame fence timed out: not every rank arrived (counter " + std::to_string(w[0]) + ")");
    return 0;
}
int blokb_frame_fence(blokb_ctx* c, void* stream) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if (!c->fenceRanks) return fail(BLOKB_E_STATE, "no fence bound (blokb_bind_fence)");
    CK(cudaSetDevice(c->device));
    // sticky: a fence that timed out (its kernel has finished by now) has left the counters short; further fences cannot be trusted
    if (*reinterpret_cast<volatile uint32_t*>(c->fenceFlagHost))
        return fail(BLOKB_E_STATE, "an earlier frame fence timed out (a rank was missing); rebind the fence on every rank (blokb_bind_fence) to resynchronise");
    c->fenceEpoch++;
    CK(launch_fence(c->fenceLocal, c->fenceMc, c->fenceRanks * c->fenceEpoch, c->fenceFlagDev, pick(c, stream)));
    c->launches++;
    return dispatched(c, pick(c, stream));
}
int blokb_bind_accum_for_reduce(blokb_ctx* c, void* local, void* multicast) {
    if (int rc = blokb_bind_image(c, BLOKB_IMG_ACCUM, local, multicast)) return rc;
    if (local) c->bound[2].readOnlyMc = true;
    return 0;
}
int blokb_trace_primary(blokb_ctx* c, void* stream) {
    if (int rc = ready(c, true)) return rc;
    CK(launch_primary(c->dw, c->frame, false, out_for(c, 0, c->buf.hit), own_only(c->buf.shadow), st(c), pick(c, stream)));
    c->launches++;
    return dispatched(c, pick(c, stream));
}
int blokb_trace_primary_shadow(blokb_ctx* c, void* stream) {
    if (int rc = ready(c, true)) return rc;
    CK(launch_primary(c->dw, c->frame, true, out_for(c, 0, c->buf.hit), out_for(c, 1, c->buf.shadow), st(c), pick(c, stream)));
    c->launches++;
    return dispatched(c, pick(c, stream));
}
int blokb_trace_primary_packed(blokb_ctx* c, void* stream) {
    if (int rc = ready(c, true)) return rc;
    if (c->stats.nodes > BLOKB_PACKED_MAX_NODES) return fail(BLOKB_E_UNSUPPORTED, "world has more nodes than a packed record can address");
    CK(launch_primary_packed(c->dw, c->frame, out_for(c, 3, c->buf.hitPacked), pick(c, stream)));
    c->launches++;
    return dispatched(c, pick(c, stream));
}
static int scratch_enter(blokb_ctx::ScratchGuard& g, cudaStream_t q) {
    if (g.used && g.last != q) CK(cudaStreamWaitEvent(q, g.ev, 0));
    return 0;
}
static int scratch_leave(blokb_ctx::ScratchGuard& g, cudaStream_t q) {
    if (!g.ev) CK(cudaEventCreateWithFlags(&g.ev, cudaEventDisableTiming));
    CK(cudaEventRecord(g.ev, q));
    g.last = q; g.used = true;
    return 0;
}
// Scratch of the wavefront schedule for the current launch size (frame size / strip set): grow-only, sized for the configured
// batch; when the device cannot hold that, the batch is halved until it fits and the smaller batch is remembered.
static int wave_ready(blokb_ctx* c, uint32_t samples, uint32_t* batchOut) {
    const size_t nT = std::max<size_t>(wave_threads(c->frame), 1);
    const uint32_t maxBatch = uint32_t(std::min<size_t>(0xFFFFFFFFull / nT, 1u << 16));
    if (maxBatch == 0) return fail(BLOKB_E_UNSUPPORTED, "frame too large for the wavefront schedule (use BLOKB_PATH_FUSED)");
    const uint32_t cfg = std::max(1u, std::min(c->pathOpt.batchSamples, maxBatch));
    uint32_t want = std::max(1u, std::min(cfg, samples));
    if (c->dWave && c->waveLimited && nT >= c->waveThreads) want = std::min(want, std::max(1u, uint32_t(size_t(c->waveBatch) * c->waveThreads / nT)));
    if (c->dWave && wave_scratch_bytes(nT, want) <= c->waveCap) { *batchOut = want; return 0; }
    CK(cudaDeviceSynchronize());   // the old scratch may still be in use on any stream
    if (c->dWave) cudaFree(c->dWave);
    c->dWave = nullptr; c->waveCap = 0; c->waveBatch = 0; c->waveThreads = 0; c->waveLimited = false;
    for (uint32_t b = cfg;; b = (b + 1) / 2) {
        const size_t need = wave_scratch_bytes(nT, b);
        if (cudaMalloc(&c->dWave, need) == cudaSuccess) { c->waveCap = need; c->waveBatch = b; c->waveThreads = nT; c->waveLimited = b < cfg; break; }
        cudaGetLastError();
        c->dWave = nullptr;
        if (b == 1) return fail(BLOKB_E_CUDA, "out of device memory for the wavefront scratch (" + std::to_string(need) + " bytes at batch 1)");
    }
    *batchOut = std::min(want, c->waveBatch);
    return 0;
}
int blokb_render(blokb_ctx* c, void* stream) {
    if (int rc = ready(c, true)) return rc;
    if (c->frame.sampleCount == 0) return fail(BLOKB_E_ARG, "frame.sampleCount is 0");
    if (c->pathOpt.mode == BLOKB_PATH_FUSED) {
        CK(launch_render(c->dw, c->frame, c->buf.color, c->buf.worldPos, c->buf.normalRough, c->buf.albedoMetal, c->buf.motion, st(c), pick(c, stream)));
        c->launches += 2;  // k_gbuffer + k_path<.., kColor>
        return dispatched(c, pick(c, stream));
    }
    uint32_t batch = 0;
    if (int rc = wave_ready(c, c->frame.sampleCount, &batch)) return rc;
    CK(launch_gbuffer(c->dw, c->frame, c->buf.worldPos, c->buf.normalRough, c->buf.albedoMetal, c->buf.motion, pick(c, stream)));
    c->launches++;
    const WaveKey K = make_wave_key(c->dw, c->pathOpt.dirCells, c->pathOpt.originCellShift, c->pathOpt.mortonBitsPerAxis);
    if (int rc = scratch_enter(c->waveGuard, pick(c, stream))) return rc;
    CK(launch_path_wavefront(c->dw, c->frame, 0u, c->frame.sampleCount, true, own_only(c->buf.color), c->dWave, batch, K, c->pathOpt.refillBelow, st(c), &c->launches, pick(c, stream)));
    if (int rc = scratch_leave(c->waveGuard, pick(c, stream))) return rc;
    return dispatched(c, pick(c, stream));
}
int blokb_trace_path(blokb_ctx* c, uint32_t sampleBegin, uint32_t sampleEnd, void* stream) {
    if (int rc = ready(c, true)) return rc;
    if (sampleEnd < sampleBegin) return fail(BLOKB_E_ARG, "sampleEnd < sampleBegin");
    if (sampleEnd == sampleBegin) return 0;
    if (c->pathOpt.mode == BLOKB_PATH_FUSED) {
        CK(launch_path(c->dw, c->frame, sampleBegin, sampleEnd, out_for(c, 2, c->buf.accum), st(c), pick(c, stream)));
        c->launches++;
        return dispatched(c, pick(c, stream));
    }
    uint32_t batch = 0;
    if (int rc = wave_ready(c, sampleEnd - sampleBegin, &batch)) return rc;
    const WaveKey K = make_wave_key(c->dw, c->pathOpt.dirCells, c->pathOpt.originCellShift, c->pathOpt.mortonBitsPerAxis);
    if (int rc = scratch_enter(c->waveGuard, pick(c, stream))) return rc;
    CK(launch_path_wavefront(c->dw, c->frame, sampleBegin, sampleEnd, false, out_for(c, 2, c->buf.accum), c->dWave, batch, K, c->pathOpt.refillBelow, st(c), &c->launches,
                             pick(c, stream)));
    if (int rc = scratch_leave(c->waveGuard, pick(c, stream))) return rc;
    return dispatched(c, pick(c, stream));
}
void blokb_path_options_default(blokb_path_options* o) {
    if (o) *o = blokb_path_options{BLOKB_PATH_WAVEFRONT, 8, 8, 6, 2, kRefillDefault};
}
int blokb_set_path_options(blokb_ctx* c, const blokb_path_options* o) {
    if (!c || !o) return fail(BLOKB_E_ARG, "null argument");
    if (o->mode != BLOKB_PATH_WAVEFRONT && o->mode != BLOKB_PATH_FUSED) return fail(BLOKB_E_ARG, "unknown path mode");
    if (o->dirCells > 16 || o->mortonBitsPerAxis > 6 || o->originCellShift < 0 || o->originCellShift > 20) return fail(BLOKB_E_ARG, "sort-key shape out of range");
    blokb_path_options d;
    blokb_path_options_default(&d);
    c->pathOpt.mode = o->mode;
    c->pathOpt.batchSamples = o->batchSamples ? o->batchSamples : d.batchSamples;
    c->pathOpt.dirCells = o->dirCells ? o->dirCells : d.dirCells;
    c->pathOpt.originCellShift = o->originCellShift ? o->originCellShift : d.originCellShift;
    c->pathOpt.mortonBitsPerAxis = o->mortonBitsPerAxis ? o->mortonBitsPerAxis : d.mortonBitsPerAxis;
    if (o->refillBelow != BLOKB_REFILL_KEEP) {
        if (o->refillBelow > 32) return fail(BLOKB_E_ARG, "refillBelow must be 0..32");
        c->pathOpt.refillBelow = o->refillBelow;
    }
    return 0;
}
int blokb_get_path_options(blokb_ctx* c, blokb_path_options* o) {
    if (!c || !o) return fail(BLOKB_E_ARG, "null argument");
    *o = c->pathOpt;
    return 0;
}
int blokb_resolve(blokb_ctx* c, void* stream) {
    if (int rc = ready(c, true)) return rc;
    CK(launch_resolve(c->buf.accum, c->buf.color, size_t(c->buf.width) * c->buf.height, pick(c, stream)));
    c->launches++;
    return dispatched(c, pick(c, stream));
}
int blokb_reset_accum(blokb_ctx* c, void* stream) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if (!c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame)");
    CK(cudaSetDevice(c->device));
    CK(cudaMemsetAsync(c->buf.accum, 0, size_t(c->buf.width) * c->buf.height * 16, pick(c, stream)));
    return dispatched(c, pick(c, stream));
}
int blokb_trace_rays(blokb_ctx* c, const float* dRays, size_t n, int /*anyHit*/, blokb_hit* dOut, void* stream) {
    // The walk is strictly front to back: the first filled voxel answers both closest-hit and any-hit queries.
    if (int rc = ready(c, false)) return rc;
    if (n && (!dRays || !dOut)) return fail(BLOKB_E_ARG, "null ray / output pointer");
    CK(launch_trace_rays(c->dw, dRays, n, dOut, st(c), pick(c, stream)));
    if (n) c->launches++;
    return dispatched(c, pick(c, stream));
}
int blokb_trace_rays_reordered(blokb_ctx* c, const float* dRays, size_t n, int /*anyHit*/, blokb_hit* dOut, void* stream) {
    if (int rc = ready(c, false)) return rc;
    if (n && (!dRays || !dOut)) return fail(BLOKB_E_ARG, "null ray / output pointer");
    if (n == 0) return 0;
    if (n > 0xFFFFFFFFull) return fail(BLOKB_E_ARG, "more than 2^32 rays in one list");
    const size_t need = reorder_scratch_bytes(n, nullptr);
    if (need > c->reorderCap) {
        CK(cudaDeviceSynchronize());   // the old scratch may still be in use on any stream
        if (c->dReorder) cudaFree(c->dReorder);
        c->dReorder = nullptr; c->reorderCap = 0;
        CK(cudaMalloc(&c->dReorder, need));
        c->reorderCap = need;
    }
    if (int rc = scratch_enter(c->reorderGuard, pick(c, stream))) return rc;
    CK(launch_trace_rays_reordered(c->dw, dRays, n, dOut, c->dReorder, st(c), pick(c, stream)));
    if (int rc = scratch_leave(c->reorderGuard, pick(c, stream))) return rc;
    c->launches += 3;  // keys, radix sort (counted as one), ordered trace
    return dispatched(c, pick(c, stream));
}
// blokb_trace_rays_reordered for hosts that cannot allocate (include/blokb200.h): scratch lives in a file-scope __device__ array of
// kernels.cu, nothing is allocated or freed here — not even the per-stream event `dispatched` would create for a caller stream.
int blokb_trace_rays_reordered_static(blokb_ctx* c, const float* dRays, size_t n, int /*anyHit*/, blokb_hit* dOut, void* stream) {
    if (int rc = ready(c, false)) return rc;
    if (n && (!dRays || !dOut)) return fail(BLOKB_E_ARG, "null ray / output pointer");
    if (n == 0) return 0;
    if (n > size_t(BLOKB_STATIC_MAX_RAYS))
        return fail(BLOKB_E_ARG, "ray list of " + std::to_string(n) + " exceeds BLOKB_STATIC_MAX_RAYS (" + std::to_string(BLOKB_STATIC_MAX_RAYS) +
                                     "): the static scratch holds no more; split the list or use blokb_trace_rays_reordered");
    const cudaError_t e = launch_trace_rays_reordered_static(c->dw, dRays, n, dOut, c->device, st(c), pick(c, stream));
    if (e != cudaSuccess) return fail(BLOKB_E_CUDA, std::string("static-scratch reordered trace: ") + cudaGetErrorString(e));
    c->launches += 4;  // keys + histogram, scan, scatter, ordered trace
    return 0;
}
int blokb_trace_rays_host(blokb_ctx* c, const float* rays, size_t n, int anyHit, blokb_hit* out) {
    if (int rc = ready(c, false)) return rc;
    if (n == 0) return 0;
    if (!rays || !out) return fail(BLOKB_E_ARG, "null ray / output pointer");
    if (n > c->rayCap) {
        CK(cudaStreamSynchronize(c->stream));
        if (c->dRays) cudaFree(c->dRays);
        if (c->dHits) cudaFree(c->dHits);
        c->dRays = nullptr; c->dHits = nullptr; c->rayCap = 0;
        CK(cudaMalloc((void**)&c->dRays, n * 32));
        CK(cudaMalloc((void**)&c->dHits, n * sizeof(blokb_hit)));
        c->rayCap = n;
    }
    CK(cudaMemcpyAsync(c->dRays, rays, n * 32, cudaMemcpyHostToDevice, c->stream));
    if (int rc = blokb_trace_rays(c, c->dRays, n, anyHit, c->dHits, c->stream)) return rc;
    CK(cudaMemcpyAsync(out, c->dHits, n * sizeof(blokb_hit), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

// ---- denoiser + post-process ------------------------------------------------------
void blokb_post_settings_default(blokb_post_settings* s) {
    if (!s) return;
    // Denoiser::Settings (renderer_denoising.hpp:57-76), PostProcess::Settings (renderer_postprocess.hpp:102-119)
    *s = blokb_post_settings{4, 1, 0.93f, 0.98f, 1, 1.0f, 1.15f, 1, 0.0f, 1, 0.5f};
}
static int post_zero(blokb_ctx* c, cudaStream_t s) {
    auto& P = c->post;
    const size_t n = size_t(c->buf.width) * c->buf.height;
    for (int k = 0; k < 2; k++) {
        CK(cudaMemsetAsync(P.histColor[k], 0, n * 16, s));
        CK(cudaMemsetAsync(P.histMoments[k], 0, n * 8, s));
        CK(cudaMemsetAsync(P.histLen[k], 0, n * 2, s));
        CK(cudaMemsetAsync(P.taaHist[k], 0, n * 16, s));
    }
    CK(cudaMemsetAsync(P.prevWorldPos, 0, n * 16, s));
    CK(cudaMemsetAsync(P.prevNormalRough, 0, n * 8, s));
    P.i = P.j = 0;
    P.denoised = nullptr; P.final = nullptr;
    return 0;
}
static int post_ready(blokb_ctx* c, const blokb_post_settings* s, cudaStream_t stream) {
    if (!c || !s) return fail(BLOKB_E_ARG, "null argument");
    if (!c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame)");
    CK(cudaSetDevice(c->device));
    auto& P = c->post;
    if (P.alloc) return 0;
    const size_t n = size_t(c->buf.width) * c->buf.height;
    for (int k = 0; k < 2; k++) {
        CK(cudaMalloc((void**)&P.histColor[k], n * 16));
        CK(cudaMalloc((void**)&P.histMoments[k], n * 8));
        CK(cudaMalloc((void**)&P.histLen[k], n * 2));
        CK(cudaMalloc((void**)&P.taaHist[k], n * 16));
    }
    CK(cudaMalloc((void**)&P.prevWorldPos, n * 16));
    CK(cudaMalloc((void**)&P.prevNormalRough, n * 8));
    CK(cudaMalloc((void**)&P.variance, n * 4));
    CK(cudaMalloc((void**)&P.ping, n * 16));
    CK(cudaMalloc((void**)&P.pong, n * 16));
    CK(cudaMalloc((void**)&P.tonemapOut, n * 4));
    CK(cudaMalloc((void**)&P.sharpenOut, n * 4));
    P.alloc = true;
    return post_zero(c, stream);
}
int blokb_denoise(blokb_ctx* c, const blokb_post_settings* s, void* stream) {
    cudaStream_t q = c ? pick(c, stream) : nullptr;
    if (int rc = post_ready(c, s, q)) return rc;
    if (s->atrousIterations < 1 || s->atrousIterations > 5) return fail(BLOKB_E_ARG, "atrousIterations must be 1..5 (MAX_ATROUS_ITERATIONS)");
    auto& P = c->post;
    const post::PostFrame F = post::make_post_frame(c->ubo, *s);
    const int i = P.i, p = 1 - P.i;
    // image-to-binding wiring of Denoiser::updateDescriptorSets (renderer_denoising.cpp:432-556)
    post::TemporalIO io{c->buf.color, c->buf.worldPos, c->buf.normalRough, c->buf.motion, P.histColor[p], P.histMoments[p], P.histLen[p],
                        P.prevWorldPos, P.prevNormalRough, P.histColor[i], P.histMoments[i], P.histLen[i], P.variance};
    CK(post::launch_temporal_variance(F, io, q));
    c->launches++;
    const float* src = P.histColor[i];
    for (int it = 0; it < s->atrousIterations; it++) {
        float* dst = (it % 2 == 0) ? P.ping : P.pong;
        CK(post::launch_atrous(F, 1 << it, src, P.variance, c->buf.worldPos, c->buf.normalRough, dst, q));
        c->launches++;
        src = dst;
    }
    P.denoised = src;
    return dispatched(c, q);
}
int blokb_postprocess(blokb_ctx* c, const blokb_post_settings* s, void* stream) {
    cudaStream_t q = c ? pick(c, stream) : nullptr;
    if (int rc = post_ready(c, s, q)) return rc;
    auto& P = c->post;
    if (!P.denoised) return fail(BLOKB_E_STATE, "blokb_postprocess needs this frame's blokb_denoise output");
    if (!s->enableTAA && !s->enableTonemapping) return fail(BLOKB_E_UNSUPPORTED, "neither TAA nor tonemapping enabled (the reference has no passthrough either)");
    const post::PostFrame F = post::make_post_frame(c->ubo, *s);
    const size_t n = size_t(c->buf.width) * c->buf.height;
    if (!s->enableTonemapping && !P.taaOut) CK(cudaMalloc((void**)&P.taaOut, n * 16));
    const int j = P.j, pj = 1 - P.j;
    CK(post::launch_taa_tonemap(F, s->enableTAA != 0, s->enableTonemapping != 0, P.denoised, P.taaHist[pj], c->buf.motion, P.taaHist[j], P.taaOut,
                                reinterpret_cast<uint32_t*>(P.tonemapOut), q));
    c->launches++;
    if (s->enableTonemapping) {
        P.final = P.tonemapOut; P.finalBytesPerPixel = 4;
        if (s->enableSharpening) {
            CK(post::launch_sharpen(F, P.tonemapOut, reinterpret_cast<uint32_t*>(P.sharpenOut), q));
            c->launches++;
            P.final = P.sharpenOut;
        }
    } else {
        P.final = P.taaOut; P.finalBytesPerPixel = 16;
    }
    return dispatched(c, q);
}
int blokb_end_frame(blokb_ctx* c, void* /*stream*/) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if (!c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame)");
    auto& P = c->post;
    if (!P.alloc) return 0;
    // copyCurrentGeometryToHistory without the copy: this frame's geometry images become "previous", the old
    // previous ones receive the next frame. Everything is stream-ordered, so no synchronisation is needed.
    std::swap(c->buf.worldPos, P.prevWorldPos);
    std::swap(c->buf.normalRough, P.prevNormalRough);
    P.i = 1 - P.i;
    P.j = 1 - P.j;
    P.denoised = nullptr; P.final = nullptr;
    return 0;
}
int blokb_reset_history(blokb_ctx* c, void* stream) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if (!c->post.alloc) return 0;
    CK(cudaSetDevice(c->device));
    if (int rc = post_zero(c, pick(c, stream))) return rc;
    return dispatched(c, pick(c, stream));
}

int blokb_get_buffers(blokb_ctx* c, blokb_buffers* out) {
    if (!c || !out) return fail(BLOKB_E_ARG, "null argument");
    if (!c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame)");
    *out = c->buf;
    return 0;
}
// Device address and whole-image byte size of an image id (nullptr: not produced yet). Returns false for an unknown id.
static bool image_lookup(blokb_ctx* c, int image, const void** srcOut, size_t* szOut) {
    const size_t n = size_t(c->buf.width) * c->buf.height;
    const void* src = nullptr;
    size_t sz = 0;
    bool known = true;
    switch (image) {
        case BLOKB_IMG_HIT: src = c->buf.hit; sz = n * 16; break;
        case BLOKB_IMG_HIT_PACKED: src = c->buf.hitPacked; sz = n * 8; break;
        case BLOKB_IMG_SHADOW: src = c->buf.shadow; sz = n; break;
        case BLOKB_IMG_COLOR: src = c->buf.color; sz = n * 16; break;
        case BLOKB_IMG_WORLDPOS: src = c->buf.worldPos; sz = n * 16; break;
        case BLOKB_IMG_NORMALROUGH: src = c->buf.normalRough; sz = n * 8; break;
        case BLOKB_IMG_ALBEDOMETAL: src = c->buf.albedoMetal; sz = n * 4; break;
        case BLOKB_IMG_MOTION: src = c->buf.motion; sz = n * 4; break;
        case BLOKB_IMG_ACCUM: src = c->buf.accum; sz = n * 16; break;
        case BLOKB_IMG_HISTORY_COLOR: src = c->post.histColor[c->post.i]; sz = n * 16; break;
        case BLOKB_IMG_MOMENTS: src = c->post.histMoments[c->post.i]; sz = n * 8; break;
        case BLOKB_IMG_HISTORY_LENGTH: src = c->post.histLen[c->post.i]; sz = n * 2; break;
        case BLOKB_IMG_VARIANCE: src = c->post.variance; sz = n * 4; break;
        case BLOKB_IMG_DENOISED: src = c->post.denoised; sz = n * 16; break;
        case BLOKB_IMG_TAA_HISTORY: src = c->post.taaHist[c->post.j]; sz = n * 16; break;
        case BLOKB_IMG_TAA_OUTPUT: src = c->post.taaOut; sz = n * 16; break;
        case BLOKB_IMG_TONEMAPPED: src = c->post.tonemapOut; sz = n * 4; break;
        case BLOKB_IMG_FINAL: src = c->post.final; sz = n * c->post.finalBytesPerPixel; break;
        default: known = false; break;
    }
    *srcOut = src; *szOut = sz;
    return known;
}
// Rows of the strips (first, first + stride, ...) of an image with `rowBytes` per pixel row: device -> host at the same offsets.
static int copy_strips(const void* dev, void* host, size_t rowBytes, uint32_t height, uint32_t first, uint32_t stride, cudaStream_t q) {
    const uint32_t strips = (height + 7) / 8;
    if (first >= strips) return 0;
    const size_t stripBytes = 8 * rowBytes;
    uint32_t mine = (strips - first + stride - 1) / stride;
    const uint32_t last = first + (mine - 1) * stride;
    const uint32_t lastRows = std::min<uint32_t>(8, height - last * 8);
    uint32_t full = lastRows == 8 ? mine : mine - 1;   // strips of 8 whole rows
    const char* d = static_cast<const char*>(dev);
    char* h = static_cast<char*>(host);
    if (full) CK(cudaMemcpy2DAsync(h + size_t(first) * stripBytes, size_t(stride) * stripBytes, d + size_t(first) * stripBytes, size_t(stride) * stripBytes,
                                   stripBytes, full, cudaMemcpyDeviceToHost, q));
    if (full != mine) CK(cudaMemcpyAsync(h + size_t(last) * stripBytes, d + size_t(last) * stripBytes, size_t(lastRows) * rowBytes, cudaMemcpyDeviceToHost, q));
    return 0;
}
int blokb_readback_strips(blokb_ctx* c, int image, void* host, size_t bytes) {
    if (!c || !host) return fail(BLOKB_E_ARG, "null argument");
    if (!c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame)");
    const void* src; size_t sz;
    if (!image_lookup(c, image, &src, &sz)) return fail(BLOKB_E_ARG, "unknown image id");
    if (!src) return fail(BLOKB_E_STATE, "image not produced yet (blokb_denoise / blokb_postprocess of the current frame)");
    if (bytes != sz) return fail(BLOKB_E_ARG, "host buffer size mismatch: expected " + std::to_string(sz) + " (the whole image; only this context's strips are written)");
    CK(cudaSetDevice(c->device));
    if (int rc = join_dispatches(c)) return rc;
    if (int rc = copy_strips(src, host, sz / c->buf.height, c->buf.height, c->frame.stripFirst, c->frame.stripStride, c->stream)) return rc;
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

// ---- cross-GPU reduction of the accumulation buffers -----------------------------------------------------
static const NcclApi* nccl_or_fail() {
    std::string err;
    const NcclApi* a = nccl_api(err);
    if (!a) fail(BLOKB_E_UNSUPPORTED, "NCCL unavailable: " + err);
    return a;
}
#define NK(expr)                                                                                                         \
    do {                                                                                                                 \
        int r_ = (expr);                                                                                                 \
        if (r_ != 0) return fail(BLOKB_E_CUDA, std::string(#expr) + ": " + (N->GetErrorString ? N->GetErrorString(r_) : "nccl error")); \
    } while (0)
static void drop_comm(blokb_ctx* c) {
    if (c->comm && c->commOwned) {
        std::string err;
        if (const NcclApi* N = nccl_api(err)) N->CommDestroy(c->comm);
    }
    c->comm = nullptr; c->commOwned = false; c->commRank = 0; c->commSize = 1;
}
int blokb_comm_unique_id(void* id128) {
    if (!id128) return fail(BLOKB_E_ARG, "null argument");
    const NcclApi* N = nccl_or_fail();
    if (!N) return BLOKB_E_UNSUPPORTED;
    NcclId id;
    NK(N->GetUniqueId(&id));
    std::memcpy(id128, &id, sizeof id);
    return 0;
}
int blokb_comm_init(blokb_ctx* c, const void* id128, int rank, int nranks) {
    if (!c || !id128) return fail(BLOKB_E_ARG, "null argument");
    if (nranks < 1 || rank < 0 || rank >= nranks || size_t(nranks) > kSlicePad) return fail(BLOKB_E_ARG, "bad rank / communicator size (at most 64 ranks)");
    const NcclApi* N = nccl_or_fail();
    if (!N) return BLOKB_E_UNSUPPORTED;
    CK(cudaSetDevice(c->device));
    if (int rc = idle(c)) return rc;
    drop_comm(c);
    NcclId id;
    std::memcpy(&id, id128, sizeof id);
    NcclComm comm = nullptr;
    NK(N->CommInitRank(&comm, nranks, id, rank));
    c->comm = comm; c->commOwned = true; c->commRank = rank; c->commSize = nranks;
    return 0;
}
int blokb_set_comm(blokb_ctx* c, void* ncclComm) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    CK(cudaSetDevice(c->device));
    if (int rc = idle(c)) return rc;
    drop_comm(c);
    if (!ncclComm) return 0;
    const NcclApi* N = nccl_or_fail();
    if (!N) return BLOKB_E_UNSUPPORTED;
    int rank = 0, size = 1;
    NK(N->CommUserRank(ncclComm, &rank));
    NK(N->CommCount(ncclComm, &size));
    if (size_t(size) > kSlicePad) return fail(BLOKB_E_ARG, "communicators of more than 64 ranks are not supported");
    c->comm = ncclComm; c->commOwned = false; c->commRank = rank; c->commSize = size;
    return 0;
}
int blokb_comm_rank(blokb_ctx* c, int* rank, int* nranks) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if (rank) *rank = c->commRank;
    if (nranks) *nranks = c->commSize;
    return 0;
}
static void slice_of(blokb_ctx* c, size_t* per, size_t* first, size_t* count) {
    const size_t n = size_t(c->buf.width) * c->buf.height, G = size_t(c->commSize);
    *per = (n + G - 1) / G;
    *first = std::min(n, *per * size_t(c->commRank));
    *count = std::min(n, *first + *per) - *first;
}
int blokb_reduce_slice(blokb_ctx* c, uint64_t* firstPixel, uint64_t* nPixels) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if (!c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame)");
    size_t per, first, count;
    slice_of(c, &per, &first, &count);
    if (firstPixel) *firstPixel = first;
    if (nPixels) *nPixels = count;
    return 0;
}
int blokb_reduce_accum(blokb_ctx* c, int mode, void* stream) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    if (!c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame)");
    if (mode != BLOKB_REDUCE_ALL && mode != BLOKB_REDUCE_SLICE) return fail(BLOKB_E_ARG, "unknown reduce mode");
    CK(cudaSetDevice(c->device));
    cudaStream_t q = pick(c, stream);
    const size_t n = size_t(c->buf.width) * c->buf.height;
    if (c->commSize == 1 || !c->comm) {   // one GPU: the reduction is the identity
        if (c->commSize > 1) return fail(BLOKB_E_STATE, "no communicator (blokb_set_comm / blokb_comm_init)");
        CK(launch_resolve(c->buf.accum, c->buf.color, n, q));
        c->launches++;
        return dispatched(c, q);
    }
    const NcclApi* N = nccl_or_fail();
    if (!N) return BLOKB_E_UNSUPPORTED;
    size_t per, first, count;
    slice_of(c, &per, &first, &count);
    const bool accumMc = c->bound[2].on && c->bound[2].mc, colorMc = c->bound[4].on && c->bound[4].mc;
    if (accumMc) {
        // in-switch sum + resolve in one kernel over peer memory, fenced by two tiny all-reduces: every rank's samples are in
        // its accum before anyone reads through the multicast address, and nobody resets accum while a peer may still read it
        // refused before anything is enqueued: a rank that bailed out between the collectives would leave its peers waiting in NCCL
        if (mode == BLOKB_REDUCE_ALL && !colorMc && c->bound[4].on && n % size_t(c->commSize) != 0)
            return fail(BLOKB_E_UNSUPPORTED, "colour image in caller-owned memory without a multicast address and w*h not divisible by the ranks: bind it with one, or use BLOKB_REDUCE_SLICE");
        if (!c->dFence) { CK(cudaMalloc((void**)&c->dFence, 16)); CK(cudaMemsetAsync(c->dFence, 0, 16, q)); }
        NK(N->AllReduce(c->dFence, c->dFence, 1, kNcclFloat32, kNcclSum, c->comm, q));
        CK(launch_reduce_resolve_mc(c->bound[2].mc, c->buf.color, (mode == BLOKB_REDUCE_ALL && colorMc) ? c->bound[4].mc : nullptr, first, count, q));
        c->launches++;
        if (mode == BLOKB_REDUCE_ALL && !colorMc) {
            // in-place all-gather: the send pointer is recvbuf + rank * per even where `first` was clamped to n (tiny frames with many
            // ranks) — the context's own colour image carries the slack, and a caller-owned one got here only with n % G == 0
            NK(N->AllGather(c->buf.color + 4 * size_t(c->commRank) * per, c->buf.color, per * 4, kNcclFloat32, c->comm, q));
        }
        NK(N->AllReduce(c->dFence, c->dFence, 1, kNcclFloat32, kNcclSum, c->comm, q));
        return dispatched(c, q);
    }
    const bool padded = !c->bound[2].on && !c->bound[4].on;   // the context's own allocations carry kSlicePad pixels of slack
    if (!padded && n % size_t(c->commSize) != 0) {
        // caller-owned images of exactly n pixels cannot be cut into G equal slices: all-reduce in place, resolve everywhere
        NK(N->AllReduce(c->buf.accum, c->buf.accum, n * 4, kNcclFloat32, kNcclSum, c->comm, q));
        CK(launch_resolve(c->buf.accum, c->buf.color, n, q));
        c->launches++;
        return dispatched(c, q);
    }
    if (per > c->sliceCap) {
        CK(cudaDeviceSynchronize());
        if (c->dSlice) cudaFree(c->dSlice);
        c->dSlice = nullptr; c->sliceCap = 0;
        CK(cudaMalloc((void**)&c->dSlice, per * 16));
        c->sliceCap = per;
    }
    NK(N->ReduceScatter(c->buf.accum, c->dSlice, per * 4, kNcclFloat32, kNcclSum, c->comm, q));
    CK(launch_resolve(c->dSlice, c->buf.color + 4 * size_t(c->commRank) * per, count, q));
    c->launches++;
    if (mode == BLOKB_REDUCE_ALL) NK(N->AllGather(c->buf.color + 4 * size_t(c->commRank) * per, c->buf.color, per * 4, kNcclFloat32, c->comm, q));
    return dispatched(c, q);
}
int blokb_readback_slice(blokb_ctx* c, int image, void* host, size_t bytes) {
    if (!c || !host) return fail(BLOKB_E_ARG, "null argument");
    if (!c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame)");
    const void* src; size_t sz;
    if (!image_lookup(c, image, &src, &sz)) return fail(BLOKB_E_ARG, "unknown image id");
    if (!src) return fail(BLOKB_E_STATE, "image not produced yet (blokb_denoise / blokb_postprocess of the current frame)");
    size_t per, first, count;
    slice_of(c, &per, &first, &count);
    const size_t bpp = sz / (size_t(c->buf.width) * c->buf.height);
    if (bytes != count * bpp) return fail(BLOKB_E_ARG, "host buffer size mismatch: expected " + std::to_string(count * bpp));
    CK(cudaSetDevice(c->device));
    if (int rc = join_dispatches(c)) return rc;
    if (count) CK(cudaMemcpyAsync(host, static_cast<const char*>(src) + first * bpp, count * bpp, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}

int blokb_readback(blokb_ctx* c, int image, void* host, size_t bytes) {
    if (!c || !host) return fail(BLOKB_E_ARG, "null argument");
    if (!c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame)");
    const void* src = nullptr;
    size_t sz = 0;
    if (!image_lookup(c, image, &src, &sz)) return fail(BLOKB_E_ARG, "unknown image id");
    if (!src) return fail(BLOKB_E_STATE, "image not produced yet (blokb_denoise / blokb_postprocess of the current frame)");
    if (bytes != sz) return fail(BLOKB_E_ARG, "host buffer size mismatch: expected " + std::to_string(sz));
    CK(cudaSetDevice(c->device));
    if (int rc = join_dispatches(c)) return rc;   // ordered after dispatches enqueued on caller streams
    CK(cudaMemcpyAsync(host, src, sz, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}
int blokb_write_image(blokb_ctx* c, int image, const void* host, size_t bytes) {
    if (!c || !host) return fail(BLOKB_E_ARG, "null argument");
    if (!c->haveFrame) return fail(BLOKB_E_STATE, "no frame constants (blokb_set_frame)");
    const size_t n = size_t(c->buf.width) * c->buf.height;
    void* dst = nullptr;
    size_t sz = 0;
    switch (image) {
        case BLOKB_IMG_COLOR: dst = c->buf.color; sz = n * 16; break;
        case BLOKB_IMG_WORLDPOS: dst = c->buf.worldPos; sz = n * 16; break;
        case BLOKB_IMG_NORMALROUGH: dst = c->buf.normalRough; sz = n * 8; break;
        case BLOKB_IMG_ALBEDOMETAL: dst = c->buf.albedoMetal; sz = n * 4; break;
        case BLOKB_IMG_MOTION: dst = c->buf.motion; sz = n * 4; break;
        case BLOKB_IMG_ACCUM: dst = c->buf.accum; sz = n * 16; break;
        case BLOKB_IMG_HIT: dst = c->buf.hit; sz = n * sizeof(blokb_hit); break;
        case BLOKB_IMG_SHADOW: dst = c->buf.shadow; sz = n; break;
        case BLOKB_IMG_HIT_PACKED: dst = c->buf.hitPacked; sz = n * sizeof(blokb_hit_packed); break;
        default: return fail(BLOKB_E_ARG, "image is not writable");
    }
    if (bytes != sz) return fail(BLOKB_E_ARG, "host buffer size mismatch: expected " + std::to_string(sz));
    CK(cudaSetDevice(c->device));
    if (int rc = join_dispatches(c)) return rc;
    CK(cudaMemcpyAsync(dst, host, sz, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return 0;
}
int blokb_sync(blokb_ctx* c) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    CK(cudaSetDevice(c->device));
    return idle(c);   // the context's own stream and every dispatch enqueued on a caller stream
}
uint64_t blokb_launch_count(blokb_ctx* c) { return c ? c->launches : 0; }

// Profiling aid: when enabled, kernels add {DDA steps, rays} to two device counters.
int blokb_stats_enable(blokb_ctx* c, int on) {
    if (!c) return fail(BLOKB_E_ARG, "ctx is null");
    c->statsOn = on != 0;
    CK(cudaSetDevice(c->device));
    if (int rc = join_dispatches(c)) return rc;
    CK(cudaMemsetAsync(c->dStats, 0, 4 * sizeof(unsigned long long), c->stream));
    CK(cudaStreamSynchronize(c->stream));   // zeroed before a dispatch on any stream can add to them
    return 0;
}
int blokb_stats_read(blokb_ctx* c, uint64_t* stepsAndRays /*[2]*/) {
    if (!c || !stepsAndRays) return fail(BLOKB_E_ARG, "null argument");
    CK(cudaSetDevice(c->device));
    unsigned long long v[2];
    if (int rc = join_dispatches(c)) return rc;
    CK(cudaMemcpyAsync(v, c->dStats, sizeof v, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    stepsAndRays[0] = v[0]; stepsAndRays[1] = v[1];
    return 0;
}
void* blokb_stream(blokb_ctx* c) { return c ? c->stream : nullptr; }

// ---- frames in flight ---------------------------------------------------------------------------------------
struct blokb_pipeline {
    blokb_ctx* ctx = nullptr;
    int depth = 0;
    uint32_t width = 0, height = 0;
    std::vector<blokb_hit*> dHit;
    std::vector<cudaEvent_t> traced, copied;
    cudaStream_t copyStream = nullptr;
    uint64_t submitted = 0;
    bool packed = false;  // 8-byte blokb_hit_packed records instead of blokb_hit
    size_t recordBytes() const { return packed ? sizeof(blokb_hit_packed) : sizeof(blokb_hit); }
};

static int pipeline_create(blokb_ctx* c, int depth, uint32_t width, uint32_t height, bool packed, blokb_pipeline** out) {
    if (!c || !out) return fail(BLOKB_E_ARG, "null argument");
    *out = nullptr;
    if (depth < 1 || depth > 8 || width == 0 || height == 0 || width > 65536 || height > 65536) return fail(BLOKB_E_ARG, "bad pipeline shape");
    CK(cudaSetDevice(c->device));
    blokb_pipeline* p = new (std::nothrow) blokb_pipeline;
    if (!p) return fail(BLOKB_E_ARG, "out of host memory");
    p->ctx = c; p->depth = depth; p->width = width; p->height = height; p->packed = packed;
    auto bail = [&](cudaError_t e) { blokb_pipeline_destroy(p); return fail(BLOKB_E_CUDA, cudaGetErrorString(e)); };
    cudaError_t e = cudaStreamCreateWithFlags(&p->copyStream, cudaStreamNonBlocking);
    if (e != cudaSuccess) return bail(e);
    for (int i = 0; i < depth; i++) {
        blokb_hit* d = nullptr;
        cudaEvent_t a = nullptr, b = nullptr;
        if ((e = cudaMalloc((void**)&d, size_t(width) * height * p->recordBytes())) != cudaSuccess) return bail(e);
        p->dHit.push_back(d);
        if ((e = cudaEventCreateWithFlags(&a, cudaEventDisableTiming)) != cudaSuccess) return bail(e);
        p->traced.push_back(a);
        if ((e = cudaEventCreateWithFlags(&b, cudaEventDisableTiming)) != cudaSuccess) return bail(e);
        p->copied.push_back(b);
    }
    *out = p;
    return 0;
}

int blokb_pipeline_create(blokb_ctx* c, int depth, uint32_t width, uint32_t height, blokb_pipeline** out) {
    return pipeline_create(c, depth, width, height, false, out);
}
int blokb_pipeline_create_packed(blokb_ctx* c, int depth, uint32_t width, uint32_t height, blokb_pipeline** out) {
    return pipeline_create(c, depth, width, height, true, out);
}

static int pipeline_submit(blokb_pipeline* p, const blokb_frame_ubo* f, void* hostOut, bool packed, uint64_t* ticket) {
    if (!p || !f || !hostOut) return fail(BLOKB_E_ARG, "null argument");
    if (p->packed != packed) return fail(BLOKB_E_ARG, packed ? "pipeline was not created with blokb_pipeline_create_packed"
                                                             : "packed pipeline: use blokb_pipeline_submit_packed");
    blokb_ctx* c = p->ctx;
    if (packed && c->stats.nodes > BLOKB_PACKED_MAX_NODES) return fail(BLOKB_E_UNSUPPORTED, "world has more nodes than a packed record can address");
    if (!c->haveWorld) return fail(BLOKB_E_STATE, "no world uploaded (blokb_upload_world)");
    if (f->screenWidth != p->width || f->screenHeight != p->height) return fail(BLOKB_E_ARG, "frame size differs from the pipeline's");
    CK(cudaSetDevice(c->device));
    const int slot = int(p->submitted % uint64_t(p->depth));
    FrameDev d{};
    std::memcpy(d.invView, f->invView, 64);
    std::memcpy(d.invProj, f->invProj, 64);
    std::memcpy(d.prevViewProj, f->prevViewProj, 64);
    std::memcpy(d.camPos, f->camPos, 12);
    d.frameCount = f->frameCount; d.sampleCount = f->sampleCount; d.width = f->screenWidth; d.height = f->screenHeight;
    d.stripFirst = c->frame.stripFirst; d.stripStride = c->frame.stripStride;   // a tiled frame: this GPU's strips only, traced AND copied
    if (p->submitted >= uint64_t(p->depth)) CK(cudaStreamWaitEvent(c->stream, p->copied[slot], 0));  // slot's previous copy must have drained
    if (packed) CK(launch_primary_packed(c->dw, d, own_only(p->dHit[slot]), c->stream));
    else CK(launch_primary(c->dw, d, false, own_only(p->dHit[slot]), own_only(nullptr), st(c), c->stream));
    c->launches++;
    CK(cudaEventRecord(p->traced[slot], c->stream));
    CK(cudaStreamWaitEvent(p->copyStream, p->traced[slot], 0));
    if (d.stripStride == 1 && d.stripFirst == 0)
        CK(cudaMemcpyAsync(hostOut, p->dHit[slot], size_t(p->width) * p->height * p->recordBytes(), cudaMemcpyDeviceToHost, p->copyStream));
    else if (int rc = copy_strips(p->dHit[slot], hostOut, size_t(p->width) * p->recordBytes(), p->height, d.stripFirst, d.stripStride, p->copyStream)) return rc;
    CK(cudaEventRecord(p->copied[slot], p->copyStream));
    if (ticket) *ticket = p->submitted;
    p->submitted++;
    return 0;
}

int blokb_pipeline_submit(blokb_pipeline* p, const blokb_frame_ubo* f, blokb_hit* hostOut, uint64_t* ticket) {
    return pipeline_submit(p, f, hostOut, false, ticket);
}
int blokb_pipeline_submit_packed(blokb_pipeline* p, const blokb_frame_ubo* f, blokb_hit_packed* hostOut, uint64_t* ticket) {
    return pipeline_submit(p, f, hostOut, true, ticket);
}

int blokb_pipeline_wait(blokb_pipeline* p, uint64_t ticket) {
    if (!p) return fail(BLOKB_E_ARG, "pipeline is null");
    if (ticket >= p->submitted || ticket + uint64_t(p->depth) < p->submitted) return fail(BLOKB_E_ARG, "ticket not in flight");
    CK(cudaSetDevice(p->ctx->device));
    CK(cudaEventSynchronize(p->copied[ticket % uint64_t(p->depth)]));
    return 0;
}

int blokb_pipeline_drain(blokb_pipeline* p) {
    if (!p) return fail(BLOKB_E_ARG, "pipeline is null");
    CK(cudaSetDevice(p->ctx->device));
    CK(cudaStreamSynchronize(p->ctx->stream));
    CK(cudaStreamSynchronize(p->copyStream));
    return 0;
}

void blokb_pipeline_destroy(blokb_pipeline* p) {
    if (!p) return;
    cudaSetDevice(p->ctx->device);
    if (p->copyStream) cudaStreamSynchronize(p->copyStream);
    cudaStreamSynchronize(p->ctx->stream);
    for (auto d : p->dHit) cudaFree(d);
    for (auto e : p->traced) cudaEventDestroy(e);
    for (auto e : p->copied) cudaEventDestroy(e);
    if (p->copyStream) cudaStreamDestroy(p->copyStream);
    delete p;
}

int blokb_node_prims(const blokb_svo_node* nodes, size_t nNodes, const blokb_subchunk* subs, size_t nSubs, uint32_t* primOfNode) {
    if (node_prims(nodes, nNodes, subs, nSubs, primOfNode) != 0) return fail(BLOKB_E_ARG, "null array with non-zero count");
    return 0;
}
int blokb_unpack_hits(const blokb_hit_packed* in, size_t n, const blokb_svo_node* nodes, size_t nNodes, const uint32_t* primOfNode,
                      blokb_hit* out) {
    if (n && (!in || !out || !nodes)) return fail(BLOKB_E_ARG, "null array with non-zero count");
    if (unpack_hits(in, n, nodes, nNodes, primOfNode, out) != 0) return fail(BLOKB_E_ARG, "packed record names a node outside the array (stored as a miss)");
    return 0;
}

void* blokb_host_alloc(size_t bytes) {
    void* h = nullptr;
    if (cudaHostAlloc(&h, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) { g_err = "cudaHostAlloc failed"; return nullptr; }
    return h;
}
void blokb_host_free(void* h) { if (h) cudaFreeHost(h); }
}
