// Launchers implemented in kernels.cu (device code); called from the C-ABI layer in api.cpp.
#pragma once
#include <cstddef>
#include <cuda_runtime.h>

#include "../../include/blokb200.h"
#include "scene.h"

namespace blokb {
// The fields of FrameUBO that raygen.rgen reads (resources.hpp:103-150).
struct FrameDev {
    float invView[16], invProj[16], prevViewProj[16];
    float camPos[3];
    uint32_t frameCount, sampleCount, width, height;
    // tile scheduler: of the frame's 8-row strips (32-pixel-wide blocks, 8 / BLOKB_BLOCK_ROWS of them stacked per strip) this launch renders strips
    // stripFirst, stripFirst + stripStride, ... — the whole frame by default (blokb_set_strips)
    uint32_t stripFirst = 0, stripStride = 1;
};
// Destinations of one output image: dst[0] is this GPU's own buffer; dst[1..n) are the same image in peer GPUs' memory
// (CUDA IPC mappings, blokb_set_peers). A tiled frame's kernels store every record to all of them, so the exchange of
// strips rides NVLink under the traversal instead of following it as a separate all-gather.
// mc != nullptr: the image lives in symmetric memory bound to an NVLink multicast object (blokb_bind_image); ONE
// multimem.st to mc is replicated by the NVSwitch into every GPU's copy (this GPU's included), so the sender's egress is
// 1x the strip instead of (G-1)x. dst[0] then only serves reads of the local copy (accumulate).
constexpr uint32_t kMaxPeers = 8;
struct PeerOut {
    void* dst[kMaxPeers] = {};
    uint32_t n = 1;
    void* mc = nullptr;
};
cudaError_t launch_trace_rays(const DeviceWorld& W, const float* rays, size_t n, blokb_hit* out, unsigned long long* stats, cudaStream_t st);
// Incoherent ray lists: sort by (direction cell, Morton code of the origin cell), trace in that order, write each record to the
// ray's own slot. scratch: reorder_scratch_bytes(n) bytes of device memory.
size_t reorder_scratch_bytes(size_t n, size_t* cubBytes);
cudaError_t launch_trace_rays_reordered(const DeviceWorld& W, const float* rays, size_t n, blokb_hit* out, void* scratch, unsigned long long* stats,
                                        cudaStream_t st);
// The same without any allocation: scratch is a file-scope __device__ array in kernels.cu (one per GPU, shared by all callers), the
// ordering a counting sort of its own. n <= BLOKB_STATIC_MAX_RAYS; `device` must be the current device.
cudaError_t launch_trace_rays_reordered_static(const DeviceWorld& W, const float* rays, size_t n, blokb_hit* out, int device, unsigned long long* stats,
                                               cudaStream_t st);
cudaError_t launch_primary(const DeviceWorld& W, const FrameDev& F, bool shadow, const PeerOut& out, const PeerOut& shadowOut, unsigned long long* stats,
                           cudaStream_t st);
cudaError_t launch_primary_packed(const DeviceWorld& W, const FrameDev& F, const PeerOut& out, cudaStream_t st);
cudaError_t launch_render(const DeviceWorld& W, const FrameDev& F, float* color, float* worldPos, uint16_t* normalRough, uint8_t* albedoMetal,
                          uint16_t* motion, unsigned long long* stats, cudaStream_t st);
cudaError_t launch_path(const DeviceWorld& W, const FrameDev& F, uint32_t sBegin, uint32_t sEnd, const PeerOut& accum, unsigned long long* stats,
                        cudaStream_t st);
// Wavefront path tracing (kernels.cu): slot arrays of one batch of samples. nT = threads of the pixel grid of the launch.
struct WaveBufs {
    float4* front = nullptr;     // [batch * nT]     (radiance.rgb of iteration 0, 1 = a bounce ray follows)
    float4* rec = nullptr;       // [batch * nT * 3] bounce ray: origin.xyz dir.x | dir.yz thr.xy | thr.z
    uint32_t* keys = nullptr;    // [batch * nT]     sort key, 0xFFFFFFFF = no bounce ray
    uint32_t* order = nullptr;   // [batch * nT]     slot index (the radix sort's values)
    float4* partial = nullptr;   // [nT]             running per-pixel sum across the batches of one call
    uint32_t nT = 0;
};
// key = direction cell (dirCells^3, x fastest) << mortonBits | Morton code of the origin's (1 << cellShift)-voxel cell
struct WaveKey {
    uint32_t dirCells = 8;
    float dirScale = 4.0f;
    int cellShift = 6;
    uint32_t mortonBits = 15;
    uint32_t bits = 24;  // key bits the radix sort looks at
};
WaveKey make_wave_key(const DeviceWorld& W, uint32_t dirCells, int cellShift, uint32_t mortonBitsPerAxis);
size_t wave_threads(const FrameDev& F);
size_t wave_scratch_bytes(size_t nT, uint32_t batch);
// samples [sBegin, sEnd) in batches of `batch`; toColor: colour = mean + firefly clamp into accum.dst[0] (blokb_render) instead of accumulating
cudaError_t launch_path_wavefront(const DeviceWorld& W, const FrameDev& F, uint32_t sBegin, uint32_t sEnd, bool toColor, const PeerOut& accum,
                                  void* scratch, uint32_t batch, const WaveKey& K, uint32_t refillBelow, unsigned long long* stats, uint64_t* launches,
                                  cudaStream_t st);
cudaError_t launch_gbuffer(const DeviceWorld& W, const FrameDev& F, float* worldPos, uint16_t* normalRough, uint8_t* albedoMetal, uint16_t* motion,
                           cudaStream_t st);
cudaError_t launch_resolve(const float* accum, float* color, size_t n, cudaStream_t st);
// one-thread cross-GPU barrier over a multicast word (blokb_frame_fence): word[0] counts arrivals, word[1] = 1 after a timeout, which
// is also written to *hostFlag (device address of a mapped page-locked host word; may be null)
cudaError_t launch_fence(uint32_t* local, void* multicast, uint32_t target, uint32_t* hostFlag, cudaStream_t st);
// colour[first, first + n) = resolve(sum over GPUs of accum) through an NVLink multicast address of the accumulation images
cudaError_t launch_reduce_resolve_mc(const void* accumMc, float* color, void* colorMc, size_t first, size_t n, cudaStream_t st);
}  // namespace blokb
