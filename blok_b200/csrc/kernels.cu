// Kernels of the traversal path. One thread per ray / pixel; a warp covers an 8x4 pixel tile so that its 32
// rays walk neighbouring cells (coherent mask loads, little divergence). All scene reads are read-only
// (ld.global.nc) and the compact scene is sized to stay L2-resident.
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cub/cub.cuh>

#include <algorithm>
#include <mutex>

#include "kernels.h"
#include "shade.cuh"

namespace blokb {

// Measurement variant: the wavefront's slot arrays (records, keys, order, front radiance: gigabytes per batch, touched once per pass)
// as streaming accesses (ld.global.cs / st.global.cs: evict first), so that they do not push the 42 MB scene out of the 126 MB L2.
// Measured on c4: k_bounce's DRAM reads 10.2 -> 9.7 GB per batch, time unchanged (36.4 vs 36.0 ms per 16 spp): the reads are the
// record gathers themselves, not scene misses. Off.
#ifndef BLOKB_STREAM_SLOTS
#define BLOKB_STREAM_SLOTS 0
#endif
#if BLOKB_STREAM_SLOTS
#define BLOKB_STREAM_LD(p) __ldcs(p)
#define BLOKB_STREAM_ST(p, v) __stcs(p, v)
#else
#define BLOKB_STREAM_LD(p) (*(p))
#define BLOKB_STREAM_ST(p, v) (*(p) = (v))
#endif

// Register budgets of the traversal kernels of the path tracer (min resident blocks per SM -> registers per thread): measured on
// the c4 frame with variant libraries, tools/wave_bench.py.
#ifndef BLOKB_FRONT_MINBLOCKS
#define BLOKB_FRONT_MINBLOCKS 5
#endif
#ifndef BLOKB_PATH_MINBLOCKS
#define BLOKB_PATH_MINBLOCKS 6
#endif
#ifndef BLOKB_PRIMARY_MINBLOCKS
#define BLOKB_PRIMARY_MINBLOCKS 1    // k_primary<false>: left alone ptxas stays at 45 registers = 11 blocks per SM; any bound rounds to 40 (spills) or 48 (10 blocks)
#endif
#ifndef BLOKB_SHADOW_MINBLOCKS
#define BLOKB_SHADOW_MINBLOCKS 8     // k_primary<true>
#endif
#ifndef BLOKB_BOUNCE_MINBLOCKS
#define BLOKB_BOUNCE_MINBLOCKS 1
#endif
// pixel of this thread: block = 32 x BLOKB_BLOCK_ROWS pixels (4 warps of 8x4 pixels by default), two such blocks per 8-row strip
// blockIdx.y counts the strips THIS launch renders (FrameDev::stripFirst / stripStride): pixel coordinates, seeds and
// output addresses are those of the full frame, so any partition of the strips over launches or GPUs writes the same bytes.
#ifndef BLOKB_BLOCK_ROWS
#define BLOKB_BLOCK_ROWS 4  // pixel kernels: block = 32 x BLOKB_BLOCK_ROWS pixels (4 or 8 rows of an 8-row strip; 4 measured 0.6-0.9 % faster: DESIGN.md 5)
#endif
#ifndef BLOKB_BLOCK_W
#define BLOKB_BLOCK_W 32   // pixel kernels: block width in pixels (32 = four 8x4 warp tiles side by side; 16 = two, measured: DESIGN.md 5)
#endif
constexpr uint32_t kBlockW = BLOKB_BLOCK_W;
constexpr uint32_t kBlocksPerStrip = 8u / BLOKB_BLOCK_ROWS, kPixThreads = kBlockW * BLOKB_BLOCK_ROWS;
__device__ __forceinline__ bool thread_pixel(const FrameDev& F, uint32_t& px, uint32_t& py) {
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
#ifndef BLOKB_TILE_W_LOG2
#define BLOKB_TILE_W_LOG2 3  // warp tile = 8 x 4 pixels (measured against 4x8, 16x2 and 32x1: DESIGN.md 5)
#endif
    constexpr uint32_t TW = 1u << BLOKB_TILE_W_LOG2, TH = 32u / TW, tilesX = kBlockW / TW;
    px = blockIdx.x * kBlockW + (warp % tilesX) * TW + (lane & (TW - 1u));
    py = (F.stripFirst + (blockIdx.y / kBlocksPerStrip) * F.stripStride) * 8u + (blockIdx.y % kBlocksPerStrip) * BLOKB_BLOCK_ROWS +
         (warp / tilesX) * TH + (lane >> BLOKB_TILE_W_LOG2);
    return px < F.width && py < F.height;
}

// Shared-memory node stacks: [level][thread] rows of 16 bytes, W.levels rows (launchers pass levels * 256 * 16 bytes).
extern __shared__ uint4 g_stackMem[];
__device__ __forceinline__ NodeStack thread_stack() {
    return NodeStack{uint32_t(__cvta_generic_to_shared(g_stackMem + threadIdx.x)), uint32_t(blockDim.x * sizeof(uint4))};
}

__device__ __forceinline__ void store_hit(blokb_hit* dst, bool hit, const HitRec& h) {
    uint4 v;
    if (hit) v = make_uint4(h.node, h.material, __float_as_uint(h.t), h.facePrim);
    else v = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, __float_as_uint(-1.0f), 0xFFFFFFFFu);
    *reinterpret_cast<uint4*>(dst) = v;
}

// One 16-byte record to every destination of the image (own buffer + peers over NVLink); n is launch-uniform.
// With a multicast binding (o.mc) ONE multimem.st reaches every GPU's copy through the NVSwitch.
__device__ __forceinline__ void store16_all(const PeerOut& o, size_t index, uint4 v) {
    if (o.mc) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(static_cast<uint4*>(o.mc) + index),
                     "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
                     : "memory");
        return;
    }
    reinterpret_cast<uint4*>(o.dst[0])[index] = v;
    if (o.n == 1) return;  // the single-GPU dispatch: one store, no peer loop
#pragma unroll
    for (uint32_t k = 1; k < kMaxPeers; k++)
        if (k < o.n) reinterpret_cast<uint4*>(o.dst[k])[index] = v;
}
// One byte per pixel. multimem.st has no byte form: with a multicast binding the four pixels of a lane quad (x = 4q..4q+3
// of one row; whole quads are in or out of the frame because blokb_bind_image requires width % 4 == 0) travel as one word.
__device__ __forceinline__ void store8_all(const PeerOut& o, size_t index, uint8_t b) {
    if (o.mc) {
        const uint32_t lane = threadIdx.x & 31u, quad = 0xFu << (lane & 28u);
        uint32_t w = b;
        w |= __shfl_down_sync(quad, w, 1) << 8;
        w |= __shfl_down_sync(quad, w, 2) << 16;
        if ((lane & 3u) == 0u)
            asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(static_cast<uint8_t*>(o.mc) + index), "r"(w) : "memory");
        return;
    }
    static_cast<uint8_t*>(o.dst[0])[index] = b;
    if (o.n == 1) return;
#pragma unroll
    for (uint32_t k = 1; k < kMaxPeers; k++)
        if (k < o.n) static_cast<uint8_t*>(o.dst[k])[index] = b;
}
// One 8-byte record (blokb_hit_packed) to every destination: half the NVLink bytes of the 16-byte record.
__device__ __forceinline__ void store8B_all(const PeerOut& o, size_t index, uint2 v) {
    if (o.mc) {
        asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(static_cast<uint2*>(o.mc) + index), "f"(__uint_as_float(v.x)),
                     "f"(__uint_as_float(v.y))
                     : "memory");
        return;
    }
    reinterpret_cast<uint2*>(o.dst[0])[index] = v;
    if (o.n == 1) return;
#pragma unroll
    for (uint32_t k = 1; k < kMaxPeers; k++)
        if (k < o.n) reinterpret_cast<uint2*>(o.dst[k])[index] = v;
}
__device__ __forceinline__ void store_hit_all(const PeerOut& o, size_t index, bool hit, const HitRec& h) {
    store16_all(o, index, hit ? make_uint4(h.node, h.material, __float_as_uint(h.t), h.facePrim)
                              : make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, __float_as_uint(-1.0f), 0xFFFFFFFFu));
}

template <bool kStats>
__global__ void __launch_bounds__(256) k_trace_rays(const __grid_constant__ DeviceWorld W, const float4* __restrict__ rays, size_t n,
                                                    blokb_hit* __restrict__ out, unsigned long long* stats) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = __ldg(&rays[2 * i]), b = __ldg(&rays[2 * i + 1]);
    HitRec h;
    uint32_t steps = 0;
    const bool hit = traverse<kStats>(W, thread_stack(), a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, h, steps);
    store_hit(out + i, hit, h);
    if (kStats) atomicAdd(stats, (unsigned long long)steps);
}

// ---- ray reordering for incoherent ray lists (bounce rays) -------------------------------------------------------------
// Sort key (24 bits = three radix passes) = 8^3 direction cell (9 bits) above the Morton code of the origin's coarse cell
// (3 x 5 bits, cells of >= 64 voxels; the sort is stable, so inside a key the list keeps its own — usually already spatial —
// order): rays of one warp then head the same way from neighbouring cells and walk the same nodes. Measured on 4.67 M cosine-sampled bounce
// rays of the c3 frame: 3.78 -> 5.35 Grays/s for the traversal itself.
__device__ __forceinline__ uint32_t dilate7(uint32_t v) {  // 7 bits -> every third bit
    v &= 0x7Fu;
    v = (v | (v << 8)) & 0x0000700Fu;
    v = (v | (v << 4)) & 0x000430C3u;
    v = (v | (v << 2)) & 0x00049249u;
    return v;
}
__global__ void __launch_bounds__(256) k_ray_keys(const __grid_constant__ DeviceWorld W, const float4* __restrict__ rays, size_t n, int cellShift,
                                                  uint32_t* __restrict__ keys, uint32_t* __restrict__ idx) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = __ldg(&rays[2 * i]), b = __ldg(&rays[2 * i + 1]);
    // direction as given: a.w, b.x, b.y. Not normalised on purpose (only the cell matters; any finite scale sorts consistently enough)
    const float dx = a.w, dy = b.x, dz = b.y;
    const float inv = rsqrtf(fmaxf(dx * dx + dy * dy + dz * dz, 1e-30f));
    auto cell = [inv](float d) { const int q = int((d * inv + 1.0f) * 4.0f); return uint32_t(q < 0 ? 0 : (q > 7 ? 7 : q)); };
    const uint32_t dcell = cell(dx) | (cell(dy) << 3) | (cell(dz) << 6);
    auto coarse = [cellShift](float o, int g, int dim) {
        float r = o - float(g);
        r = r < 0.0f ? 0.0f : (r > float(dim - 1) ? float(dim - 1) : r);   // rays that start outside sort by their nearest grid cell
        return uint32_t(int(r) >> cellShift);
    };
    const uint32_t m = dilate7(coarse(a.x, W.origin[0], W.dim[0])) | (dilate7(coarse(a.y, W.origin[1], W.dim[1])) << 1) |
                       (dilate7(coarse(a.z, W.origin[2], W.dim[2])) << 2);
    keys[i] = (dcell << 15) | (m & 0x7FFFu);
    idx[i] = uint32_t(i);
}
// Thread i traces ray order[i]; the record goes back to the ray's own slot, so the output is that of k_trace_rays.
template <bool kStats>
__global__ void __launch_bounds__(256) k_trace_rays_ordered(const __grid_constant__ DeviceWorld W, const float4* __restrict__ rays,
                                                            const uint32_t* __restrict__ order, size_t n, blokb_hit* __restrict__ out,
                                                            unsigned long long* stats) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t r = __ldg(order + i);
    const float4 a = __ldg(&rays[2 * size_t(r)]), b = __ldg(&rays[2 * size_t(r) + 1]);
    HitRec h;
    uint32_t steps = 0;
    const bool hit = traverse<kStats>(W, thread_stack(), a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, h, steps);
    store_hit(out + r, hit, h);
    if (kStats) atomicAdd(stats, (unsigned long long)steps);
}

// Sample-0 primary rays (raygen.rgen:193-229) -> hit records; optionally the sun shadow ray of :280-298.
template <bool kShadow, bool kStats>
__global__ void __launch_bounds__(kPixThreads, kShadow ? BLOKB_SHADOW_MINBLOCKS : BLOKB_PRIMARY_MINBLOCKS) k_primary(const __grid_constant__ DeviceWorld W, const __grid_constant__ FrameDev F,
                                                 const __grid_constant__ PeerOut out, const __grid_constant__ PeerOut shadow,
                                                 unsigned long long* stats) {
    uint32_t px, py;
    if (!thread_pixel(F, px, py)) return;
    const F3 d = primary_dir(F, float(px) + 0.5f, float(py) + 0.5f);
    uint32_t steps = 0;
    NodeStack stk = thread_stack();
    const size_t pi = size_t(py) * F.width + px;
    if constexpr (!kShadow) {   // the camera ray alone: no loop state to carry (45 registers = 11 blocks per SM)
        HitRec h;
        const bool hit = traverse<kStats>(W, stk, F.camPos[0], F.camPos[1], F.camPos[2], d.x, d.y, d.z, 0.001f, 10000.0f, h, steps, /*locate=*/false);
        store_hit_all(out, pi, hit, h);
        if (kStats) atomicAdd(stats, (unsigned long long)steps);
        return;
    }
    // camera ray, then (kShadow) the sun shadow ray of a lit hit: two passes around ONE inlined copy of the walk, so that the lanes
    // re-converge between the rays (shade.cuh, path_rays); the shadow ray starts from the camera ray's node stack (WalkResume)
    WalkResume resume;
    F3 o = f3(F.camPos[0], F.camPos[1], F.camPos[2]), dir = d;
    float tmax = 10000.0f;
    uint8_t sv = 0;  // 0 = no shadow ray cast, 1 = lit, 2 = shadowed
    int pass = 0;
#pragma unroll 1
    for (;;) {
        HitRec h;
        const bool hit = traverse<kStats>(W, stk, o.x, o.y, o.z, dir.x, dir.y, dir.z, 0.001f, tmax, h, steps, /*locate=*/pass > 0, kShadow ? &resume : nullptr);
        if (pass == 1) { sv = hit ? 2 : 1; break; }
        store_hit_all(out, pi, hit, h);
        if (!kShadow || !hit) break;
        F3 N = face_normal(h.facePrim & 7u);
        if (dot3(N, d) > 0.0f) N = -N;
        const F3 sunDir = sun_dir();
        if (!(fmaxf(dot3(N, sunDir), 0.0f) > 0.0f)) break;
        o = f3(F.camPos[0], F.camPos[1], F.camPos[2]) + d * h.t + N * 0.001f;
        dir = sunDir;
        tmax = 1000.0f;
        pass = 1;
    }
    if (kShadow) store8_all(shadow, pi, sv);
    if (kStats) atomicAdd(stats, (unsigned long long)steps);
}

// k_primary<false> with the 8-byte record of blokb_hit_packed (the readback is PCIe-bound: half the bytes per frame).
__global__ void __launch_bounds__(kPixThreads) k_primary_packed(const __grid_constant__ DeviceWorld W, const __grid_constant__ FrameDev F,
                                                        const __grid_constant__ PeerOut out) {
    uint32_t px, py;
    if (!thread_pixel(F, px, py)) return;
    const F3 d = primary_dir(F, float(px) + 0.5f, float(py) + 0.5f);
    HitRec h;
    uint32_t steps = 0;
    const bool hit = traverse<false>(W, thread_stack(), F.camPos[0], F.camPos[1], F.camPos[2], d.x, d.y, d.z, 0.001f, 10000.0f, h, steps, false);
    store8B_all(out, size_t(py) * F.width + px, hit ? make_uint2(h.node | (h.facePrim & 7u) << 29, __float_as_uint(h.t))
                                                    : make_uint2(0xFFFFFFFFu, __float_as_uint(-1.0f)));
}

__device__ __forceinline__ uint32_t unorm8(float v) { return __float2uint_rn(__saturatef(v) * 255.0f); }

// raygen.rgen as two kernels (blokb_render): the four G-buffer images that describe the FIRST hit of sample 0 (:253-263,
// :394-414) come from k_gbuffer — one unjittered primary ray per pixel — and the colour image from k_path<.., kColor>
// running the frame's samples. One kernel doing both kept the first-hit record alive across every sample's traversals:
// 103 registers, 2 blocks per SM, 26.4 ms for 8 spp at 4K against 21.4 ms for the same samples in k_path. Sample 0's
// primary ray is traced twice (0.7 ms); every stored value is computed by the same expressions as before.
__global__ void __launch_bounds__(kPixThreads) k_gbuffer(const __grid_constant__ DeviceWorld W, const __grid_constant__ FrameDev F, float4* __restrict__ worldPos,
                                                 uint2* __restrict__ normalRough, uint32_t* __restrict__ albedoMetal, uint32_t* __restrict__ motion) {
    uint32_t px, py;
    if (!thread_pixel(F, px, py)) return;
    const F3 cam = f3(F.camPos[0], F.camPos[1], F.camPos[2]);
    const F3 d = primary_dir(F, float(px) + 0.5f, float(py) + 0.5f);
    HitRec h;
    uint32_t steps = 0;
    const bool had = traverse<false>(W, thread_stack(), cam.x, cam.y, cam.z, d.x, d.y, d.z, 0.001f, 10000.0f, h, steps, /*locate=*/false);
    F3 pos, N, fa;
    float depth, roughness = 0.f, metallic = 0.f;
    if (had) {
        const Surface m = fetch_surface(W, h.material);
        N = face_normal(h.facePrim & 7u);
        pos = cam + d * h.t;
        if (dot3(N, d) > 0.0f) N = -N;
        depth = h.t; roughness = m.roughness; metallic = m.metallic;
        fa = is_emissive(m.emission) ? m.emission : m.albedo;
    } else {
        depth = 10000.0f;
        const F4 fw = mat_mul(F.invView, 0.f, 0.f, -1.f, 0.f);
        pos = cam + normalize3(f3(fw.x, fw.y, fw.z)) * 10000.0f;
        N = f3(0.f, 1.f, 0.f);
        fa = sky_color(normalize3(pos - cam));
    }
    const size_t pi = size_t(py) * F.width + px;
    worldPos[pi] = make_float4(pos.x, pos.y, pos.z, depth);
    {
        const __half2 a = __floats2half2_rn(N.x, N.y), b = __floats2half2_rn(N.z, roughness);
        normalRough[pi] = make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
    }
    albedoMetal[pi] = unorm8(fa.x) | (unorm8(fa.y) << 8) | (unorm8(fa.z) << 16) | (unorm8(metallic) << 24);
    float mx = 0.f, my = 0.f;
    if (had && depth < 9999.0f) {
        const F4 pc = mat_mul(F.prevViewProj, pos.x, pos.y, pos.z, 1.0f);
        const float nx = pc.x / pc.w, ny = pc.y / pc.w;
        mx = (float(px) + 0.5f) / float(F.width) - (nx * 0.5f + 0.5f);
        my = (float(py) + 0.5f) / float(F.height) - (ny * 0.5f + 0.5f);
    }
    const __half2 m = __floats2half2_rn(mx, my);
    motion[pi] = *reinterpret_cast<const uint32_t*>(&m);
}

// Progressive path tracing: samples [sBegin, sEnd) of every pixel added into accum = (sum.rgb, spp)
// (the reference CUDA tracer's accumulator, cuda_tracer.cu:372-380), with raygen.rgen's per-(pixel, frame,
// sample) seeding so that any partition of the sample range over launches or GPUs draws the same paths.
// kColor: the launch covers ALL of the frame's samples and stores color = sum / count with the firefly clamp
// (raygen.rgen:383-392) into accum.dst[0] — the colour image — instead of accumulating.
// No minBlocksPerSM on purpose: left alone, ptxas settles on 64 registers (4 blocks per SM, a few spilled words) = 21.4 ms for
// 8 spp at 4K; __launch_bounds__(256, 1) gives 90 registers / 25.9 ms, (256, 2) 22.2, (256, 3) 22.6, (256, 5) 24.1.
template <bool kStats, bool kColor>
__global__ void __launch_bounds__(kPixThreads, BLOKB_PATH_MINBLOCKS) k_path(const __grid_constant__ DeviceWorld W, const __grid_constant__ FrameDev F, uint32_t sBegin, uint32_t sEnd,
                                              const __grid_constant__ PeerOut accum, unsigned long long* stats) {
    uint32_t px, py;
    if (!thread_pixel(F, px, py)) return;
    F3 acc = f3(0.f);
    uint32_t rays = 0, steps = 0;
    const NodeStack stk = thread_stack();
    for (uint32_t s = sBegin; s < sEnd; s++) acc = acc + path_sample<kStats>(W, stk, F, px, py, s, nullptr, rays, steps);
    const size_t pi = size_t(py) * F.width + px;
    if (kColor) {
        F3 c = acc / float(sEnd - sBegin);
        const float mv = max3(c);
        if (mv > 100.0f) c = c * (100.0f / mv);
        static_cast<float4*>(accum.dst[0])[pi] = make_float4(c.x, c.y, c.z, 1.0f);
        if (kStats) { atomicAdd(stats, (unsigned long long)steps); atomicAdd(stats + 1, (unsigned long long)rays); }
        return;
    }
    float4 a = static_cast<const float4*>(accum.dst[0])[pi];  // the pixel's owner accumulates locally, then replicates
    a.x += acc.x; a.y += acc.y; a.z += acc.z; a.w += float(sEnd - sBegin);
    store16_all(accum, pi, make_uint4(__float_as_uint(a.x), __float_as_uint(a.y), __float_as_uint(a.z), __float_as_uint(a.w)));
    if (kStats) { atomicAdd(stats, (unsigned long long)steps); atomicAdd(stats + 1, (unsigned long long)rays); }
}

// ---- wavefront path tracing: bounce rays regrouped before they are traced ------------------------------------------------
// k_path walks camera, shadow and bounce ray of a sample back to back in one thread; the bounce rays of a warp's 32 pixels
// scatter over the hemisphere, and the warp then issues with ~12 of 32 lanes (profiles/r01t_ncu_k_path_c4_8spp.txt). Here a batch
// of `nS` samples of every pixel runs as three stages over one slot array (slot = sample-in-batch * nT + launch-wide thread id):
//   k_path_front   iteration 0 of raygen's loop for every (pixel, sample) of the batch — coherent camera + sun shadow rays —
//                  stores the radiance gathered so far (front[slot].rgb, .w = "a bounce ray follows") and, for paths that go
//                  on, the bounce ray {origin, direction, throughput} and its sort key;
//   radix sort     of (key, slot) over the whole batch: key = direction cell above the Morton code of the origin's coarse
//                  cell, so that consecutive rays leave neighbouring surface points in about the same direction; slots
//                  without a bounce ray carry the key 0xFFFFFFFF and end up behind every ray;
//   k_bounce       thread i walks ray order[i] (iteration 1: only a miss or an emissive hit still changes the radiance,
//                  shade.cuh:path_bounce) and adds its term into front[slot] — one writer per slot, no atomics;
//   k_path_gather  per pixel, the batch's slots are summed in sample order into the running sum of the call; the last batch
//                  applies k_path's epilogue (accumulate into accum, or colour = mean with the firefly clamp).
// Same expressions, same order of additions as k_path: the accumulated image is bit-identical (tests/test_gpu_wavefront.py).
constexpr uint32_t kDeadKey = 0xFFFFFFFFu;
__device__ __forceinline__ uint32_t dilate10(uint32_t v) {  // 10 bits -> every third bit
    v &= 0x3FFu;
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
__device__ __forceinline__ uint32_t wave_key(const DeviceWorld& W, const WaveKey& K, const F3& o, const F3& d) {
    // d is unit length (sample_cosine / reflect3 of unit vectors)
    auto cell = [&K](float v) { const int q = int((v + 1.0f) * K.dirScale); return uint32_t(q < 0 ? 0 : (q >= int(K.dirCells) ? int(K.dirCells) - 1 : q)); };
    const uint32_t dcell = cell(d.x) + K.dirCells * (cell(d.y) + K.dirCells * cell(d.z));
    auto coarse = [&K](float p, int g, int dim) {
        float r = p - float(g);
        r = r < 0.0f ? 0.0f : (r > float(dim - 1) ? float(dim - 1) : r);
        return uint32_t(int(r) >> K.cellShift);
    };
    const uint32_t m = dilate10(coarse(o.x, W.origin[0], W.dim[0])) | (dilate10(coarse(o.y, W.origin[1], W.dim[1])) << 1) |
                       (dilate10(coarse(o.z, W.origin[2], W.dim[2])) << 2);
    return (dcell << K.mortonBits) | (m & ((1u << K.mortonBits) - 1u));
}
__device__ __forceinline__ uint32_t launch_thread_id() { return (blockIdx.y * gridDim.x + blockIdx.x) * kPixThreads + threadIdx.x; }

template <bool kStats>
__global__ void __launch_bounds__(kPixThreads, BLOKB_FRONT_MINBLOCKS) k_path_front(const __grid_constant__ DeviceWorld W, const __grid_constant__ FrameDev F, uint32_t s0, uint32_t nS,
                                                            const __grid_constant__ WaveBufs B, const __grid_constant__ WaveKey K, unsigned long long* stats) {
    const uint32_t gtid = launch_thread_id();
    uint32_t px, py;
    const bool inFrame = thread_pixel(F, px, py);
    uint32_t rays = 0, steps = 0;
    const NodeStack stk = thread_stack();
    for (uint32_t k = 0; k < nS; k++) {
        const size_t slot = size_t(k) * B.nT + gtid;
        F3 rad = f3(0.f);
        uint32_t key = kDeadKey;
        BLOKB_STREAM_ST(B.order + slot, uint32_t(slot));
        if (inFrame)
            path_rays<kStats, false>(W, stk, F, px, py, s0 + k, nullptr, rays, steps, rad, [&](const BounceRay& br) {
                // called before the sun shadow ray is traced: the bounce ray leaves the registers here
                key = wave_key(W, K, br.origin, br.dir);
                float4* r = B.rec + 3 * slot;
                BLOKB_STREAM_ST(r + 0, make_float4(br.origin.x, br.origin.y, br.origin.z, br.dir.x));
                BLOKB_STREAM_ST(r + 1, make_float4(br.dir.y, br.dir.z, br.thr.x, br.thr.y));
                BLOKB_STREAM_ST(r + 2, make_float4(br.thr.z, 0.f, 0.f, 0.f));
                BLOKB_STREAM_ST(B.keys + slot, key);
            });
        if (key == kDeadKey) BLOKB_STREAM_ST(B.keys + slot, kDeadKey);
        BLOKB_STREAM_ST(B.front + slot, make_float4(rad.x, rad.y, rad.z, key != kDeadKey ? 1.0f : 0.0f));

    }
    if (kStats) { atomicAdd(stats, (unsigned long long)steps); atomicAdd(stats + 1, (unsigned long long)rays); }
}

template <bool kStats>
__global__ void __launch_bounds__(256, BLOKB_BOUNCE_MINBLOCKS) k_bounce(const __grid_constant__ DeviceWorld W, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ order,
                                                size_t n, const float4* __restrict__ rec, float4* __restrict__ front, unsigned long long* stats) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (BLOKB_STREAM_LD(keys + i) == kDeadKey) return;  // slots without a bounce ray sort behind every ray: whole warps leave here
    const uint32_t slot = BLOKB_STREAM_LD(order + i);
    const float4 a = BLOKB_STREAM_LD(rec + 3 * size_t(slot)), b = BLOKB_STREAM_LD(rec + 3 * size_t(slot) + 1), c = BLOKB_STREAM_LD(rec + 3 * size_t(slot) + 2);
    BounceRay br;
    br.origin = f3(a.x, a.y, a.z); br.dir = f3(a.w, b.x, b.y); br.thr = f3(b.z, b.w, c.x);
    uint32_t rays = 0, steps = 0;
    bool add;
    const F3 t = path_bounce<kStats>(W, thread_stack(), br, rays, steps, add);
    if (add) {
        float4 f = BLOKB_STREAM_LD(front + slot);
        f.x += t.x; f.y += t.y; f.z += t.z;
        BLOKB_STREAM_ST(front + slot, f);
    }
    if (kStats) { atomicAdd(stats, (unsigned long long)steps); atomicAdd(stats + 1, (unsigned long long)rays); }
}

// k_bounce with idle lanes refilled. Sorted or not, the bounce rays of a warp end at very different times (10.1 of 32 lanes per issued
// instruction in k_bounce: the warp lasts as long as its longest ray). Here a warp is persistent: lanes whose ray has ended wait, and
// as soon as fewer than `refillBelow` lanes are still walking, ALL idle lanes fetch the next rays of the sorted list together (one
// atomicAdd per warp), set up their walks together, and the warp goes on — the walk's loop is the product's while-while loop, run to
// the next refill point instead of to the end of the longest ray. Which thread walks a ray does not change its result: same records.
template <bool kStats>
__global__ void __launch_bounds__(256, BLOKB_BOUNCE_MINBLOCKS) k_bounce_refill(const __grid_constant__ DeviceWorld W, const uint32_t* __restrict__ keys,
                                                                               const uint32_t* __restrict__ order, uint32_t n, const float4* __restrict__ rec,
                                                                               float4* __restrict__ front, unsigned* counter, uint32_t refillBelow,
                                                                               unsigned long long* stats) {
    const uint32_t lane = threadIdx.x & 31u;
    const NodeStack stk = thread_stack();
    Walker wk;
    BounceRay br;
    uint32_t slot = 0, steps = 0, rays = 0;
    bool have = false;   // this lane is in the middle of a walk
    bool more = true;    // (warp-uniform) the list may still hold rays
    auto settle = [&](bool hit) {   // iteration 1's term into the ray's slot (shade.cuh, bounce_term)
        bool add;
        const F3 t = bounce_term(W, br, hit, hit ? wk.hit_material(W) : 0u, add);
        if (add) {
            float4 f = front[slot];
            f.x += t.x; f.y += t.y; f.z += t.z;
            front[slot] = f;
        }
    };
    for (;;) {
        if (more) {
            const unsigned idle = __ballot_sync(0xFFFFFFFFu, !have);
            uint32_t start = 0;
            if (lane == 0) start = atomicAdd(counter, (unsigned)__popc(idle));
            start = __shfl_sync(0xFFFFFFFFu, start, 0);
            const uint32_t i = start + __popc(idle & ((1u << lane) - 1u));
            bool dry = false;
            if (!have) {
                if (i < n && __ldg(keys + i) != kDeadKey) {
                    slot = __ldg(order + i);
                    const float4 a = __ldg(rec + 3 * size_t(slot)), b = __ldg(rec + 3 * size_t(slot) + 1), c = __ldg(rec + 3 * size_t(slot) + 2);
                    br.origin = f3(a.x, a.y, a.z); br.dir = f3(a.w, b.x, b.y); br.thr = f3(b.z, b.w, c.x);
                    rays++;
                    if (wk.init(W, br.origin.x, br.origin.y, br.origin.z, br.dir.x, br.dir.y, br.dir.z, 0.001f, 10000.0f, /*locate=*/true)) have = true;
                    else settle(false);   // the ray never enters the grid
                } else {
                    dry = true;           // past the last ray (slots without one sort behind every ray)
                }
            }
            more = !__any_sync(0xFFFFFFFFu, dry);
        }
        if (!__any_sync(0xFFFFFFFFu, have)) {
            if (!more) break;
            continue;
        }
        bool fin = false, hit = false;
        for (;;) {
            if (have && !fin) {
                while (wk.occupied()) {
                    if (wk.sh == 0) { hit = true; break; }
                    wk.descend(W, stk);
                }
                if (hit) fin = true;
                else {
                    if (kStats) steps++;
                    if (!wk.step(W, stk)) fin = true;
                }
            }
            const unsigned walking = __ballot_sync(0xFFFFFFFFu, have && !fin);
            if (walking == 0u || (more && uint32_t(__popc(walking)) < refillBelow)) break;
        }
        if (have && fin) {
            settle(hit);
            have = false;
        }
    }
    if (kStats) { atomicAdd(stats, (unsigned long long)steps); atomicAdd(stats + 1, (unsigned long long)rays); }
}

// firstBatch / lastBatch: position of this batch inside one blokb_trace_path / blokb_render call (sTotal samples in all).
template <bool kColor>
__global__ void __launch_bounds__(kPixThreads) k_path_gather(const __grid_constant__ FrameDev F, uint32_t nS, bool firstBatch, bool lastBatch, uint32_t sTotal,
                                                             const __grid_constant__ WaveBufs B, const __grid_constant__ PeerOut accum) {
    const uint32_t gtid = launch_thread_id();
    uint32_t px, py;
    if (!thread_pixel(F, px, py)) return;
    F3 acc = f3(0.f);
    if (!firstBatch) { const float4 p = B.partial[gtid]; acc = f3(p.x, p.y, p.z); }
    for (uint32_t k = 0; k < nS; k++) {
        const float4 r = B.front[size_t(k) * B.nT + gtid];
        acc = acc + f3(r.x, r.y, r.z);
    }
    if (!lastBatch) { B.partial[gtid] = make_float4(acc.x, acc.y, acc.z, 0.f); return; }
    const size_t pi = size_t(py) * F.width + px;
    if (kColor) {
        F3 c = acc / float(sTotal);
        const float mv = max3(c);
        if (mv > 100.0f) c = c * (100.0f / mv);
        static_cast<float4*>(accum.dst[0])[pi] = make_float4(c.x, c.y, c.z, 1.0f);
        return;
    }
    float4 a = static_cast<const float4*>(accum.dst[0])[pi];
    a.x += acc.x; a.y += acc.y; a.z += acc.z; a.w += float(sTotal);
    store16_all(accum, pi, make_uint4(__float_as_uint(a.x), __float_as_uint(a.y), __float_as_uint(a.z), __float_as_uint(a.w)));
}

// color = sum / spp, firefly clamp (raygen.rgen:383-392).
__global__ void __launch_bounds__(256) k_resolve(const float4* __restrict__ accum, float4* __restrict__ color, size_t n) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = accum[i];
    F3 c = f3(a.x, a.y, a.z) / a.w;
    const float mv = max3(c);
    if (mv > 100.0f) c = c * (100.0f / mv);
    color[i] = make_float4(c.x, c.y, c.z, 1.0f);
}

// Cross-GPU sum + resolve of a slice of the accumulation image in ONE pass over peer memory: every GPU's accum lives in
// symmetric memory bound to one NVLink multicast object (blokb_bind_image), and multimem.ld_reduce returns the sum of all
// GPUs' copies of a 16-byte texel — the NVSwitch adds them on the way — so rank g resolves pixels [first, first + n) without a
// separate reduce-scatter. colour = sum.rgb / sum.a with the firefly clamp (raygen.rgen:383-392), written to this GPU's colour
// image (and, when `colorMc` is set, to every GPU's through a multicast store: the all-gather of the slices).
__global__ void __launch_bounds__(256) k_reduce_resolve_mc(const float4* __restrict__ accumMc, float4* __restrict__ color, float4* __restrict__ colorMc,
                                                           size_t first, size_t n) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "l"(accumMc + first + i) : "memory");
    F3 c = f3(a.x, a.y, a.z) / a.w;
    const float mv = max3(c);
    if (mv > 100.0f) c = c * (100.0f / mv);
    if (colorMc) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(colorMc + first + i), "f"(c.x), "f"(c.y), "f"(c.z), "f"(1.0f) : "memory");
    } else {
        color[first + i] = make_float4(c.x, c.y, c.z, 1.0f);
    }
}

// Cross-GPU frame fence (blokb_frame_fence): one multimem.red adds this rank's arrival to the counter on every GPU at once; the
// release makes everything this GPU did before (earlier kernels of the stream: their multicast stores) visible first; the acquire
// loop on the LOCAL copy ends when all ranks of this round have arrived. Bounded: ~1 s of polling, then the error word is set.
// A fence that timed out leaves the counters short for good, so later fences on this rank could only time out again: once the error
// word is set they still add their arrival (peers keep counting this rank) but return without polling. `hostFlag` is a word of mapped
// page-locked host memory the C-ABI layer reads before it enqueues the next fence (may be null).
__global__ void k_fence(uint32_t* local, void* mc, uint32_t target, uint32_t* hostFlag) {
    asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc), "r"(1u) : "memory");
    if (*reinterpret_cast<volatile uint32_t*>(local + 1)) return;
    uint32_t v = 0;
    for (uint32_t spin = 0; spin < 8000000u; spin++) {   // ~1 s: a poll is a system-scope load plus a 64 ns sleep
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(local) : "memory");
        if (int32_t(v - target) >= 0) return;
        __nanosleep(64);
    }
    local[1] = 1u;
    if (hostFlag) { *reinterpret_cast<volatile uint32_t*>(hostFlag) = 1u; __threadfence_system(); }
}

// ---- launchers ------------------------------------------------------------------------------------
static inline dim3 pixel_grid(const FrameDev& F) {
    const uint32_t strips = (F.height + 7) / 8;
    const uint32_t mine = F.stripFirst < strips ? (strips - F.stripFirst + F.stripStride - 1) / F.stripStride : 0;
    return dim3((F.width + kBlockW - 1) / kBlockW, mine * kBlocksPerStrip);
}
static inline size_t pix_stack_bytes(const DeviceWorld& W) { return size_t(W.levels) * kPixThreads * sizeof(uint4); }
static inline size_t stack_bytes(const DeviceWorld& W) { return size_t(W.levels) * 256 * sizeof(uint4); }  // <= 48 KB (kMaxLevels = 12)

cudaError_t launch_trace_rays(const DeviceWorld& W, const float* rays, size_t n, blokb_hit* out, unsigned long long* stats, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    if (stats) k_trace_rays<true><<<blocks, 256, stack_bytes(W), st>>>(W, reinterpret_cast<const float4*>(rays), n, out, stats);
    else k_trace_rays<false><<<blocks, 256, stack_bytes(W), st>>>(W, reinterpret_cast<const float4*>(rays), n, out, nullptr);
    return cudaGetLastError();
}
size_t reorder_scratch_bytes(size_t n, size_t* cubBytes) {
    size_t tb = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tb, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int64_t)n, 0, 24);
    if (cubBytes) *cubBytes = tb;
    const size_t a = (n * 4 + 255) & ~size_t(255), b = a;
    return 2 * a + 2 * b + tb;
}
cudaError_t launch_trace_rays_reordered(const DeviceWorld& W, const float* rays, size_t n, blokb_hit* out, void* scratch, unsigned long long* stats,
                                        cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    if (n > 0xFFFFFFFFull) return cudaErrorInvalidValue;
    size_t tb = 0;
    reorder_scratch_bytes(n, &tb);
    const size_t a = (n * 4 + 255) & ~size_t(255), b = a;
    char* p = static_cast<char*>(scratch);
    uint32_t* k0 = reinterpret_cast<uint32_t*>(p); uint32_t* k1 = reinterpret_cast<uint32_t*>(p + a);
    uint32_t* i0 = reinterpret_cast<uint32_t*>(p + 2 * a); uint32_t* i1 = reinterpret_cast<uint32_t*>(p + 2 * a + b);
    void* tmp = p + 2 * a + 2 * b;
    // coarse origin cells: 5 bits per axis must cover the grid, and cells are at least 64 voxels wide
    int shift = 6;
    while ((std::max(W.dim[0], std::max(W.dim[1], W.dim[2])) >> shift) > 32) shift++;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    k_ray_keys<<<blocks, 256, 0, st>>>(W, reinterpret_cast<const float4*>(rays), n, shift, k0, i0);
    cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tb, k0, k1, i0, i1, (int64_t)n, 0, 24, st);
    if (e != cudaSuccess) return e;
    if (stats) k_trace_rays_ordered<true><<<blocks, 256, stack_bytes(W), st>>>(W, reinterpret_cast<const float4*>(rays), i1, n, out, stats);
    else k_trace_rays_ordered<false><<<blocks, 256, stack_bytes(W), st>>>(W, reinterpret_cast<const float4*>(rays), i1, n, out, nullptr);
    return cudaGetLastError();
}
// ---- reordered ray lists for callers that cannot allocate (blokb_trace_rays_reordered_static) --------------------------------------
// Same contract as launch_trace_rays_reordered — every record lands in its ray's own slot, so the output bytes do not depend on the
// order — but the scratch (bin counters, keys, order) is the file-scope array below instead of memory the context allocates, and the
// ordering is a counting sort written here (15-bit key, 32768 bins) instead of cub's radix sort: no library call on this path at all.
// The array exists once per GPU that loads this library and is shared by every context and every call on that GPU: calls that
// overlap on the device race on it. It keeps the last call's contents, so the bin counters are zeroed at the start of every call.
constexpr size_t kStaticMaxRays = size_t(BLOKB_STATIC_MAX_RAYS);
constexpr uint32_t kStaticBins = 1u << 15;   // 8^3 direction cells x 4^3 origin cells
constexpr size_t kStaticBinBytes = size_t(kStaticBins) * 4;
constexpr size_t kStaticBytes = kStaticBinBytes + 2 * kStaticMaxRays * 4;   // 128 KiB + 16 MiB
static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

__global__ void __launch_bounds__(256) k_static_keys(const __grid_constant__ DeviceWorld W, const float4* __restrict__ rays, uint32_t n, int cellShift,
                                                     uint32_t* __restrict__ bins, uint32_t* __restrict__ keys) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = __ldg(&rays[2 * size_t(i)]), b = __ldg(&rays[2 * size_t(i) + 1]);
    const float dx = a.w, dy = b.x, dz = b.y;
    const float inv = rsqrtf(fmaxf(dx * dx + dy * dy + dz * dz, 1e-30f));
    auto dcell = [inv](float d) { const int q = int((d * inv + 1.0f) * 4.0f); return uint32_t(q < 0 ? 0 : (q > 7 ? 7 : q)); };
    auto ocell = [cellShift](float o, int g, int dim) {
        float r = o - float(g);
        r = r < 0.0f ? 0.0f : (r > float(dim - 1) ? float(dim - 1) : r);
        const uint32_t q = uint32_t(int(r) >> cellShift);
        return q > 3u ? 3u : q;
    };
    const uint32_t key = (dcell(dx) | (dcell(dy) << 3) | (dcell(dz) << 6)) << 6 |
                         (ocell(a.x, W.origin[0], W.dim[0]) | (ocell(a.y, W.origin[1], W.dim[1]) << 2) | (ocell(a.z, W.origin[2], W.dim[2]) << 4));
    keys[i] = key;
    atomicAdd(bins + key, 1u);
}
// counts -> exclusive prefix sums, in place; one block of 1024 threads, 32 consecutive bins each
__global__ void __launch_bounds__(1024) k_static_scan(uint32_t* __restrict__ bins) {
    __shared__ uint32_t warpSum[32];
    const uint32_t t = threadIdx.x, lane = t & 31u, w = t >> 5;
    uint32_t sum = 0;
    for (int k = 0; k < 32; k++) sum += bins[t * 32 + k];
    uint32_t incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= uint32_t(d)) incl += o; }
    if (lane == 31) warpSum[w] = incl;
    __syncthreads();
    if (w == 0) {
        uint32_t x = warpSum[lane];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, x, d); if (lane >= uint32_t(d)) x += o; }
        warpSum[lane] = x;
    }
    __syncthreads();
    uint32_t run = incl - sum + (w ? warpSum[w - 1] : 0u);
    for (int k = 0; k < 32; k++) { const uint32_t v = bins[t * 32 + k]; bins[t * 32 + k] = run; run += v; }
}
// order[slot of ray i inside its bin] = i. Rays of one bin land in whatever order the atomics resolve: harmless, each record goes to out[i].
__global__ void __launch_bounds__(256) k_static_scatter(const uint32_t* __restrict__ keys, uint32_t n, uint32_t* __restrict__ bins, uint32_t* __restrict__ order) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    order[atomicAdd(bins + keys[i], 1u)] = i;
}
// Device address of g_static_scratch on the CURRENT device (which must be `device`). In host code the array's name is only a handle
// for the runtime, not an address a kernel can use: the address is asked for once per device and remembered.
static cudaError_t static_scratch_address(int device, unsigned char** out) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static unsigned char* known[kMaxDevices] = {};
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!known[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess) return e;
        if (cur != device) return cudaErrorInvalidDevice;
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        known[device] = static_cast<unsigned char*>(p);
    }
    *out = known[device];
    return cudaSuccess;
}
cudaError_t launch_trace_rays_reordered_static(const DeviceWorld& W, const float* rays, size_t n, blokb_hit* out, int device, unsigned long long* stats,
                                               cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    if (n > kStaticMaxRays) return cudaErrorInvalidValue;
    unsigned char* p = nullptr;
    cudaError_t e = static_scratch_address(device, &p);
    if (e != cudaSuccess) return e;
    uint32_t* bins = reinterpret_cast<uint32_t*>(p);
    uint32_t* keys = reinterpret_cast<uint32_t*>(p + kStaticBinBytes);
    uint32_t* order = keys + kStaticMaxRays;
    int shift = 0;   // four origin cells per axis must cover the grid
    while ((std::max(W.dim[0], std::max(W.dim[1], W.dim[2])) >> shift) > 4) shift++;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    e = cudaMemsetAsync(bins, 0, kStaticBinBytes, st);   // the previous call left its running offsets here
    if (e != cudaSuccess) return e;
    k_static_keys<<<blocks, 256, 0, st>>>(W, reinterpret_cast<const float4*>(rays), uint32_t(n), shift, bins, keys);
    k_static_scan<<<1, 1024, 0, st>>>(bins);
    k_static_scatter<<<blocks, 256, 0, st>>>(keys, uint32_t(n), bins, order);
    if (stats) k_trace_rays_ordered<true><<<blocks, 256, stack_bytes(W), st>>>(W, reinterpret_cast<const float4*>(rays), order, n, out, stats);
    else k_trace_rays_ordered<false><<<blocks, 256, stack_bytes(W), st>>>(W, reinterpret_cast<const float4*>(rays), order, n, out, nullptr);
    return cudaGetLastError();
}
cudaError_t launch_primary(const DeviceWorld& W, const FrameDev& F, bool shadow, const PeerOut& out, const PeerOut& shadowOut,
                           unsigned long long* stats, cudaStream_t st) {
    const dim3 g = pixel_grid(F);
    if (g.y == 0) return cudaSuccess;
    if (stats) {
        if (shadow) k_primary<true, true><<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, out, shadowOut, stats);
        else k_primary<false, true><<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, out, shadowOut, stats);
    } else {
        if (shadow) k_primary<true, false><<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, out, shadowOut, nullptr);
        else k_primary<false, false><<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, out, shadowOut, nullptr);
    }
    return cudaGetLastError();
}
cudaError_t launch_primary_packed(const DeviceWorld& W, const FrameDev& F, const PeerOut& out, cudaStream_t st) {
    const dim3 g = pixel_grid(F);
    if (g.y == 0) return cudaSuccess;
    k_primary_packed<<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, out);
    return cudaGetLastError();
}
cudaError_t launch_render(const DeviceWorld& W, const FrameDev& F, float* color, float* worldPos, uint16_t* normalRough, uint8_t* albedoMetal,
                          uint16_t* motion, unsigned long long* stats, cudaStream_t st) {
    const dim3 g = pixel_grid(F);
    if (g.y == 0) return cudaSuccess;
    k_gbuffer<<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, (float4*)worldPos, (uint2*)normalRough, (uint32_t*)albedoMetal, (uint32_t*)motion);
    PeerOut out;
    out.dst[0] = color;
    if (stats) k_path<true, true><<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, 0u, F.sampleCount, out, stats);
    else k_path<false, true><<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, 0u, F.sampleCount, out, nullptr);
    return cudaGetLastError();
}
cudaError_t launch_path(const DeviceWorld& W, const FrameDev& F, uint32_t sBegin, uint32_t sEnd, const PeerOut& accum, unsigned long long* stats,
                        cudaStream_t st) {
    const dim3 g = pixel_grid(F);
    if (g.y == 0) return cudaSuccess;
    if (stats) k_path<true, false><<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, sBegin, sEnd, accum, stats);
    else k_path<false, false><<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, sBegin, sEnd, accum, nullptr);
    return cudaGetLastError();
}
// ---- wavefront launchers --------------------------------------------------------------------------------------------
WaveKey make_wave_key(const DeviceWorld& W, uint32_t dirCells, int cellShift, uint32_t mortonBitsPerAxis) {
    WaveKey K;
    K.dirCells = std::min(std::max(dirCells, 1u), 16u);
    K.dirScale = 0.5f * float(K.dirCells);
    K.mortonBits = 3u * std::min(std::max(mortonBitsPerAxis, 1u), 6u);
    // the coarse origin cells must cover the grid with mortonBitsPerAxis bits per axis
    int shift = std::max(cellShift, 0);
    const int maxDim = std::max(W.dim[0], std::max(W.dim[1], W.dim[2]));
    while (((maxDim - 1) >> shift) >= (1 << (K.mortonBits / 3))) shift++;
    K.cellShift = shift;
    uint32_t dbits = 0;
    while ((1u << dbits) < K.dirCells * K.dirCells * K.dirCells) dbits++;
    K.bits = dbits + K.mortonBits;
    return K;
}
size_t wave_threads(const FrameDev& F) {
    const dim3 g = pixel_grid(F);
    return size_t(g.x) * g.y * kPixThreads;
}
static inline size_t al256(size_t b) { return (b + 255) & ~size_t(255); }
size_t wave_scratch_bytes(size_t nT, uint32_t batch) {
    const size_t n = nT * batch;
    size_t tb = 0;
    cub::DoubleBuffer<uint32_t> dk(nullptr, nullptr), dv(nullptr, nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, dv, (int64_t)n, 0, 32);
    return al256(n * 16) + al256(n * 48) + 4 * al256(n * 4) + al256(nT * 16) + 256 + al256(tb);
}
cudaError_t launch_path_wavefront(const DeviceWorld& W, const FrameDev& F, uint32_t sBegin, uint32_t sEnd, bool toColor, const PeerOut& accum,
                                  void* scratch, uint32_t batch, const WaveKey& K, uint32_t refillBelow, unsigned long long* stats, uint64_t* launches,
                                  cudaStream_t st) {
    const dim3 g = pixel_grid(F);
    if (g.y == 0 || sEnd <= sBegin) return cudaSuccess;
    const size_t nT = size_t(g.x) * g.y * kPixThreads;
    if (nT * batch > 0xFFFFFFFFull) return cudaErrorInvalidValue;
    const size_t cap = nT * batch;
    char* p = static_cast<char*>(scratch);
    WaveBufs B;
    B.nT = uint32_t(nT);
    B.front = reinterpret_cast<float4*>(p); p += al256(cap * 16);
    B.rec = reinterpret_cast<float4*>(p); p += al256(cap * 48);
    uint32_t* k0 = reinterpret_cast<uint32_t*>(p); p += al256(cap * 4);
    uint32_t* k1 = reinterpret_cast<uint32_t*>(p); p += al256(cap * 4);
    uint32_t* i0 = reinterpret_cast<uint32_t*>(p); p += al256(cap * 4);
    uint32_t* i1 = reinterpret_cast<uint32_t*>(p); p += al256(cap * 4);
    B.partial = reinterpret_cast<float4*>(p); p += al256(nT * 16);
    unsigned* counter = reinterpret_cast<unsigned*>(p); p += 256;
    void* tmp = p;
    B.keys = k0; B.order = i0;
    const uint32_t sTotal = sEnd - sBegin;
    for (uint32_t s0 = sBegin; s0 < sEnd; s0 += batch) {
        const uint32_t nS = std::min(batch, sEnd - s0);
        const size_t n = nT * nS;
        if (stats) k_path_front<true><<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, s0, nS, B, K, stats);
        else k_path_front<false><<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, s0, nS, B, K, nullptr);
        size_t tb = 0;
        cub::DoubleBuffer<uint32_t> dk(k0, k1), dv(i0, i1);
        // one bit more than the key uses: live keys are < 2^K.bits, the dead key is 0xFFFFFFFF, so bit K.bits puts every dead slot strictly
        // behind every live one. With K.bits alone a live key whose bits are all ones (reachable with 1, 2 or 4 direction cells per axis)
        // ties with the dead key and the stable sort interleaves them — the refill variant stops at the first dead key it meets.
        const int sortBits = int(std::min(K.bits + 1u, 32u));
        cub::DeviceRadixSort::SortPairs(nullptr, tb, dk, dv, (int64_t)n, 0, sortBits);
        cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tb, dk, dv, (int64_t)n, 0, sortBits, st);
        if (e != cudaSuccess) return e;
        const unsigned blocks = (unsigned)((n + 255) / 256);
        if (refillBelow > 0) {
            // persistent warps: as many blocks as stay resident, each warp pulls rays off the sorted list until it is empty
            static int perSM[2] = {0, 0}, sms = 0;
            if (!sms) {
                int dev = 0;
                cudaGetDevice(&dev);
                cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            }
            int& occ = perSM[stats ? 1 : 0];
            if (!occ) {
                if (stats) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_bounce_refill<true>, 256, stack_bytes(W));
                else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_bounce_refill<false>, 256, stack_bytes(W));
                if (occ < 1) occ = 1;
            }
            const unsigned pblocks = std::min<unsigned>(blocks, unsigned(sms * occ));
            cudaMemsetAsync(counter, 0, sizeof(unsigned), st);
            if (stats) k_bounce_refill<true><<<pblocks, 256, stack_bytes(W), st>>>(W, dk.Current(), dv.Current(), uint32_t(n), B.rec, B.front, counter, refillBelow, stats);
            else k_bounce_refill<false><<<pblocks, 256, stack_bytes(W), st>>>(W, dk.Current(), dv.Current(), uint32_t(n), B.rec, B.front, counter, refillBelow, nullptr);
        } else if (stats) k_bounce<true><<<blocks, 256, stack_bytes(W), st>>>(W, dk.Current(), dv.Current(), n, B.rec, B.front, stats);
        else k_bounce<false><<<blocks, 256, stack_bytes(W), st>>>(W, dk.Current(), dv.Current(), n, B.rec, B.front, nullptr);
        const bool first = s0 == sBegin, last = s0 + nS == sEnd;
        if (toColor) k_path_gather<true><<<g, kPixThreads, 0, st>>>(F, nS, first, last, sTotal, B, accum);
        else k_path_gather<false><<<g, kPixThreads, 0, st>>>(F, nS, first, last, sTotal, B, accum);
        if (launches) *launches += 4;  // front, radix sort (counted as one), bounce, gather
    }
    return cudaGetLastError();
}
cudaError_t launch_gbuffer(const DeviceWorld& W, const FrameDev& F, float* worldPos, uint16_t* normalRough, uint8_t* albedoMetal, uint16_t* motion,
                           cudaStream_t st) {
    const dim3 g = pixel_grid(F);
    if (g.y == 0) return cudaSuccess;
    k_gbuffer<<<g, kPixThreads, pix_stack_bytes(W), st>>>(W, F, (float4*)worldPos, (uint2*)normalRough, (uint32_t*)albedoMetal, (uint32_t*)motion);
    return cudaGetLastError();
}
cudaError_t launch_reduce_resolve_mc(const void* accumMc, float* color, void* colorMc, size_t first, size_t n, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    k_reduce_resolve_mc<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(static_cast<const float4*>(accumMc), reinterpret_cast<float4*>(color),
                                                                    static_cast<float4*>(colorMc), first, n);
    return cudaGetLastError();
}
cudaError_t launch_fence(uint32_t* local, void* multicast, uint32_t target, uint32_t* hostFlag, cudaStream_t st) {
    k_fence<<<1, 1, 0, st>>>(local, multicast, target, hostFlag);
    return cudaGetLastError();
}
cudaError_t launch_resolve(const float* accum, float* color, size_t n, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    k_resolve<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const float4*)accum, (float4*)color, n);
    return cudaGetLastError();
}

}  // namespace blokb
