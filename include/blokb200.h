/* blokb200.h — C-ABI of the B200-native voxel ray-traversal path (drop-in boundary).
 *
 * The reference (collinlongoria/blok) has no plugin/FFI interface: its traversal sits behind three C++
 * seams inside one executable. Each entry point below replaces one of those seams (paths relative to
 * the reference root; see SURVEY.md §8(b) and INTEGRATION.md for the reference-side binding):
 *
 *   scene in      Renderer::addWorld / updateWorld(WorldSvoGpu&)        blok/include/renderer.hpp:40-71
 *                 uploadSvoBuffers / uploadMaterialBuffer                 blok/src/renderer_upload.cpp:237-312
 *   frame in      Renderer::render(const Camera&, float dt) -> FrameUBO   blok/include/renderer.hpp:34,
 *                 Denoiser::fillFrameUBO                                  blok/src/renderer_denoising.cpp:655-704
 *   dispatch      RayTracing::dispatchRayTracing(cmd, w, h, frame)        blok/src/renderer_raytracing.cpp:666-685
 *   images out    storage images, bindings 4-8                            assets/shaders/raygen.rgen:55-59
 *   accumulate    CudaTracer float4(sum.xyz, spp) + resetAccum            blok/src/cuda_tracer.cu:372-380, 478-482
 *
 * Conventions: every function returns 0 on success, <0 on error (BLOKB_E_*); blokb_last_error() gives the
 * thread-local message. No exceptions cross the boundary. A context is single-owner (not thread-safe);
 * one context per GPU. "Enqueue" calls are asynchronous on the given CUDA stream (a cudaStream_t passed as
 * void*; NULL = the context's own stream) exactly like recording into the reference's command buffer.
 * Streams: a dispatch on a caller stream is ordered after earlier work on THAT stream only. Calls that move data for the
 * caller — blokb_readback, blokb_write_image, blokb_stats_read, blokb_world_export, the uploads, blokb_set_frame when it
 * reallocates, blokb_set_peers / blokb_bind_image — run on the context's own stream, which first waits (cudaStreamWaitEvent)
 * for every dispatch this context enqueued on any caller stream; so "dispatch on my stream, then blokb_readback" returns that
 * dispatch's output without a synchronisation by the caller, and blokb_sync waits for caller-stream dispatches too. The
 * reverse direction needs nothing: those calls have completed when they return.
 * Dispatches that share a context's scratch — blokb_render / blokb_trace_path in wavefront mode (one slot-array scratch per context) and
 * blokb_trace_rays_reordered (one sort scratch) — are serialised by the library when they arrive on different streams: the later one
 * waits (cudaStreamWaitEvent) for the earlier one's last kernel. They never run concurrently; use two contexts for that.
 * There is no CPU fallback: without a CUDA device every compute entry point fails with BLOKB_E_CUDA.
 */
#ifndef BLOKB200_H
#define BLOKB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BLOKB_E_ARG (-1)         /* bad argument */
#define BLOKB_E_CUDA (-2)        /* CUDA runtime error / no device */
#define BLOKB_E_UNSUPPORTED (-3) /* world layout outside what the reference's builder produces */
#define BLOKB_E_STATE (-4)       /* call order (e.g. trace before upload / set_frame) */

/* blok::SvoNode — blok/include/svo.hpp:23-28 (mirrored in assets/shaders/intersect.rint:11-16). */
typedef struct blokb_svo_node {
    uint32_t childMask;  /* bits 0-7: child i present; i = x | y<<1 | z<<2 */
    uint32_t firstChild; /* chunk-relative index of child 0; 0xFFFFFFFF = none */
    uint32_t materialId;
    float occupancy; /* > 0 => filled leaf */
} blokb_svo_node;

/* blok::SubChunkGpu — blok/include/resources.hpp:170-184 (intersect.rint:18-27). One per non-empty 16^3 region;
 * its array index is the reference's gl_PrimitiveID. */
typedef struct blokb_subchunk {
    uint32_t nodeOffset, rootNodeIndex, nodeCount, startDepth;
    float worldMin[3];
    float subChunkSize;
    float worldMax[3];
    float pad0;
} blokb_subchunk;

/* blok::MaterialGpu — blok/include/material.hpp:88-114 (hit.rchit:28-33). */
typedef struct blokb_material {
    float albedo[3];
    uint32_t flags; /* metal8<<24 | rough8<<16 | type4<<12 | alpha4<<8 | spec8 */
    float emission[3];
    float ior;
} blokb_material;

/* blok::FrameUBO — blok/include/resources.hpp:103-150 (raygen.rgen:13-53), std140, 560 bytes, column-major. */
typedef struct blokb_frame_ubo {
    float view[16], proj[16], invView[16], invProj[16], prevView[16], prevProj[16], prevViewProj[16];
    float camPos[3];
    float deltaTime;
    float prevCamPos[3];
    uint32_t depth;
    uint32_t frameCount, sampleCount, screenWidth, screenHeight;
    float temporalAlpha, momentAlpha, varianceClipGamma, depthThreshold;
    float normalThreshold, phiColor, phiNormal, phiDepth;
    int32_t atrousIteration, stepSize;
    float varianceBoost;
    int32_t minHistoryLength;
    float jitterOffset[2];
    float _padding[2];
} blokb_frame_ubo;

/* blok::Camera — blok/include/camera.hpp:15-59. */
typedef struct blokb_camera {
    float position[3];
    float yaw, pitch, fov; /* degrees */
} blokb_camera;

/* Per-ray result. node = the reference's global SVO node index of the hit leaf (sub.nodeOffset + relative index,
 * intersect.rint:196), material = that node's materialId, t = the reported hit distance (item.tEntry),
 * facePrim = face (0..5: +X,-X,+Y,-Y,+Z,-Z; intersect.rint:58-68) | gl_PrimitiveID << 3.
 * Miss: node = material = facePrim = 0xFFFFFFFF, t = -1. */
typedef struct blokb_hit {
    uint32_t node;
    uint32_t material;
    float t;
    uint32_t facePrim;
} blokb_hit;

/* Packed per-ray result for bandwidth-bound consumers (PCIe readback, NVLink exchange of a tiled frame): 8 bytes instead of 16.
 * nodeFace = node | face << 29 (node and face as in blokb_hit); a miss is nodeFace = 0xFFFFFFFF, t = -1. material and
 * gl_PrimitiveID are functions of the leaf node and are re-derived on the host from the arrays the caller uploaded
 * (blokb_node_prims + blokb_unpack_hits). Worlds with more than BLOKB_PACKED_MAX_NODES nodes are refused. */
typedef struct blokb_hit_packed {
    uint32_t nodeFace;
    float t;
} blokb_hit_packed;
#define BLOKB_PACKED_MAX_NODES 0x1FFFFFFFu

/* Device pointers owned by the context (valid until the next set_frame with a different size / destroy). */
typedef struct blokb_buffers {
    blokb_hit* hit;        /* w*h       primary-hit records (trace_primary / trace_primary_shadow) */
    uint8_t* shadow;       /* w*h       0 = no shadow ray cast, 1 = lit, 2 = sun-shadowed (trace_primary_shadow) */
    float* color;          /* w*h*4 f32 outColor          (binding 4, rgba32f) */
    float* worldPos;       /* w*h*4 f32 outWorldPosition  (binding 5, rgba32f) */
    uint16_t* normalRough; /* w*h*4 f16 outNormalRoughness(binding 6, rgba16f) */
    uint8_t* albedoMetal;  /* w*h*4 u8  outAlbedoMetallic (binding 7, rgba8 unorm) */
    uint16_t* motion;      /* w*h*2 f16 outMotionVectors  (binding 8, rg16f) */
    float* accum;          /* w*h*4 f32 (sum.rgb, spp)    (cuda_tracer.cu:372-380) */
    uint32_t width, height;
    blokb_hit_packed* hitPacked; /* w*h   packed primary-hit records, written by blokb_trace_primary_packed */
} blokb_buffers;

/* Denoiser::Settings::atrousIterations (blok/include/renderer_denoising.hpp:71) + PostProcess::Settings
 * (blok/include/renderer_postprocess.hpp:102-119). The remaining Denoiser settings travel in the FrameUBO, exactly as
 * Denoiser::fillFrameUBO puts them there (renderer_denoising.cpp:686-703). */
typedef struct blokb_post_settings {
    int32_t atrousIterations; /* 1..5 (DenoiserPipeline::MAX_ATROUS_ITERATIONS) */
    int32_t enableTAA;
    float feedbackMin, feedbackMax;
    int32_t enableTonemapping;
    float exposure, saturationBoost;
    int32_t tonemapOperator; /* 0 = soft clip, 1 = Khronos PBR neutral */
    float whitePoint;
    int32_t enableSharpening;
    float sharpenStrength;
} blokb_post_settings;

enum blokb_image {
    BLOKB_IMG_HIT = 0,
    BLOKB_IMG_SHADOW = 1,
    BLOKB_IMG_COLOR = 2,
    BLOKB_IMG_WORLDPOS = 3,
    BLOKB_IMG_NORMALROUGH = 4,
    BLOKB_IMG_ALBEDOMETAL = 5,
    BLOKB_IMG_MOTION = 6,
    BLOKB_IMG_ACCUM = 7,
    /* denoiser / post-process images (valid after blokb_denoise / blokb_postprocess of the current frame) */
    BLOKB_IMG_HISTORY_COLOR = 8,  /* currentHistory()        rgba32f  temporal accumulation output */
    BLOKB_IMG_MOMENTS = 9,        /* currentMoments()        rg32f */
    BLOKB_IMG_HISTORY_LENGTH = 10, /* currentHistoryLength() r16f */
    BLOKB_IMG_VARIANCE = 11,      /* gbuffer.variance        r32f */
    BLOKB_IMG_DENOISED = 12,      /* Denoiser::getOutputImage() rgba32f (last a-trous iteration) */
    BLOKB_IMG_TAA_HISTORY = 13,   /* PostProcess currentHistory() rgba32f */
    BLOKB_IMG_TAA_OUTPUT = 14,    /* taaOutput rgba32f — only materialised when tonemapping is disabled */
    BLOKB_IMG_TONEMAPPED = 15,    /* tonemapOutput rgba8 */
    BLOKB_IMG_FINAL = 16,         /* PostProcess::getOutputImage(): sharpenOutput / tonemapOutput rgba8 (or taaOutput rgba32f) */
    BLOKB_IMG_HIT_PACKED = 17     /* w*h packed 8-byte hit records, written by blokb_trace_primary_packed */
};

/* Sizes of the GPU-resident compact layout built by blokb_upload_world. */
typedef struct blokb_world_stats {
    uint64_t nodes, subChunks, materials; /* as uploaded */
    uint64_t bricks, blocks, voxels;      /* 4^3 bricks, non-empty 2^3 leaf blocks, filled voxels */
    uint64_t compactBytes;                /* bytes of the traversal structure (masks, bricks, tables) */
    int32_t gridOrigin[3], gridDim[3];    /* voxel coordinates of the top grid, extent in voxels */
    int32_t levels;                       /* 4^levels >= max extent */
    float buildMs, uploadMs;
} blokb_world_stats;

typedef struct blokb_ctx blokb_ctx;

const char* blokb_last_error(void);
int blokb_create(int device, blokb_ctx** out);
void blokb_destroy(blokb_ctx* ctx);

/* Scene in (Renderer::addWorld / updateWorld). Host pointers are borrowed for the call only. Builds the compact
 * 64-bit-mask hierarchy and copies it (and the materials) to the device. A second call replaces the world. */
int blokb_upload_world(blokb_ctx* ctx, const blokb_svo_node* nodes, size_t nNodes, const blokb_subchunk* subs, size_t nSubs,
                       const blokb_material* materials, size_t nMaterials);
int blokb_world_stats_get(blokb_ctx* ctx, blokb_world_stats* out);
/* Scene in, built on the GPU. Equivalent to ChunkManager(128, 1.0f)::setVoxelMaterial(p, id, 1.0f) for
 * every listed voxel (a voxel listed twice keeps the last id), rebuildDirtyChunks (blok/src/chunk_manager.cpp:106-140;
 * SvoTree::insertVoxel, svo.cpp:59-101), packChunksToGpuSvo (:234-314) and blokb_upload_world — but the voxel list goes
 * to the device as is and the compact hierarchy is built there; the reference's 16-byte node arrays are never
 * materialised, only their numbering (hit records carry the same global node indices as after blokb_upload_world of the
 * reference's arrays; chunks are ordered by (z, y, x) like blokb_worldgen_pack). xyz: n x int32[3] world voxel
 * coordinates (|coordinate| < 2^20), materialId: n x uint32; host pointers, borrowed for the call. */
int blokb_upload_voxels(blokb_ctx* ctx, const int32_t* xyz, const uint32_t* materialId, size_t n, const blokb_material* materials,
                        size_t nMaterials);
/* The same from the reference's dense representation: one entry per Chunk (blok/include/chunk.hpp:33-42), density and
 * materialIds being its two 128^3 arrays (index x + 128 y + 16384 z); a cell is filled iff density > 0
 * (buildSvoFromDensity, chunk_manager.cpp:106-119). Replaces rebuildDirtyChunks + packChunksToGpuSvo + addWorld for a
 * ChunkManager(128, 1.0f): pass every chunk of mgr.chunks. Chunks are ordered by (z, y, x); listing one twice is an error. */
typedef struct blokb_dense_chunk {
    int32_t cx, cy, cz; /* Chunk::cx, cy, cz */
    const float* density;
    const uint32_t* materialIds;
} blokb_dense_chunk;
int blokb_upload_dense(blokb_ctx* ctx, const blokb_dense_chunk* chunks, size_t nChunks, const blokb_material* materials, size_t nMaterials);
/* Editing (applyBrush marks chunks dirty, brush.cpp:13-63; rebuildDirtyChunks rebuilds them and the reference then re-packs and
 * re-uploads the WHOLE world, chunk_manager.cpp:106-140, renderer.hpp:55-71): after blokb_upload_dense the voxel list stays on the
 * device, and this call takes only the dirty chunks — their voxels are replaced in the list (chunks not seen before are added,
 * chunks that became empty keep their lone root node) and the world is rebuilt on the GPU. The result is byte-identical to
 * blokb_upload_dense of all chunks; the host sends 16 MB per dirty chunk instead of the world. Materials stay as uploaded.
 * BLOKB_E_STATE unless the current world came from blokb_upload_dense. */
int blokb_update_dense(blokb_ctx* ctx, const blokb_dense_chunk* chunks, size_t nChunks);
/* MagicaVoxel .vox import: loadVoxFile + importVoxMaterials + importVoxToChunks of the reference
 * (blok/src/vox_loader.cpp:151-430) restated as "file -> voxel list + material table", the inputs of blokb_upload_voxels.
 * Material id i = palette entry i (1..255) and id 0 = the library's default grey, as in a fresh MaterialLibrary
 * (material.cpp:19-29; MaterialGpu::pack, material.hpp:95-111); world voxel = floor(offset + (x, z, y)) (VOX z is up).
 * Host only. offset may be NULL (= 0,0,0). Errors: code < 0 and blokb_vox_last_error(). */
typedef struct blokb_vox blokb_vox;
int blokb_vox_load(const char* path, uint32_t modelIndex, const float* offset3, blokb_vox** out);
int blokb_vox_parse(const void* fileBytes, size_t n, uint32_t modelIndex, const float* offset3, blokb_vox** out);
void blokb_vox_counts(const blokb_vox* v, uint64_t* nVoxels, uint64_t* nMaterials, uint32_t* nModels, uint32_t* modelSize3);
const int32_t* blokb_vox_xyz(const blokb_vox* v);               /* nVoxels x int32[3] */
const uint32_t* blokb_vox_material_ids(const blokb_vox* v);     /* nVoxels */
const blokb_material* blokb_vox_materials(const blokb_vox* v);  /* 256 entries */
void blokb_vox_free(blokb_vox* v);
const char* blokb_vox_last_error(void);
/* loadAndImportVox + rebuildDirtyChunks + packChunksToGpuSvo + addWorld (blok/src/app.cpp:100-126) in one call: the file's voxels
 * go straight to the device builder. */
int blokb_upload_vox(blokb_ctx* ctx, const char* path, uint32_t modelIndex, const float* offset3);
/* Device-resident world arrays -> host (parity checks, tools). host == NULL: only *bytesOut is set. */
enum blokb_world_array {
    BLOKB_WORLD_NODES = 0,          /* 16-byte {u64 mask, u32 base, u32 aux} records of the compact 64-tree */
    BLOKB_WORLD_BLOCKFIRSTLEAF = 1, /* u32 reference node index of leaf 0 of every non-empty 2^3 block */
    BLOKB_WORLD_VOXELMAT = 2,       /* u16 (u32 if any id > 65535) material id per filled voxel */
    BLOKB_WORLD_SUBCHUNKS = 3       /* blokb_subchunk records (worlds built by blokb_upload_voxels) */
};
/* Scene dump: the GPU-resident compact world — nodes, blockFirstLeaf, voxelMat, materials and, for
 * device-built worlds, the SubChunkGpu records — written to / read from one file (64-bit FNV-1a checked), so an asset built
 * once (from the reference's arrays, a voxel list, dense chunks or a .vox file) loads without any builder: file -> device.
 * The reference has no scene file of its own (it re-imports .vox files at start-up, app.cpp:100-126). A loaded world traces
 * exactly like the one that was saved (same hit records). Bad files are BLOKB_E_ARG with the reason in blokb_last_error(). */
int blokb_world_save(blokb_ctx* ctx, const char* path);
int blokb_world_load(blokb_ctx* ctx, const char* path);
int blokb_world_export(blokb_ctx* ctx, int which, void* host, size_t bytes, size_t* bytesOut);

/* Frame in (binding 3). (Re)allocates the output images when the size changes. */
int blokb_set_frame(blokb_ctx* ctx, const blokb_frame_ubo* frame);
/* Host helper: Camera::view()/projection() + fillFrameUBO for a first frame without TAA jitter
 * (camera.hpp:42-59; renderer_draw.cpp:52-89; renderer_denoising.cpp:655-704). Pure host code. */
int blokb_camera_frame(const blokb_camera* cam, uint32_t width, uint32_t height, uint32_t frameCount, uint32_t sampleCount,
                       blokb_frame_ubo* out);

/* The same for frame n of a sequence: prevCam = the previous frame's camera (NULL on the first frame) fills
 * prevView / prevProj / prevViewProj / prevCamPos as Denoiser::updatePreviousFrameData left them (non-jittered,
 * renderer_draw.cpp:314-318); jitterIndex >= 0 applies PostProcess::getJitteredProjection with Halton(2,3) sample
 * jitterIndex % 16 (renderer_postprocess.cpp:208-268), < 0 = TAA jitter off. */
int blokb_camera_frame_seq(const blokb_camera* cam, const blokb_camera* prevCam, uint32_t width, uint32_t height, uint32_t frameCount,
                           uint32_t sampleCount, int32_t jitterIndex, blokb_frame_ubo* out);

/* Tile scheduler. The reference dispatches the whole image in one traceRaysKHR(width, height, 1)
 * (renderer_raytracing.cpp:666-685); to split ONE frame over launches or GPUs the frame is cut into 8-row strips
 * (strip s = rows [8s, 8s + 8)) and the dispatch calls below render strips first, first + stride, first + 2*stride, ...
 * only. Pixel coordinates, RNG seeds (raygen.rgen:92-99) and output addresses stay those of the full frame, so the union
 * of the strip sets {(g, G) : g < G} writes exactly the bytes of the undivided dispatch; interleaving (stride = G)
 * balances sky and ground rows across GPUs. Default (0, 1) = the whole frame. stride == 0 is BLOKB_E_ARG. */
int blokb_set_strips(blokb_ctx* ctx, uint32_t first, uint32_t stride);

/* Fused exchange for a frame tiled over the GPUs of one box (one process per GPU). Each rank exports its hit / shadow /
 * accumulation image as a CUDA IPC handle (64 bytes), the ranks swap handles by any means (torch.distributed
 * all_gather in the Python layer), and blokb_set_peers(image, handles[n], n, self) maps the peers' images. From then on
 * every dispatch that writes that image stores each record into ALL n images — its own and, over NVLink, the peers' —
 * at the full-frame address, so when every rank's kernel has finished (one barrier) every rank holds the complete image:
 * the all-gather of the strips is fused into the traversal kernel instead of following it (the one
 * exchange step of the pixel-partitioned path). n <= 8; n == 0 unmaps. Images must keep their size while mapped
 * (blokb_set_frame with another resolution unmaps this side). The reference has no multi-GPU path. */
/* The same exchange through NVLink multicast (NVLS): the image is placed in caller-owned symmetric memory — `local` is this
 * GPU's address of it, `multicast` the address of the multicast object all ranks' copies are bound to (both e.g. from
 * torch.distributed._symmetric_memory: tensor.data_ptr() and rendezvous(...).multicast_ptr). The context then uses
 * `local` as the image (readback, get_buffers, resolve ...) and its kernels store each record ONCE, with multimem.st to
 * `multicast`; the NVSwitch replicates it into every GPU's copy, so a rank's NVLink egress is its own strips, not
 * (G-1) copies of them. multicast == NULL just substitutes the memory; local == NULL restores the context's own
 * allocation. The memory must hold the whole image and outlive the binding. */
/* Exchangeable images: BLOKB_IMG_HIT, _SHADOW, _ACCUM, _HIT_PACKED (stores of the dispatches) and _COLOR (the resolved slices of
 * blokb_reduce_accum). */
int blokb_bind_image(blokb_ctx* ctx, int image, void* local, void* multicast);
/* The accumulation image in symmetric memory for SAMPLE-sharded frames: every rank accumulates all pixels, so dispatches keep storing
 * into the LOCAL copy only, and the multicast address serves blokb_reduce_accum's multimem.ld_reduce (with blokb_bind_image the stores
 * of a dispatch would be replicated into every rank's copy — right for a tiled frame whose pixels have one owner, wrong here). */
int blokb_bind_accum_for_reduce(blokb_ctx* ctx, void* local, void* multicast);
int blokb_ipc_export(blokb_ctx* ctx, int image, void* handle64);
/* Frame fence of a tiled frame without a collective library: `local` / `multicast` name a 16-byte, zero-filled word pair in symmetric
 * memory bound to an NVLink multicast object (like the images above), shared by the `nranks` contexts of the frame. blokb_frame_fence
 * enqueues a one-thread kernel that adds 1 to the word on EVERY GPU with one multimem.red.release and then waits (ld.acquire, bounded)
 * until its own copy has counted all ranks of this round: when it completes on a rank, every rank's earlier work on its fence stream
 * — the multicast stores of its traversal kernel — has been performed. Every rank must enqueue the same number of fences. A fence that
 * times out (a rank missing for ~1 s) sets the second word to 1; from then on blokb_fence_status returns BLOKB_E_STATE, fence kernels
 * already enqueued pass without waiting, and every blokb_frame_fence called after the timed-out kernel has finished returns
 * BLOKB_E_STATE without enqueuing anything — until the fence is bound again (all ranks, fresh zero-filled words), which restarts the
 * count. nranks == 0 unbinds. */
int blokb_bind_fence(blokb_ctx* ctx, void* local, void* multicast, uint32_t nranks);
int blokb_frame_fence(blokb_ctx* ctx, void* stream);
int blokb_fence_status(blokb_ctx* ctx);
int blokb_set_peers(blokb_ctx* ctx, int image, const void* handles, uint32_t n, uint32_t self);

/* Cross-GPU reduction of the per-GPU accumulation buffers — the one exchange step of sample-sharded path tracing (SURVEY.md
 * 8(e); BASELINE configs[3]): GPU g of G traced samples [spp*g/G, spp*(g+1)/G) of every pixel into its own accum image, and
 * colour = (sum over GPUs of sum.rgb) / (sum over GPUs of spp) with the firefly clamp (raygen.rgen:383-392). The reference has no
 * multi-GPU path; behind Renderer::render (blok/include/renderer.hpp:34) a C++ host calls, per frame and per GPU:
 *   blokb_set_frame, blokb_reset_accum, blokb_trace_path(s0, s1), blokb_reduce_accum(mode), then reads the colour image.
 * The communicator is NCCL's, reached through dlopen("libnccl.so.2") — no link-time dependency, and a process that already
 * carries an NCCL (its own, or torch's) gets that one:
 *   blokb_set_comm            adopts a caller-owned ncclComm_t (one rank per context / GPU); never destroyed by the library.
 *   blokb_comm_unique_id /    for hosts without an NCCL of their own: rank 0 obtains a 128-byte ncclUniqueId, ships it to the
 *   blokb_comm_init           other ranks by any means, and every rank creates its (library-owned) communicator with it.
 *   blokb_reduce_accum        enqueue only. BLOKB_REDUCE_SLICE: reduce-scatter of accum (fp32 sum) into the slice of pixels this
 *                             rank owns, resolved into that slice of the colour image — each rank ends with 1/G of the frame
 *                             (blokb_reduce_slice tells which pixels; blokb_readback_slice copies just those).
 *                             BLOKB_REDUCE_ALL: the same, then an all-gather of the resolved slices: every rank holds the frame.
 *                             When accum is bound to an NVLink multicast address (blokb_bind_image) the sum and the resolve are
 *                             ONE kernel over peer memory (multimem.ld_reduce: the NVSwitch adds the G copies of a texel on
 *                             the way), fenced by two 4-byte all-reduces; with BLOKB_IMG_COLOR bound too, BLOKB_REDUCE_ALL
 *                             broadcasts each resolved texel with one multicast store instead of the all-gather.
 * Pixels are sliced in row-major order: rank g owns [g*ceil(n/G), min(n, (g+1)*ceil(n/G))). accum keeps the LOCAL sums (a later
 * blokb_trace_path keeps accumulating); only colour holds the reduced result. */
#define BLOKB_REDUCE_ALL 0
#define BLOKB_REDUCE_SLICE 1
int blokb_comm_unique_id(void* id128);
int blokb_comm_init(blokb_ctx* ctx, const void* id128, int rank, int nranks);
int blokb_set_comm(blokb_ctx* ctx, void* ncclComm);
int blokb_comm_rank(blokb_ctx* ctx, int* rank, int* nranks);
int blokb_reduce_accum(blokb_ctx* ctx, int mode, void* stream);
int blokb_reduce_slice(blokb_ctx* ctx, uint64_t* firstPixel, uint64_t* nPixels);
int blokb_readback_slice(blokb_ctx* ctx, int image, void* host, size_t bytes); /* image's pixels [first, first + n) -> host (n * bytes-per-pixel) */
/* Only the strips this context renders (blokb_set_strips) of an image, copied to their full-frame positions in `host` (a buffer
 * of the whole image; other rows are left untouched): what one GPU of a tiled frame hands to the host. */
int blokb_readback_strips(blokb_ctx* ctx, int image, void* host, size_t bytes);

/* Dispatch (enqueue only). */
int blokb_trace_primary(blokb_ctx* ctx, void* stream);        /* sample-0 primary rays -> buffers.hit */
int blokb_trace_primary_shadow(blokb_ctx* ctx, void* stream); /* + sun shadow ray per lit hit -> buffers.shadow */
int blokb_trace_primary_packed(blokb_ctx* ctx, void* stream); /* sample-0 primary rays -> buffers.hitPacked (8-byte records) */
int blokb_render(blokb_ctx* ctx, void* stream);               /* full raygen.rgen: frame.sampleCount spp -> 5 images */
int blokb_trace_path(blokb_ctx* ctx, uint32_t sampleBegin, uint32_t sampleEnd, void* stream); /* accum += samples */
int blokb_resolve(blokb_ctx* ctx, void* stream);              /* color = accum.rgb / accum.a, firefly clamp */
int blokb_reset_accum(blokb_ctx* ctx, void* stream);          /* CudaTracer::resetAccum */
/* How blokb_render / blokb_trace_path schedule the rays of a sample (the reference leaves this to the driver's ray-tracing
 * pipeline, renderer_raytracing.cpp:666-685). Either way the images are those of raygen.rgen:188-392, bit for bit the same.
 *   BLOKB_PATH_WAVEFRONT (default)  batches of `batchSamples` samples of every pixel: camera + sun-shadow rays per pixel tile,
 *                                   then the bounce rays of the whole batch sorted by a key — direction cell (dirCells^3) above
 *                                   the Morton code (mortonBitsPerAxis bits per axis) of the origin's 2^originCellShift-voxel
 *                                   cell — and traced in that order, so a warp's rays walk the same nodes.
 *                                   Needs scratch memory: 80 bytes x pixels x batchSamples, allocated on first use, grow-only;
 *                                   the batch is halved if the allocation fails.
 *   BLOKB_PATH_FUSED                one thread walks camera, shadow and bounce ray of a sample back to back; no scratch.
 * Zero fields mean "default". Takes effect from the next dispatch. */
#define BLOKB_PATH_WAVEFRONT 0
#define BLOKB_PATH_FUSED 1
typedef struct blokb_path_options {
    int32_t mode;
    uint32_t batchSamples;      /* default 8 */
    uint32_t dirCells;          /* 1..16, default 8 */
    int32_t originCellShift;    /* default 6 (64-voxel cells); raised as needed to cover the grid */
    uint32_t mortonBitsPerAxis; /* 1..6, default 2: 9 + 6 = 15 key bits = two radix passes (finer origin cells measured no faster) */
    uint32_t refillBelow;       /* bounce pass: 0 = one sorted ray per thread; k in 1..32 = persistent warps whose idle lanes fetch the next
                                   rays of the sorted list whenever fewer than k lanes are still walking. Default: BLOKB_REFILL_DEFAULT */
} blokb_path_options;
#define BLOKB_REFILL_KEEP 0xFFFFFFFFu /* in blokb_set_path_options: leave refillBelow as it is (0 is a value here, not "default") */
void blokb_path_options_default(blokb_path_options* out);
int blokb_set_path_options(blokb_ctx* ctx, const blokb_path_options* opt);
int blokb_get_path_options(blokb_ctx* ctx, blokb_path_options* out);

/* Denoiser + post-process, enqueue only. Frame order as in Renderer::render
 * (blok/src/renderer_draw.cpp:141-213, 314-328): blokb_set_frame, blokb_render, blokb_denoise, blokb_postprocess,
 * blokb_end_frame. History images start zero-filled; frameCount == 0 in the FrameUBO disables both reprojections.
 *   blokb_denoise      Denoiser::denoise (renderer_denoising.cpp:714-769): temporal accumulation + variance (one fused
 *                      kernel) and settings->atrousIterations a-trous passes -> BLOKB_IMG_DENOISED
 *   blokb_postprocess  PostProcess::process (renderer_postprocess.cpp:505-558): TAA + tonemap (one fused kernel) and
 *                      sharpen -> BLOKB_IMG_FINAL
 *   blokb_end_frame    copyCurrentGeometryToHistory + both swapHistoryBuffers (renderer_denoising.cpp:833-865): the
 *                      G-buffer's worldPos / normalRough images are double-buffered and the roles flip (no copy), so
 *                      blokb_get_buffers must be called again after it
 *   blokb_reset_history zero-fills every history image and restarts the ping-pong */
void blokb_post_settings_default(blokb_post_settings* out);
int blokb_denoise(blokb_ctx* ctx, const blokb_post_settings* settings, void* stream);
int blokb_postprocess(blokb_ctx* ctx, const blokb_post_settings* settings, void* stream);
int blokb_end_frame(blokb_ctx* ctx, void* stream);
int blokb_reset_history(blokb_ctx* ctx, void* stream);
/* Arbitrary rays (device pointers): rays = n x {ox,oy,oz,dx,dy,dz,tmin,tmax}; anyHit != 0 = shadow-ray rules. */
int blokb_trace_rays(blokb_ctx* ctx, const float* dRays, size_t n, int anyHit, blokb_hit* dOut, void* stream);
/* The same result for INCOHERENT lists (bounce rays): the rays are sorted on the device by an 8^3 direction cell above the
 * Morton code of their origin's coarse cell, traced in that order (a warp's rays then walk the same nodes) and every record
 * is written to its ray's own slot — dOut is byte-identical to blokb_trace_rays'. Costs one key pass + one radix sort;
 * pays off when neighbouring list entries point different ways. One list per context at a time. */
int blokb_trace_rays_reordered(blokb_ctx* ctx, const float* dRays, size_t n, int anyHit, blokb_hit* dOut, void* stream);
/* No counterpart in the reference (it never allocates on its traversal path: the driver's ray-tracing pipeline owns all scratch).
 * blokb_trace_rays_reordered once more, for hosts that must not allocate device memory when they call it. Two kinds of host are meant:
 * an engine that sizes its device memory once at start-up and hands the rest of the GPU to its own pool, so that a cudaMalloc
 * arriving later from inside this library either fails or breaks the engine's budget; and a frame loop with a latency bound, because
 * blokb_trace_rays_reordered grows its scratch with cudaDeviceSynchronize + cudaFree + cudaMalloc whenever a list is longer than any
 * before it, and those calls stall every stream of the process, not only this context's.
 * This entry point calls no cudaMalloc / cudaFree and no library routine that could (the ordering is a counting sort of the library's
 * own, not cub's radix sort; on a caller stream it does not register the stream with the context either, so blokb_sync does not wait
 * for it — order later work on that stream yourself). Its scratch — bin counters, keys and the traversal order, 16 MiB + 128 KiB — is a
 * fixed array inside the library's device image: every GPU that loads the library carries it from then until the process exits,
 * whether or not this function is ever called. Hence the limit: n <= BLOKB_STATIC_MAX_RAYS, more is BLOKB_E_ARG (split the list;
 * records are per ray, so the pieces can go one after another). dOut is byte-identical to blokb_trace_rays' and
 * blokb_trace_rays_reordered's; only the order in which rays are walked differs (15-bit key: 8^3 direction cells above 4^3 origin cells).
 * The array is ONE per GPU, shared by all contexts and all calls: two calls that overlap on the device (different streams, or different
 * contexts on the same GPU) race on it and return garbage — keep them on one stream or order them with events. Nothing needs clearing
 * by the caller: the function zeroes the counters it needs at the start of every call.
 * The first call on each GPU looks up the array's device address (cudaGetSymbolAddress, with that GPU current; it is remembered); a
 * failure there is BLOKB_E_CUDA. Make that first call during set-up, before any stream capture is begun. */
#define BLOKB_STATIC_MAX_RAYS (1u << 21)
int blokb_trace_rays_reordered_static(blokb_ctx* ctx, const float* dRays, size_t n, int anyHit, blokb_hit* dOut, void* stream);
/* Same with HOST buffers: copies in, traces, copies out, synchronises. */
int blokb_trace_rays_host(blokb_ctx* ctx, const float* rays, size_t n, int anyHit, blokb_hit* out);

/* Images out. */
int blokb_get_buffers(blokb_ctx* ctx, blokb_buffers* out);
int blokb_readback(blokb_ctx* ctx, int image, void* host, size_t bytes); /* copy + sync, ordered after every dispatch of this context (any stream) */
/* Inverse of blokb_readback for the G-buffer images (COLOR, WORLDPOS, NORMALROUGH, ALBEDOMETAL, MOTION, ACCUM): lets a
 * caller that produced the G-buffer elsewhere (or a test) feed blokb_denoise / blokb_postprocess. Copy + sync. */
int blokb_write_image(blokb_ctx* ctx, int image, const void* host, size_t bytes);
int blokb_sync(blokb_ctx* ctx);
/* Number of kernel launches this context has enqueued so far. */
uint64_t blokb_launch_count(blokb_ctx* ctx);
/* Profiling aid: while enabled, kernels add {DDA steps, rays traced by render/path kernels} to two device
 * counters (slower instrumented kernel variants); blokb_stats_read copies them out ([0] steps, [1] rays). */
int blokb_stats_enable(blokb_ctx* ctx, int on);
int blokb_stats_read(blokb_ctx* ctx, uint64_t* stepsAndRays);
/* The context's own CUDA stream (cudaStream_t). */
void* blokb_stream(blokb_ctx* ctx);

/* Frames in flight (the reference keeps 2: blok/include/renderer.hpp:185). A pipeline owns `depth` device hit images, a
 * copy stream and events: submit() enqueues the primary pass of one frame on the context's stream and the device->host
 * copy of its hit records into `hostOut` (w*h blokb_hit; pinned memory from blokb_host_alloc for full PCIe speed) on the
 * copy stream, so the copy of frame i overlaps the tracing of frame i+1. Nothing blocks the host until wait()/drain().
 * Tickets count submissions from 0; wait(t) is valid for the last `depth` tickets. */
typedef struct blokb_pipeline blokb_pipeline;
int blokb_pipeline_create(blokb_ctx* ctx, int depth, uint32_t width, uint32_t height, blokb_pipeline** out);
int blokb_pipeline_submit(blokb_pipeline* p, const blokb_frame_ubo* frame, blokb_hit* hostOut, uint64_t* ticket);
int blokb_pipeline_wait(blokb_pipeline* p, uint64_t ticket);
int blokb_pipeline_drain(blokb_pipeline* p);
void blokb_pipeline_destroy(blokb_pipeline* p);
/* The same pipeline with blokb_hit_packed records (8 bytes per ray: the readback is PCIe-bound). */
int blokb_pipeline_create_packed(blokb_ctx* ctx, int depth, uint32_t width, uint32_t height, blokb_pipeline** out);
int blokb_pipeline_submit_packed(blokb_pipeline* p, const blokb_frame_ubo* frame, blokb_hit_packed* hostOut, uint64_t* ticket);
/* Host helpers (no GPU involved). blokb_node_prims: primOfNode[i] = index of the sub-chunk whose subtree holds node i
 * (the walk of intersect.rint:114-196: root = nodeOffset + rootNodeIndex, child = nodeOffset + firstChild + octant),
 * 0xFFFFFFFF for nodes no sub-chunk reaches. blokb_unpack_hits: packed -> blokb_hit; primOfNode may be NULL, in which
 * case facePrim carries the face only (prim bits zero). Returns BLOKB_E_ARG on a record whose node is out of range. */
int blokb_node_prims(const blokb_svo_node* nodes, size_t nNodes, const blokb_subchunk* subs, size_t nSubs, uint32_t* primOfNode);
int blokb_unpack_hits(const blokb_hit_packed* in, size_t n, const blokb_svo_node* nodes, size_t nNodes, const uint32_t* primOfNode,
                      blokb_hit* out);
/* Page-locked host memory (cudaHostAlloc / cudaFreeHost). */
void* blokb_host_alloc(size_t bytes);
void blokb_host_free(void* p);

/* Host helper (no GPU involved): voxel list -> the reference's upload arrays, restating ChunkManager(128, 1.0f) +
 * rebuildDirtyChunks + packChunksToGpuSvo (chunk_manager.cpp:106-140, 234-314; svo.cpp:59-101). Chunks are emitted
 * sorted by (z, y, x); the reference's order is that of a std::unordered_map (unspecified). */
void* blokb_worldgen_pack(const int32_t* xyz, const uint32_t* materialId, size_t n);
void blokb_worldgen_counts(void* world, uint64_t* nNodes, uint64_t* nSubs);
void blokb_worldgen_copy(void* world, blokb_svo_node* nodes, blokb_subchunk* subs);
const blokb_svo_node* blokb_worldgen_nodes(void* world);
const blokb_subchunk* blokb_worldgen_subs(void* world);
void blokb_worldgen_free(void* world);

#ifdef __cplusplus
}
#endif
#endif /* BLOKB200_H */
