// Device scene build (build.cu): voxel list -> brick nodes, blockFirstLeaf, voxelMat and the sub-chunk records on the GPU.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/blokb200.h"
#include "scene.h"

namespace blokb {
namespace build {

struct DeviceBuild {
    // device arrays, owned by the caller after a successful build (cudaFree)
    Node* bricks = nullptr;            // nBricks level-1 nodes, in sub-chunk order then brick-bit order
    uint32_t* blockFirstLeaf = nullptr;  // nBlocks reference node indices
    void* voxelMat = nullptr;          // nVox material ids (u16, or u32 when matWide)
    blokb_subchunk* dSubs = nullptr;   // nSubs SubChunkGpu records (reference format)
    uint64_t nBricks = 0, nBlocks = 0, nVox = 0, nSubs = 0, nChunks = 0;
    uint64_t refNodes = 0;             // nodes the reference's packed array would hold: sum over chunks of 1 + 8 * interior
    bool matWide = false;
    // host copies for the top of the tree (build_top)
    std::vector<blokb_subchunk> subs;
    std::vector<SubRec> recs;
    float sortMs = 0.f, buildMs = 0.f;
};

// xyz: n x int32[3] world voxel coordinates, mat: n x u32 (host pointers; a voxel listed twice keeps the last material,
// like ChunkManager::setVoxelMaterial). Returns 0 or BLOKB_E_*; on failure nothing is left allocated.
int from_voxels(const int32_t* xyz, const uint32_t* mat, size_t n, cudaStream_t st, DeviceBuild& out, std::string& err);
// The same from device-resident arrays (borrowed; the call synchronises the stream before returning).
// emptyChunkKeys (sorted, optional): chunks the reference would pack as a lone root node; they shift later chunks' node offsets.
int from_device_voxels(const int32_t* dXyz, const uint32_t* dMat, size_t n, cudaStream_t st, DeviceBuild& out, std::string& err,
                       const std::vector<uint64_t>* emptyChunkKeys = nullptr);
// The voxel list a world was built from, kept on the device so that dirty chunks can be replaced later
// (Renderer::updateWorld after ChunkManager::rebuildDirtyChunks) without the host resending the rest of the world.
struct VoxelList {
    int32_t* xyz = nullptr;            // n x int32[3], in write order (later entries win)
    uint32_t* mat = nullptr;
    size_t n = 0;
    std::vector<uint64_t> emptyKeys;   // sorted keys of chunks that exist but hold no filled cell
    void release();
};
// Replaces every voxel of the listed chunks by the chunks' cells with density > 0 (chunks new to L are added).
int replace_chunks(VoxelList& L, const blokb_dense_chunk* chunks, size_t nChunks, cudaStream_t st, std::string& err);
int from_list(const VoxelList& L, cudaStream_t st, DeviceBuild& out, std::string& err);
// Dense 128^3 chunks in the reference's Chunk layout (host pointers): cells with density > 0 are filled. keep != nullptr
// receives the voxel list (caller releases it).
int from_dense(const blokb_dense_chunk* chunks, size_t nChunks, cudaStream_t st, DeviceBuild& out, std::string& err, VoxelList* keep = nullptr);
void release(DeviceBuild& b);

}  // namespace build
}  // namespace blokb
