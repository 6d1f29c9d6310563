// Compact GPU-resident world layout ("64-tree": every level is a 4x4x4 cell with a 64-bit occupancy mask).
//
// Built at upload (scene_build.cpp) from the arrays the reference uploads (WorldSvoGpu, resources.hpp:195-210):
//   level 0  voxel                    1^3
//   level 1  brick                    4^3    Brick{voxelMask, blockBase, matBase}     (sparse, per sub-chunk)
//   level 2  sub-chunk                16^3   SubRec{brickMask, brickBase}             (sparse, index = gl_PrimitiveID)
//   level 3+ dense mask arrays        64^3, 256^3, ... up to the single root cell     (dense, tiny)
// plus topGrid[16^3 cell] -> sub-chunk index, blockFirstLeaf[] (reference global node index of leaf 0 of each
// non-empty 2^3 block, so hit records carry the reference's node indices bit-exactly) and voxelMat[].
// Bit order inside every mask: bit = x + 4*y + 16*z of the child cell.
#pragma once
#include <cstdint>
#include <vector_types.h>

namespace blokb {

constexpr int kMaxLevels = 12;  // 4^12 = 16.7M voxels per axis (float-exact plane limit)

struct SubRec {
    uint64_t brickMask;
    uint32_t brickBase;  // index of this sub-chunk's first brick (bricks are popcount-ordered)
    uint32_t pad;
};
static_assert(sizeof(SubRec) == 16, "SubRec");

struct Brick {
    uint64_t voxelMask;
    uint32_t blockBase;  // index into blockFirstLeaf of this brick's first non-empty 2^3 block
    uint32_t matBase;    // index into voxelMat of this brick's first filled voxel
};
static_assert(sizeof(Brick) == 16, "Brick");

// Everything the traversal kernels read; passed by value as a kernel parameter.
struct DeviceWorld {
    int32_t origin[3];  // world voxel coordinate of grid cell (0,0,0); multiple of 16
    int32_t dim[3];     // grid extent in voxels; multiple of 16
    int32_t levels;     // root cell = level `levels` (4^levels >= max dim); >= 3
    int32_t matWide;    // 0: voxelMat is u16, 1: u32
    const uint64_t* dense[kMaxLevels + 1];  // [k] valid for 3 <= k <= levels; dims = ceil(dim / 4^k)
    int32_t denseDim[kMaxLevels + 1][3];
    const uint32_t* topGrid;  // dim/16 cells; 0xFFFFFFFF = empty
    const SubRec* subs;
    const Brick* bricks;
    const uint32_t* blockFirstLeaf;
    const void* voxelMat;
    const float4* materials;  // 2 x float4 per MaterialGpu
    uint32_t nMaterials;
};

}  // namespace blokb
