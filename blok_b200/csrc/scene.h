// Compact GPU-resident world layout: a sparse "64-tree" — every node is a 4x4x4 cell with a 64-bit occupancy mask
// and popcount-indexed children, the same 16-byte record at every level so that descending is one code path.
//
// Built at upload (scene_build.cpp) from the arrays the reference uploads (WorldSvoGpu, resources.hpp:195-210):
//   level 0      voxel       1^3
//   level 1      brick       4^3    Node{voxelMask, base = index of its first filled voxel in voxelMat[],
//                                        aux  = index of its first non-empty 2^3 block in blockFirstLeaf[]}
//   level 2      sub-chunk   16^3   Node{brickMask, base = node index of its first brick, aux = gl_PrimitiveID}
//   level 3..L   64^3 ... 4^L       Node{childMask, base = node index of its first child, aux = 0};  nodes[0] = root
// Children of a node are contiguous, in mask-bit order (bit = x + 4*y + 16*z of the child cell), so
//   child(node, bit) = nodes[node.base + popcount(node.mask & ((1 << bit) - 1))].
// Nodes are stored breadth first (root, level L-1, ..., sub-chunks, bricks): the upper levels are a few KB and stay
// in L1, the whole array (7.7 MB for the 2048^3 terrain) in L2.
// blockFirstLeaf[] holds the reference's global node index of leaf 0 of each non-empty 2^3 block, so hit records
// carry the reference's node indices bit-exactly; voxelMat[] the material id of every filled voxel.
#pragma once
#include <cstdint>
#include <vector_types.h>

namespace blokb {

constexpr int kMaxLevels = 12;  // 4^12 = 16.7M voxels per axis (float-exact plane limit)

struct Node {
    uint64_t mask;
    uint32_t base;
    uint32_t aux;
};
static_assert(sizeof(Node) == 16, "Node");

// Host-side intermediates of the builder (not uploaded).
struct SubRec {
    uint64_t brickMask;
    uint32_t brickBase;  // index of this sub-chunk's first brick (bricks are popcount-ordered)
    uint32_t pad;
};
struct Brick {
    uint64_t voxelMask;
    uint32_t blockBase;  // index into blockFirstLeaf of this brick's first non-empty 2^3 block
    uint32_t matBase;    // index into voxelMat of this brick's first filled voxel
};

// Everything the traversal kernels read; passed by value as a kernel parameter.
struct DeviceWorld {
    int32_t origin[3];  // world voxel coordinate of grid cell (0,0,0); multiple of 16
    int32_t dim[3];     // grid extent in voxels; multiple of 16
    int32_t levels;     // root cell = level `levels` (4^levels >= max dim); >= 3
    int32_t matWide;    // 0: voxelMat is u16, 1: u32
    const uint4* nodes;  // Node records, nodes[0] = root
    const uint32_t* blockFirstLeaf;
    const void* voxelMat;
    const float4* materials;  // 2 x float4 per MaterialGpu
    uint32_t nMaterials;
    uint32_t nNodes;  // records in nodes[]
};

}  // namespace blokb
