// Host-side scene upload: reference upload arrays -> compact 64-tree (see scene.h).
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/blokb200.h"
#include "scene.h"

namespace blokb {

struct HostWorld {
    int32_t origin[3] = {0, 0, 0}, dim[3] = {0, 0, 0};
    int32_t levels = 3;
    bool matWide = false;
    std::vector<std::vector<uint64_t>> dense;  // [k], k = 0..levels (only 3..levels used)
    int32_t denseDim[kMaxLevels + 1][3] = {};
    std::vector<uint32_t> topGrid;
    std::vector<Node> nodes;  // the uploaded tree (scene.h); everything above except blockFirstLeaf / mats is an intermediate
    std::vector<SubRec> subs;
    std::vector<Brick> bricks;
    std::vector<uint32_t> blockFirstLeaf;
    std::vector<uint16_t> mat16;
    std::vector<uint32_t> mat32;
    uint64_t voxels = 0;
    size_t compactBytes() const;
};

// Returns 0 or a BLOKB_E_* code with `err` filled.
int build_host_world(const blokb_svo_node* nodes, size_t nNodes, const blokb_subchunk* subs, size_t nSubs, HostWorld& out,
                     std::string& err);

int build_top(const blokb_subchunk* subs, size_t nSubs, uint64_t nBricks, HostWorld& out, uint64_t& brickStart, std::string& err);

// Host half of the packed readback (unpack.cpp); return 0 or BLOKB_E_ARG.
int node_prims(const blokb_svo_node* nodes, size_t nNodes, const blokb_subchunk* subs, size_t nSubs, uint32_t* primOfNode);
int unpack_hits(const blokb_hit_packed* in, size_t n, const blokb_svo_node* nodes, size_t nNodes, const uint32_t* primOfNode, blokb_hit* out);

}  // namespace blokb
