// Host helper: voxel list -> the reference's upload arrays (WorldSvoGpu: globalNodes + globalSubChunks).
// Lets a caller (bench, tests, or an application without the reference's ChunkManager) produce worlds in the
// exact format blokb_upload_world consumes. Restates, in our own code:
//   ChunkManager::globalVoxelToChunk / globalVoxelToLocal   blok/src/chunk_manager.cpp:41-55 (floor division by C=128)
//   buildSvoFromDensity insertion order (x fastest, then y, z)  chunk_manager.cpp:106-119
//   SvoTree::insertVoxel / ensureChildren                       blok/src/svo.cpp:36-101 (children allocated 8 at a time,
//                                                               octant = x | y<<1 | z<<2, MSB-first by level: morton.hpp:55-58)
//   packChunksToGpuSvo + subChunkHasGeometry + findSubChunkRootNode   chunk_manager.cpp:144-314
// The reference iterates chunks in std::unordered_map order (chunk_manager.cpp:249), which is unspecified; this
// builder emits chunks sorted by (z, y, x). Single-chunk worlds are therefore bit-identical to the reference's
// arrays; multi-chunk worlds are identical up to that chunk permutation (tests/test_host.py::test_worldgen_* check both).
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/blokb200.h"

namespace {
constexpr int32_t kC = 128;       // ChunkManager(128, 1.0f), app.cpp:37
constexpr uint32_t kDepth = 7;    // 2^7 = 128
constexpr uint32_t kSubDiv = 8;   // SUB_CHUNK_DIVISIONS, chunk_manager.cpp:17
constexpr uint32_t kNone = 0xFFFFFFFFu;

struct Vox { int32_t cx, cy, cz; uint32_t key; uint32_t mat; uint32_t order; };  // key = local z<<14 | y<<7 | x

inline int32_t floordiv(int32_t v) { return v >= 0 ? v / kC : (v - kC + 1) / kC; }

struct World {
    std::vector<blokb_svo_node> nodes;
    std::vector<blokb_subchunk> subs;
};

inline blokb_svo_node empty_node() { return blokb_svo_node{0u, kNone, 0u, 0.0f}; }

// Build one chunk's node array by replaying insertVoxel over voxels sorted in the reference's loop order.
void build_chunk(const Vox* v, size_t n, std::vector<blokb_svo_node>& out) {
    out.clear();
    out.push_back(empty_node());
    for (size_t i = 0; i < n; i++) {
        const uint32_t x = v[i].key & 127u, y = (v[i].key >> 7) & 127u, z = (v[i].key >> 14) & 127u;
        uint32_t node = 0;
        for (uint32_t level = 0; level < kDepth; level++) {
            const uint32_t sh = kDepth - 1u - level;
            const uint32_t oct = ((x >> sh) & 1u) | (((y >> sh) & 1u) << 1) | (((z >> sh) & 1u) << 2);
            if (out[node].firstChild == kNone) {
                const uint32_t fc = (uint32_t)out.size();
                out.resize(out.size() + 8, empty_node());
                out[node].firstChild = fc;
            }
            out[node].childMask |= 1u << oct;  // same final masks as the reference's upward propagation (svo.cpp:95-100)
            node = out[node].firstChild + oct;
        }
        out[node].materialId = v[i].mat;
        out[node].occupancy = 1.0f;
    }
}

// chunk_manager.cpp:144-232: descend log2(8)=3 levels towards sub-chunk (sx,sy,sz).
bool sub_root(const std::vector<blokb_svo_node>& nodes, uint32_t sx, uint32_t sy, uint32_t sz, uint32_t& rootOut) {
    uint32_t node = 0;
    for (uint32_t level = 0; level < 3; level++) {
        const uint32_t cell = kSubDiv >> (level + 1);
        const uint32_t oct = ((sx / cell) & 1u) | (((sy / cell) & 1u) << 1) | (((sz / cell) & 1u) << 2);
        const blokb_svo_node& nd = nodes[node];
        if (!((nd.childMask >> oct) & 1u) || nd.firstChild == kNone) return false;
        node = nd.firstChild + oct;
        if (node >= nodes.size()) return false;
    }
    rootOut = node;
    return nodes[node].childMask != 0u || nodes[node].occupancy > 0.0f;
}
}  // namespace

extern "C" {

// Returns an opaque handle (NULL on allocation failure). xyz: n x int32[3] world voxel coordinates;
// materialId: n x u32. A voxel listed twice keeps the last material (like ChunkManager::setVoxelMaterial).
void* blokb_worldgen_pack(const int32_t* xyz, const uint32_t* materialId, size_t n) {
    World* w = new (std::nothrow) World;
    if (!w) return nullptr;
    std::vector<Vox> v(n);
    for (size_t i = 0; i < n; i++) {
        const int32_t x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
        const int32_t cx = floordiv(x), cy = floordiv(y), cz = floordiv(z);
        v[i] = Vox{cx, cy, cz, uint32_t(x - cx * kC) | (uint32_t(y - cy * kC) << 7) | (uint32_t(z - cz * kC) << 14), materialId[i], (uint32_t)i};
    }
    std::sort(v.begin(), v.end(), [](const Vox& a, const Vox& b) {
        if (a.cz != b.cz) return a.cz < b.cz;
        if (a.cy != b.cy) return a.cy < b.cy;
        if (a.cx != b.cx) return a.cx < b.cx;
        if (a.key != b.key) return a.key < b.key;
        return a.order < b.order;
    });
    // duplicates: keep the last write
    size_t m = 0;
    for (size_t i = 0; i < n; i++) {
        if (m && v[m - 1].cx == v[i].cx && v[m - 1].cy == v[i].cy && v[m - 1].cz == v[i].cz && v[m - 1].key == v[i].key) v[m - 1] = v[i];
        else v[m++] = v[i];
    }
    v.resize(m);
    std::vector<blokb_svo_node> chunk;
    uint32_t nodeOffset = 0;
    for (size_t b = 0; b < v.size();) {
        size_t e = b;
        while (e < v.size() && v[e].cx == v[b].cx && v[e].cy == v[b].cy && v[e].cz == v[b].cz) e++;
        build_chunk(&v[b], e - b, chunk);
        const float ox = float(v[b].cx * kC), oy = float(v[b].cy * kC), oz = float(v[b].cz * kC);
        const float ss = float(kC) / float(kSubDiv);
        for (uint32_t sz = 0; sz < kSubDiv; sz++)
            for (uint32_t sy = 0; sy < kSubDiv; sy++)
                for (uint32_t sx = 0; sx < kSubDiv; sx++) {
                    uint32_t root;
                    if (!sub_root(chunk, sx, sy, sz, root)) continue;
                    blokb_subchunk s{};
                    s.nodeOffset = nodeOffset; s.rootNodeIndex = root; s.nodeCount = (uint32_t)chunk.size(); s.startDepth = 3;
                    s.worldMin[0] = ox + float(sx) * ss; s.worldMin[1] = oy + float(sy) * ss; s.worldMin[2] = oz + float(sz) * ss;
                    s.subChunkSize = ss;
                    for (int a = 0; a < 3; a++) s.worldMax[a] = s.worldMin[a] + ss;
                    w->subs.push_back(s);
                }
        w->nodes.insert(w->nodes.end(), chunk.begin(), chunk.end());
        nodeOffset += (uint32_t)chunk.size();
        b = e;
    }
    return w;
}
void blokb_worldgen_counts(void* h, uint64_t* nNodes, uint64_t* nSubs) {
    World* w = (World*)h;
    *nNodes = w->nodes.size(); *nSubs = w->subs.size();
}
void blokb_worldgen_copy(void* h, blokb_svo_node* nodes, blokb_subchunk* subs) {
    World* w = (World*)h;
    if (!w->nodes.empty()) std::memcpy(nodes, w->nodes.data(), w->nodes.size() * sizeof(blokb_svo_node));
    if (!w->subs.empty()) std::memcpy(subs, w->subs.data(), w->subs.size() * sizeof(blokb_subchunk));
}
const blokb_svo_node* blokb_worldgen_nodes(void* h) { return ((World*)h)->nodes.data(); }
const blokb_subchunk* blokb_worldgen_subs(void* h) { return ((World*)h)->subs.data(); }
void blokb_worldgen_free(void* h) { delete (World*)h; }
}
