// Host half of the packed readback (blokb_hit_packed): the two fields the 8-byte record leaves out are
// functions of the hit leaf, so the host re-derives them from the arrays it uploaded. No GPU involved.
#include <algorithm>
#include <cstring>
#include <thread>
#include <vector>

#include "scene_build.h"

namespace blokb {

// The walk of intersect.rint:114-196 (root = nodeOffset + rootNodeIndex, child = nodeOffset + firstChild + octant in
// u32 arithmetic, indices outside [nodeOffset, nodeOffset + nodeCount) skipped) from every sub-chunk root; a node two
// sub-chunks reach keeps the first (the reference's builder never shares nodes: chunk_manager.cpp:255-303).
int node_prims(const blokb_svo_node* nodes, size_t nNodes, const blokb_subchunk* subs, size_t nSubs, uint32_t* primOfNode) {
    if ((nNodes && (!nodes || !primOfNode)) || (nSubs && !subs)) return BLOKB_E_ARG;
    if (nNodes) std::memset(primOfNode, 0xFF, nNodes * sizeof(uint32_t));
    std::vector<uint64_t> todo;
    for (size_t s = 0; s < nSubs; s++) {
        const uint64_t off = subs[s].nodeOffset;
        const uint64_t lim = off + subs[s].nodeCount < nNodes ? off + subs[s].nodeCount : nNodes;
        const uint64_t root = off + subs[s].rootNodeIndex;
        if (root >= lim) continue;
        todo.assign(1, root);
        while (!todo.empty()) {
            const uint64_t i = todo.back();
            todo.pop_back();
            if (primOfNode[i] != 0xFFFFFFFFu) continue;
            primOfNode[i] = uint32_t(s);
            const blokb_svo_node& n = nodes[i];
            for (uint32_t o = 0; o < 8; o++) {
                if (!(n.childMask >> o & 1u)) continue;
                const uint64_t ch = uint32_t(uint32_t(off) + n.firstChild + o);
                if (ch >= off && ch < lim) todo.push_back(ch);
            }
        }
    }
    return 0;
}

// Records are independent: large batches are split over the host's cores (a 4K frame is 8.3 M random lookups into the node array).
static int unpack_range(const blokb_hit_packed* in, size_t begin, size_t end, const blokb_svo_node* nodes, size_t nNodes,
                        const uint32_t* primOfNode, blokb_hit* out) {
    int rc = 0;
    for (size_t i = begin; i < end; i++) {
        const uint32_t nf = in[i].nodeFace;
        if (nf == 0xFFFFFFFFu) {
            out[i] = blokb_hit{0xFFFFFFFFu, 0xFFFFFFFFu, -1.0f, 0xFFFFFFFFu};
            continue;
        }
        const uint32_t node = nf & BLOKB_PACKED_MAX_NODES, face = nf >> 29;
        if (node >= nNodes) { out[i] = blokb_hit{0xFFFFFFFFu, 0xFFFFFFFFu, -1.0f, 0xFFFFFFFFu}; rc = BLOKB_E_ARG; continue; }
        out[i] = blokb_hit{node, nodes[node].materialId, in[i].t, face | (primOfNode ? primOfNode[node] << 3 : 0u)};
    }
    return rc;
}

int unpack_hits(const blokb_hit_packed* in, size_t n, const blokb_svo_node* nodes, size_t nNodes, const uint32_t* primOfNode,
                blokb_hit* out) {
    if (n && (!in || !out || !nodes)) return BLOKB_E_ARG;
    const size_t kMinPerThread = size_t(1) << 16;
    size_t nt = std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), 32);
    nt = std::min(nt, (n + kMinPerThread - 1) / kMinPerThread);
    if (nt <= 1) return unpack_range(in, 0, n, nodes, nNodes, primOfNode, out);
    std::vector<int> rcs(nt, 0);
    std::vector<std::thread> pool;
    for (size_t k = 0; k < nt; k++)
        pool.emplace_back([&, k] { rcs[k] = unpack_range(in, n * k / nt, n * (k + 1) / nt, nodes, nNodes, primOfNode, out); });
    for (auto& t : pool) t.join();
    for (int rc : rcs)
        if (rc) return rc;
    return 0;
}
}  // namespace blokb
