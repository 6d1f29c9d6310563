// Scene upload, host half: walks the reference's packed arrays exactly the way intersect.rint addresses them
// (root = nodeOffset + rootNodeIndex, child = nodeOffset + firstChild + octant, out-of-range indices skipped:
// intersect.rint:114, 132, 196) and emits the compact 64-tree of scene.h.
// What the reference's builder guarantees, and what is therefore required here (else BLOKB_E_UNSUPPORTED):
//   * sub-chunks are 16-unit boxes on the integer lattice with startDepth 3 (chunk_manager.cpp:17, 255-303)
//   * only depth-4 nodes below a sub-chunk root are filled leaves (svo.cpp:59-101)
#include "scene_build.h"

#include <algorithm>
#include <cmath>
#include <thread>

namespace blokb {

size_t HostWorld::compactBytes() const {
    return nodes.size() * sizeof(Node) + blockFirstLeaf.size() * 4 + mat16.size() * 2 + mat32.size() * 4;
}

namespace {
struct SubOut {
    uint64_t brickMask = 0;
    Brick bricks[64];            // indexed by brick bit, compacted later
    uint32_t firstLeaf[64][8];   // [brick][block]
    uint32_t mats[64][64];       // [brick][voxel bit]
    bool bad = false;
};

// Decode one sub-chunk (4 levels below its root) into 64 brick masks.
static void decode_sub(const blokb_svo_node* nodes, size_t nNodes, const blokb_subchunk& s, SubOut& o) {
    const uint64_t off = s.nodeOffset, lim = std::min<uint64_t>(uint64_t(s.nodeOffset) + s.nodeCount, nNodes);
    auto valid = [&](uint64_t i) { return i < lim; };
    const uint64_t ri = off + s.rootNodeIndex;
    if (!valid(ri)) return;
    const blokb_svo_node& r = nodes[ri];
    if (r.childMask == 0) { if (r.occupancy > 0.f) o.bad = true; return; }
    for (uint32_t o1 = 0; o1 < 8; o1++) {
        if (!(r.childMask >> o1 & 1)) continue;
        uint64_t i1 = uint32_t(off + r.firstChild + o1);  // u32 arithmetic, like the shader
        if (!valid(i1)) continue;
        const blokb_svo_node& n1 = nodes[i1];
        if (n1.childMask == 0) { if (n1.occupancy > 0.f) o.bad = true; continue; }
        for (uint32_t o2 = 0; o2 < 8; o2++) {
            if (!(n1.childMask >> o2 & 1)) continue;
            uint64_t i2 = uint32_t(off + n1.firstChild + o2);
            if (!valid(i2)) continue;
            const blokb_svo_node& n2 = nodes[i2];
            if (n2.childMask == 0) { if (n2.occupancy > 0.f) o.bad = true; continue; }
            const uint32_t bx = (o1 & 1) * 2 + (o2 & 1), by = (o1 >> 1 & 1) * 2 + (o2 >> 1 & 1), bz = (o1 >> 2 & 1) * 2 + (o2 >> 2 & 1);
            const uint32_t bb = bx + 4 * by + 16 * bz;
            uint64_t vm = 0;
            for (uint32_t o3 = 0; o3 < 8; o3++) {
                if (!(n2.childMask >> o3 & 1)) continue;
                uint64_t i3 = uint32_t(off + n2.firstChild + o3);
                if (!valid(i3)) continue;
                const blokb_svo_node& n3 = nodes[i3];
                if (n3.childMask == 0) { if (n3.occupancy > 0.f) o.bad = true; continue; }
                const uint32_t leaf0 = uint32_t(off + n3.firstChild);
                o.firstLeaf[bb][o3] = leaf0;
                for (uint32_t o4 = 0; o4 < 8; o4++) {
                    if (!(n3.childMask >> o4 & 1)) continue;
                    uint64_t i4 = uint32_t(leaf0 + o4);
                    if (!valid(i4)) continue;
                    const blokb_svo_node& lf = nodes[i4];
                    if (lf.childMask != 0) { o.bad = true; continue; }
                    if (!(lf.occupancy > 0.f)) continue;
                    const uint32_t vx = (o3 & 1) * 2 + (o4 & 1), vy = (o3 >> 1 & 1) * 2 + (o4 >> 1 & 1), vz = (o3 >> 2 & 1) * 2 + (o4 >> 2 & 1);
                    const uint32_t vb = vx + 4 * vy + 16 * vz;
                    vm |= 1ull << vb;
                    o.mats[bb][vb] = lf.materialId;
                }
            }
            if (vm) { o.brickMask |= 1ull << bb; o.bricks[bb].voxelMask = vm; }
        }
    }
}

// voxel-bit pattern of 2^3 block b (b = bx + 2*by + 4*bz) inside a brick mask
static inline uint64_t block_bits(uint32_t b) { return 0x0000000000330033ull << (2 * (b & 1) + 8 * (b >> 1 & 1) + 32 * (b >> 2 & 1)); }
}  // namespace

// Bounding box, level count, top grid, mask pyramid and the breadth-first upper part of the tree (root ... sub-chunk
// nodes) from the sub-chunk records alone: w.subs[i] = {brickMask, brickBase} must be filled. Brick nodes are NOT appended;
// they start at index `brickStart` (returned) in sub-chunk order. Shared by the host builder below and the device
// builder (build.cu), which produces the per-sub-chunk records and the brick nodes on the GPU.
int build_top(const blokb_subchunk* subs, size_t nSubs, uint64_t nBricks, HostWorld& w, uint64_t& brickStart, std::string& err) {
    if (nSubs == 0) {  // empty world: a single empty 16^3 grid
        w.origin[0] = w.origin[1] = w.origin[2] = 0;
        w.dim[0] = w.dim[1] = w.dim[2] = 16;
    } else {
        int64_t lo[3] = {INT64_MAX, INT64_MAX, INT64_MAX}, hi[3] = {INT64_MIN, INT64_MIN, INT64_MIN};
        for (size_t i = 0; i < nSubs; i++) {
            const blokb_subchunk& s = subs[i];
            if (s.subChunkSize != 16.0f || s.startDepth != 3u) { err = "sub-chunk " + std::to_string(i) + ": only 16-unit sub-chunks at depth 3 (ChunkManager(128, 1.0f)) are supported"; return BLOKB_E_UNSUPPORTED; }
            for (int a = 0; a < 3; a++) {
                float m = s.worldMin[a];
                if (!(std::fabs(m) < 8388608.f) || m != std::floor(m) || std::fmod(m, 16.0f) != 0.0f || s.worldMax[a] != m + 16.0f) {
                    err = "sub-chunk " + std::to_string(i) + ": bounds are not a 16-aligned integer box"; return BLOKB_E_UNSUPPORTED;
                }
                lo[a] = std::min<int64_t>(lo[a], (int64_t)m); hi[a] = std::max<int64_t>(hi[a], (int64_t)m + 16);
            }
        }
        for (int a = 0; a < 3; a++) { w.origin[a] = (int32_t)lo[a]; w.dim[a] = (int32_t)(hi[a] - lo[a]); }
    }
    const int64_t g[3] = {w.dim[0] / 16, w.dim[1] / 16, w.dim[2] / 16};
    if (g[0] * g[1] * g[2] > (int64_t(1) << 31)) { err = "world bounding box too large for the dense top grid"; return BLOKB_E_UNSUPPORTED; }
    int maxDim = std::max(w.dim[0], std::max(w.dim[1], w.dim[2]));
    w.levels = 3;
    while ((int64_t(1) << (2 * w.levels)) < maxDim) w.levels++;
    if (w.levels > kMaxLevels) { err = "world too large"; return BLOKB_E_UNSUPPORTED; }
    w.topGrid.assign(size_t(g[0] * g[1] * g[2]), 0xFFFFFFFFu);
    // top grid; later sub-chunks at the same cell do not occur in reference output (one entry per region)
    for (size_t i = 0; i < nSubs; i++) {
        int64_t c[3];
        for (int a = 0; a < 3; a++) c[a] = ((int64_t)subs[i].worldMin[a] - w.origin[a]) / 16;
        uint32_t& cell = w.topGrid[size_t((c[2] * g[1] + c[1]) * g[0] + c[0])];
        if (cell != 0xFFFFFFFFu) { err = "two sub-chunks cover the same 16^3 region"; return BLOKB_E_UNSUPPORTED; }
        cell = (uint32_t)i;
    }
    // dense mask pyramid, levels 3..levels. Level-3 bit = sub-chunk present *and* non-empty.
    w.dense.assign(w.levels + 1, {});
    for (int k = 3; k <= w.levels; k++) {
        const int64_t cs = int64_t(1) << (2 * k);
        for (int a = 0; a < 3; a++) w.denseDim[k][a] = (int32_t)((w.dim[a] + cs - 1) / cs);
        w.dense[k].assign(size_t(w.denseDim[k][0]) * w.denseDim[k][1] * w.denseDim[k][2], 0ull);
    }
    for (int64_t z = 0; z < g[2]; z++)
        for (int64_t y = 0; y < g[1]; y++)
            for (int64_t x = 0; x < g[0]; x++) {
                uint32_t p = w.topGrid[size_t((z * g[1] + y) * g[0] + x)];
                if (p == 0xFFFFFFFFu || w.subs[p].brickMask == 0) continue;
                const int32_t* d = w.denseDim[3];
                w.dense[3][size_t(((z >> 2) * d[1] + (y >> 2)) * d[0] + (x >> 2))] |= 1ull << ((x & 3) + 4 * (y & 3) + 16 * (z & 3));
            }
    for (int k = 4; k <= w.levels; k++) {
        const int32_t* dc = w.denseDim[k - 1]; const int32_t* dp = w.denseDim[k];
        for (int64_t z = 0; z < dc[2]; z++)
            for (int64_t y = 0; y < dc[1]; y++)
                for (int64_t x = 0; x < dc[0]; x++)
                    if (w.dense[k - 1][size_t((z * dc[1] + y) * dc[0] + x)])
                        w.dense[k][size_t(((z >> 2) * dp[1] + (y >> 2)) * dp[0] + (x >> 2))] |= 1ull << ((x & 3) + 4 * (y & 3) + 16 * (z & 3));
    }
    // ---- the uploaded tree: breadth-first Node array (scene.h). Upper levels from the dense pyramid, then one node per
    // non-empty sub-chunk (children of a level-3 cell contiguous, in bit order), then the bricks in their existing order.
    struct Cell { int32_t x, y, z; };
    std::vector<Cell> cur{{0, 0, 0}}, nxt;
    w.nodes.clear();
    w.nodes.push_back(Node{w.dense[w.levels][0], 0u, 0u});
    size_t levelStart = 0;  // index of the first node of the level being expanded
    uint64_t nUpper = 1;
    for (int k = w.levels; k >= 4; k--) {
        const int32_t* dc = w.denseDim[k - 1];
        nxt.clear();
        const size_t childStart = w.nodes.size();
        for (size_t i = 0; i < cur.size(); i++) {
            Node& n = w.nodes[levelStart + i];
            n.base = (uint32_t)(childStart + nxt.size());
            for (uint64_t m = n.mask; m; m &= m - 1) {
                const int b = __builtin_ctzll(m);
                nxt.push_back(Cell{cur[i].x * 4 + (b & 3), cur[i].y * 4 + ((b >> 2) & 3), cur[i].z * 4 + (b >> 4)});
            }
        }
        for (const Cell& c : nxt) w.nodes.push_back(Node{w.dense[k - 1][size_t((int64_t(c.z) * dc[1] + c.y) * dc[0] + c.x)], 0u, 0u});
        levelStart = childStart;
        cur.swap(nxt);
        nUpper = w.nodes.size();
    }
    // cur = level-3 cells; their children are sub-chunks
    uint64_t nSubNodes = 0;
    for (size_t i = 0; i < cur.size(); i++) nSubNodes += __builtin_popcountll(w.nodes[levelStart + i].mask);
    brickStart = nUpper + nSubNodes;
    if (brickStart + nBricks > 0xFFFFFFF0ull) { err = "world exceeds 32-bit node indices"; return BLOKB_E_UNSUPPORTED; }
    w.nodes.reserve(size_t(brickStart + nBricks));
    for (size_t i = 0; i < cur.size(); i++) {
        w.nodes[levelStart + i].base = (uint32_t)w.nodes.size();
        const uint64_t mask = w.nodes[levelStart + i].mask;
        for (uint64_t m = mask; m; m &= m - 1) {
            const int b = __builtin_ctzll(m);
            const int64_t x = cur[i].x * 4 + (b & 3), y = cur[i].y * 4 + ((b >> 2) & 3), z = cur[i].z * 4 + (b >> 4);
            const uint32_t p = w.topGrid[size_t((z * g[1] + y) * g[0] + x)];
            w.nodes.push_back(Node{w.subs[p].brickMask, (uint32_t)(brickStart + w.subs[p].brickBase), p});
        }
    }
    return 0;
}

int build_host_world(const blokb_svo_node* nodes, size_t nNodes, const blokb_subchunk* subs, size_t nSubs, HostWorld& w,
                     std::string& err) {
    w = HostWorld();
    w.subs.resize(nSubs);
    // pass 1 (parallel): decode each sub-chunk
    unsigned nt = std::max(1u, std::min(16u, std::thread::hardware_concurrency()));
    if (nSubs < 256) nt = 1;
    std::vector<uint8_t> bad(nt, 0);
    // sizes first, then fill: keep memory bounded by processing in slabs
    std::vector<uint32_t> nBricks(nSubs), nBlocks(nSubs), nVox(nSubs);
    std::vector<uint32_t> maxMat(nt, 0);
    auto count_range = [&](size_t b, size_t e, unsigned tid) {
        SubOut* o = new SubOut;
        for (size_t i = b; i < e; i++) {
            o->brickMask = 0; o->bad = false;
            decode_sub(nodes, nNodes, subs[i], *o);
            if (o->bad) bad[tid] = 1;
            uint32_t nb = 0, nk = 0, nv = 0;
            for (uint64_t m = o->brickMask; m; m &= m - 1) {
                uint32_t bb = __builtin_ctzll(m);
                uint64_t vm = o->bricks[bb].voxelMask;
                nb++; nv += __builtin_popcountll(vm);
                for (uint32_t k = 0; k < 8; k++) if (vm & block_bits(k)) nk++;
                for (uint64_t v = vm; v; v &= v - 1) maxMat[tid] = std::max(maxMat[tid], o->mats[bb][__builtin_ctzll(v)]);
            }
            nBricks[i] = nb; nBlocks[i] = nk; nVox[i] = nv;
        }
        delete o;
    };
    auto run = [&](auto&& fn) {
        if (nt == 1) { fn(0, nSubs, 0); return; }
        std::vector<std::thread> th;
        for (unsigned t = 0; t < nt; t++) th.emplace_back(fn, nSubs * t / nt, nSubs * (t + 1) / nt, t);
        for (auto& x : th) x.join();
    };
    run(count_range);
    for (auto b : bad) if (b) { err = "filled interior node / non-leaf at leaf depth: not a layout the reference's SvoTree::insertVoxel produces"; return BLOKB_E_UNSUPPORTED; }
    uint32_t mm = 0; for (auto m : maxMat) mm = std::max(mm, m);
    w.matWide = mm > 0xFFFFu;
    std::vector<uint64_t> brickOff(nSubs + 1, 0), blockOff(nSubs + 1, 0), voxOff(nSubs + 1, 0);
    for (size_t i = 0; i < nSubs; i++) { brickOff[i + 1] = brickOff[i] + nBricks[i]; blockOff[i + 1] = blockOff[i] + nBlocks[i]; voxOff[i + 1] = voxOff[i] + nVox[i]; }
    if (brickOff[nSubs] > 0xFFFFFFF0ull || blockOff[nSubs] > 0xFFFFFFF0ull || voxOff[nSubs] > 0xFFFFFFF0ull) { err = "world exceeds 32-bit compact indices"; return BLOKB_E_UNSUPPORTED; }
    w.bricks.resize(brickOff[nSubs]); w.blockFirstLeaf.resize(blockOff[nSubs]); w.voxels = voxOff[nSubs];
    if (w.matWide) w.mat32.resize(voxOff[nSubs]); else w.mat16.resize(voxOff[nSubs]);
    auto fill_range = [&](size_t b, size_t e, unsigned) {
        SubOut* o = new SubOut;
        for (size_t i = b; i < e; i++) {
            o->brickMask = 0; o->bad = false;
            decode_sub(nodes, nNodes, subs[i], *o);
            w.subs[i] = SubRec{o->brickMask, (uint32_t)brickOff[i], 0u};
            uint64_t bi = brickOff[i], ki = blockOff[i], vi = voxOff[i];
            for (uint64_t m = o->brickMask; m; m &= m - 1) {
                uint32_t bb = __builtin_ctzll(m);
                uint64_t vm = o->bricks[bb].voxelMask;
                w.bricks[bi++] = Brick{vm, (uint32_t)ki, (uint32_t)vi};
                for (uint32_t k = 0; k < 8; k++) if (vm & block_bits(k)) w.blockFirstLeaf[ki++] = o->firstLeaf[bb][k];
                for (uint64_t v = vm; v; v &= v - 1) {
                    uint32_t id = o->mats[bb][__builtin_ctzll(v)];
                    if (w.matWide) w.mat32[vi++] = id; else w.mat16[vi++] = (uint16_t)id;
                }
            }
        }
        delete o;
    };
    run(fill_range);
    uint64_t brickStart = 0;
    if (int rc = build_top(subs, nSubs, w.bricks.size(), w, brickStart, err)) return rc;
    for (const Brick& b : w.bricks) w.nodes.push_back(Node{b.voxelMask, b.matBase, b.blockBase});
    return 0;
}

}  // namespace blokb
