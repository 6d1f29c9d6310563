// Scene build on the device: voxel list -> compact 64-tree (scene.h) without ever materialising
// the reference's 16-byte node arrays, yet carrying the reference's node numbering exactly.
//
// What the reference does on the CPU, per dirty chunk (blok/src/chunk_manager.cpp:106-140, svo.cpp:36-101): replay
// insertVoxel over the dense 128^3 grid in (z, y, x) loop order; a node's eight children are allocated the first time a
// voxel inside it is inserted, top-down along the path. Then packChunksToGpuSvo (chunk_manager.cpp:234-314) concatenates
// the chunks and emits one SubChunkGpu per non-empty 16^3 region. Hit records must carry those node indices, so the
// numbering is part of the contract. It has a closed form:
//   * t = x + 128 y + 16384 z is a voxel's insertion time; tFirst(node) = min t inside the node;
//   * child blocks are allocated in (tFirst(node), level(node)) order, 8 nodes each, after the root at index 0;
//   * the levels at which a voxel t is the first one of its ancestor form a suffix {L0(t) .. 6}, so the block rank of the
//     level-L ancestor is   rank = #{interior nodes allocated before voxel t} + (L - L0(t)),
//     and the first term is a prefix sum of (7 - L0) over voxels in t order — only the first voxel of each 2^3 block
//     contributes. firstChild(node) = 1 + 8 rank; a chunk has 1 + 8 * (#interior nodes) nodes.
// The prefix sum runs over the 16384 (z, y) voxel rows of a chunk, then inside a row over the <= 64 blocks whose first
// voxel lies in it; tFirst for levels 5 and 6 comes straight from the 64-bit brick masks (bit = x + 4y + 16z is already
// (z, y, x)-ordered), for levels 0-4 from a small atomicMin pyramid (4681 cells per chunk).
//
// Pipeline (all on the context's stream; the few host round trips fetch counts): keys -> radix sort + last-write-wins
// dedupe (cub) -> chunk table -> brick masks + pyramid -> per-sub-chunk counts -> scans -> brick nodes, voxelMat ->
// row counts -> scan -> ranks (blockFirstLeaf, sub-chunk roots) -> sub-chunk records. The O(#sub-chunks) top of the tree
// (root ... sub-chunk nodes) is finished on the host by build_top (scene_build.cpp), shared with the host builder.
#include "build.h"

#include <cub/cub.cuh>

#include <algorithm>
#include <cstring>
#include <iterator>

namespace blokb {
namespace build {
namespace {
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr int kBias = 8192;  // chunk coordinates in [-8192, 8191] (|world coordinate| < 2^20)
__host__ __device__ inline uint32_t lvl_off(int L) { return ((1u << (3 * L)) - 1u) / 7u; }  // cells above level L: 0,1,9,73,585
constexpr uint32_t kSmall = 4681;  // levels 0..4

__device__ __forceinline__ int floordiv128(int v) { return v >> 7; }  // arithmetic shift == floor division by 128
__device__ __forceinline__ uint64_t block_bits(uint32_t k) { return 0x0000000000330033ull << (2 * (k & 1) + 8 * ((k >> 1) & 1) + 32 * (k >> 2)); }

__global__ void k_make_keys(const int32_t* __restrict__ xyz, size_t n, uint64_t* __restrict__ keys, uint32_t* __restrict__ idx, int* __restrict__ err) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    const int cx = floordiv128(x), cy = floordiv128(y), cz = floordiv128(z);
    if (cx < -kBias || cx >= kBias || cy < -kBias || cy >= kBias || cz < -kBias || cz >= kBias) { *err = 1; }
    const uint64_t t = uint64_t(x - cx * 128) | uint64_t(y - cy * 128) << 7 | uint64_t(z - cz * 128) << 14;
    keys[i] = uint64_t((cz + kBias) & 16383) << 49 | uint64_t((cy + kBias) & 16383) << 35 | uint64_t((cx + kBias) & 16383) << 21 | t;
    idx[i] = (uint32_t)i;
}
// keep the last element of every run of equal keys (setVoxelMaterial semantics: the last write wins; the sort is stable)
__global__ void k_flag_last(const uint64_t* __restrict__ keys, size_t n, uint8_t* __restrict__ flag) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    flag[i] = (i + 1 == n || keys[i] != keys[i + 1]) ? 1 : 0;
}
__global__ void k_chunk_heads(const uint64_t* __restrict__ keys, size_t m, uint32_t* __restrict__ head) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= m) return;
    head[i] = (i == 0 || (keys[i] >> 21) != (keys[i - 1] >> 21)) ? 1u : 0u;
}
// chunkIdx (inclusive scan of heads) - 1; heads record the chunk's key
__global__ void k_chunk_table(const uint64_t* __restrict__ keys, size_t m, uint32_t* __restrict__ chunkIdx, const uint32_t* __restrict__ head,
                              uint64_t* __restrict__ chunkKey) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const uint32_t c = chunkIdx[i] - 1u;
    chunkIdx[i] = c;
    if (head[i]) chunkKey[c] = keys[i] >> 21;
}
// brick masks + tFirst pyramid (levels 0..4) + largest material id
__global__ void k_scatter(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ chunkIdx, const uint32_t* __restrict__ idx,
                          const uint32_t* __restrict__ mat, size_t m, unsigned long long* __restrict__ brickMask, uint32_t* __restrict__ first,
                          uint32_t* __restrict__ maxMat) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    uint32_t mm = 0;
    if (i < m) {
        const uint32_t c = chunkIdx[i], t = uint32_t(keys[i]) & 0x1FFFFFu;
        const uint32_t x = t & 127u, y = (t >> 7) & 127u, z = t >> 14;
        atomicOr(&brickMask[size_t(c) * 32768 + ((x >> 2) + 32 * (y >> 2) + 1024 * (z >> 2))], 1ull << ((x & 3) + 4 * (y & 3) + 16 * (z & 3)));
        uint32_t* f = first + size_t(c) * kSmall;
        for (int L = 4; L >= 0; L--) {
            const int sh = 7 - L, nn = 1 << L;
            const uint32_t cell = (x >> sh) + nn * (y >> sh) + nn * nn * (z >> sh);
            if (atomicMin(&f[lvl_off(L) + cell], t) < t) break;  // an earlier voxel owns this cell and everything above it
        }
        mm = mat[idx[i]];
    }
    mm = __reduce_max_sync(0xFFFFFFFFu, mm);
    if ((threadIdx.x & 31) == 0 && mm) atomicMax(maxMat, mm);
}

// One warp per (chunk, 16^3 sub-chunk): lane l owns bricks l and l + 32 (brick bit bb = bx + 4 by + 16 bz).
struct SubWarp {
    uint32_t c, s, sx, sy, sz;
    uint64_t vm[2];
};
__device__ __forceinline__ SubWarp load_sub(const unsigned long long* __restrict__ brickMask, uint32_t gsub, int lane) {
    SubWarp w;
    w.c = gsub >> 9; w.s = gsub & 511u;
    w.sx = w.s & 7u; w.sy = (w.s >> 3) & 7u; w.sz = w.s >> 6;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint32_t bb = lane + 32 * h;
        const uint32_t bx = 4 * w.sx + (bb & 3u), by = 4 * w.sy + ((bb >> 2) & 3u), bz = 4 * w.sz + (bb >> 4);
        w.vm[h] = brickMask[size_t(w.c) * 32768 + (bx + 32 * by + 1024 * bz)];
    }
    return w;
}
__device__ __forceinline__ uint32_t nonempty_blocks(uint64_t vm) {  // 8-bit mask of non-empty 2^3 blocks
    uint32_t r = 0;
#pragma unroll
    for (uint32_t k = 0; k < 8; k++) r |= (vm & block_bits(k)) ? (1u << k) : 0u;
    return r;
}
__global__ void k_sub_counts(const unsigned long long* __restrict__ brickMask, uint32_t nSubDense, uint32_t* __restrict__ flag,
                             uint32_t* __restrict__ nBricks, uint32_t* __restrict__ nBlocks, uint32_t* __restrict__ nVox) {
    const uint32_t gsub = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (gsub >= nSubDense) return;
    const SubWarp w = load_sub(brickMask, gsub, lane);
    uint32_t b = (w.vm[0] != 0) + (w.vm[1] != 0);
    uint32_t k = __popc(nonempty_blocks(w.vm[0])) + __popc(nonempty_blocks(w.vm[1]));
    uint32_t v = __popcll(w.vm[0]) + __popcll(w.vm[1]);
    b = __reduce_add_sync(0xFFFFFFFFu, b); k = __reduce_add_sync(0xFFFFFFFFu, k); v = __reduce_add_sync(0xFFFFFFFFu, v);
    if (lane == 0) { flag[gsub] = b != 0; nBricks[gsub] = b; nBlocks[gsub] = k; nVox[gsub] = v; }
}
__device__ __forceinline__ uint32_t warp_excl_scan(uint32_t v, int lane, uint32_t& total) {
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= d) inc += o;
    }
    total = __shfl_sync(0xFFFFFFFFu, inc, 31);
    return inc - v;
}
// Brick nodes (scene.h level 1), the dense -> compact brick index, and the per-sub-chunk {brickMask, brickBase} records.
__global__ void k_brick_fill(const unsigned long long* __restrict__ brickMask, uint32_t nSubDense, const uint32_t* __restrict__ flag,
                             const uint32_t* __restrict__ primIdx, const uint32_t* __restrict__ brickOff, const uint32_t* __restrict__ blockOff,
                             const uint32_t* __restrict__ voxOff, Node* __restrict__ bricks, uint32_t* __restrict__ brickIdx, SubRec* __restrict__ recs) {
    const uint32_t gsub = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (gsub >= nSubDense) return;
    const SubWarp w = load_sub(brickMask, gsub, lane);
    const uint32_t lo = __ballot_sync(0xFFFFFFFFu, w.vm[0] != 0), hi = __ballot_sync(0xFFFFFFFFu, w.vm[1] != 0);
    uint32_t totV0, totK0, tmp;
    const uint32_t v0 = warp_excl_scan(__popcll(w.vm[0]), lane, totV0), v1 = warp_excl_scan(__popcll(w.vm[1]), lane, tmp) + totV0;
    const uint32_t k0 = warp_excl_scan(__popc(nonempty_blocks(w.vm[0])), lane, totK0), k1 = warp_excl_scan(__popc(nonempty_blocks(w.vm[1])), lane, tmp) + totK0;
    const uint32_t bo = brickOff[gsub], ko = blockOff[gsub], vo = voxOff[gsub];
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint32_t bb = lane + 32 * h;
        const uint32_t bx = 4 * w.sx + (bb & 3u), by = 4 * w.sy + ((bb >> 2) & 3u), bz = 4 * w.sz + (bb >> 4);
        uint32_t bi = kNone;
        if (w.vm[h]) {
            bi = bo + (h ? __popc(lo) + __popc(hi & ((1u << lane) - 1u)) : __popc(lo & ((1u << lane) - 1u)));
            bricks[bi] = Node{w.vm[h], vo + (h ? v1 : v0), ko + (h ? k1 : k0)};
        }
        brickIdx[size_t(w.c) * 32768 + (bx + 32 * by + 1024 * bz)] = bi;
    }
    if (lane == 0 && flag[gsub]) recs[primIdx[gsub]] = SubRec{uint64_t(hi) << 32 | lo, bo, 0u};
}
template <class T>
__global__ void k_voxel_mat(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ chunkIdx, const uint32_t* __restrict__ idx,
                            const uint32_t* __restrict__ mat, size_t m, const uint32_t* __restrict__ brickIdx, const Node* __restrict__ bricks,
                            T* __restrict__ voxelMat) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const uint32_t c = chunkIdx[i], t = uint32_t(keys[i]) & 0x1FFFFFu;
    const uint32_t x = t & 127u, y = (t >> 7) & 127u, z = t >> 14;
    const Node b = bricks[brickIdx[size_t(c) * 32768 + ((x >> 2) + 32 * (y >> 2) + 1024 * (z >> 2))]];
    const uint32_t bit = (x & 3) + 4 * (y & 3) + 16 * (z & 3);
    voxelMat[b.base + __popcll(b.mask & ((1ull << bit) - 1ull))] = (T)mat[idx[i]];
}

// One warp per block row (chunk, bz, by): the 64 2^3 blocks along x, lane l owns blocks bx = 2l and 2l + 1.
struct BlockInfo {
    uint32_t m8;    // occupancy of the block's 8 voxels, bit = x + 2y + 4z  (== (z, y, x) order)
    uint32_t tF;    // insertion time of its first voxel
    uint32_t cls;   // which of the block row's 4 voxel rows that voxel is in: zo * 2 + yo
    uint32_t L0;    // first level whose ancestor has tFirst == tF (6 if only the block itself)
};
__device__ __forceinline__ BlockInfo block_info(const unsigned long long* __restrict__ brickMask, const uint32_t* __restrict__ first, uint32_t c,
                                                uint32_t bx, uint32_t by, uint32_t bz) {
    BlockInfo b;
    const uint64_t vm = brickMask[size_t(c) * 32768 + ((bx >> 1) + 32 * (by >> 1) + 1024 * (bz >> 1))];
    const uint32_t sh = 2 * (bx & 1) + 8 * (by & 1) + 32 * (bz & 1);
    const uint64_t q = vm >> sh;  // block voxel (xo, yo, zo) sits at bit xo + 4 yo + 16 zo of q
    b.m8 = uint32_t(q & 3u) | uint32_t((q >> 4) & 3u) << 2 | uint32_t((q >> 16) & 3u) << 4 | uint32_t((q >> 20) & 3u) << 6;
    b.tF = 0; b.cls = 0; b.L0 = 7;
    if (!b.m8) return b;
    const uint32_t o = __ffs(b.m8) - 1;
    const uint32_t X = 2 * bx + (o & 1), Y = 2 * by + ((o >> 1) & 1), Z = 2 * bz + (o >> 2);
    b.tF = X + 128 * Y + 16384 * Z;
    b.cls = (o >> 1);  // zo * 2 + yo
    b.L0 = 6;
    if (uint32_t(__ffsll((long long)vm) - 1) == (X & 3) + 4 * (Y & 3) + 16 * (Z & 3)) {  // also the brick's first voxel (level 5)
        b.L0 = 5;
        const uint32_t* f = first + size_t(c) * kSmall;
        for (int L = 4; L >= 0; L--) {
            const int s = 7 - L, nn = 1 << L;
            if (f[lvl_off(L) + (X >> s) + nn * (Y >> s) + nn * nn * (Z >> s)] != b.tF) break;
            b.L0 = L;
        }
    }
    return b;
}
__global__ void k_row_counts(const unsigned long long* __restrict__ brickMask, const uint32_t* __restrict__ first, uint32_t nRows,
                             uint32_t* __restrict__ rowCnt) {
    const uint32_t g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;  // (chunk, bz, by)
    const int lane = threadIdx.x & 31;
    if (g >= nRows) return;
    const uint32_t c = g >> 12, bz = (g >> 6) & 63u, by = g & 63u;
    uint32_t cnt[4] = {0, 0, 0, 0};
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const BlockInfo b = block_info(brickMask, first, c, 2 * lane + h, by, bz);
        if (b.m8) cnt[b.cls] += 7 - b.L0;
    }
#pragma unroll
    for (int r = 0; r < 4; r++) {
        const uint32_t s = __reduce_add_sync(0xFFFFFFFFu, cnt[r]);
        if (lane == 0) rowCnt[size_t(c) * 16384 + (2 * bz + (r >> 1)) * 128 + (2 * by + (r & 1))] = s;
    }
}
// rowBase = exclusive scan of rowCnt over all chunks' rows (global). Block ranks -> blockFirstLeaf, brick ranks, levels 0..4.
__global__ void k_ranks(const unsigned long long* __restrict__ brickMask, const uint32_t* __restrict__ first, uint32_t nRows,
                        const uint32_t* __restrict__ rowBase, const uint32_t* __restrict__ brickIdx, const Node* __restrict__ bricks,
                        uint32_t* __restrict__ blockFirstLeaf, uint32_t* __restrict__ rankSmall, const uint32_t* __restrict__ extra) {
    const uint32_t g = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (g >= nRows) return;
    const uint32_t c = g >> 12, bz = (g >> 6) & 63u, by = g & 63u;
    const uint32_t chunkBase = rowBase[size_t(c) * 16384];  // interior nodes of all earlier chunks
    // sum over earlier chunks of (1 + 8 * interior); extra[c] = all-empty chunks (1 root node each) the reference packs before c
    const uint32_t nodeOffset = c + (extra ? extra[c] : 0u) + 8u * chunkBase;
    BlockInfo b[2];
    uint64_t packed = 0;  // four 16-bit per-class counters
#pragma unroll
    for (int h = 0; h < 2; h++) {
        b[h] = block_info(brickMask, first, c, 2 * lane + h, by, bz);
        if (b[h].m8) packed += uint64_t(7 - b[h].L0) << (16 * b[h].cls);
    }
    uint64_t inc = packed;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint64_t o = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if (lane >= d) inc += o;
    }
    uint64_t excl = inc - packed;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        if (b[h].m8) {
            const uint32_t bx = 2 * lane + h;
            const uint32_t row = (2 * bz + (b[h].cls >> 1)) * 128 + (2 * by + (b[h].cls & 1));
            const uint32_t base = (rowBase[size_t(c) * 16384 + row] - chunkBase) + uint32_t((excl >> (16 * b[h].cls)) & 0xFFFFu);
            // rank of the level-L ancestor = base + (L - L0), L = L0 .. 6
            const uint32_t r6 = base + (6 - b[h].L0);
            const Node bn = bricks[brickIdx[size_t(c) * 32768 + ((bx >> 1) + 32 * (by >> 1) + 1024 * (bz >> 1))]];
            const uint32_t k = (bx & 1) + 2 * (by & 1) + 4 * (bz & 1);
            blockFirstLeaf[bn.aux + __popc(nonempty_blocks(bn.mask) & ((1u << k) - 1u))] = nodeOffset + 1u + 8u * r6;
            const uint32_t X = b[h].tF & 127u, Y = (b[h].tF >> 7) & 127u, Z = b[h].tF >> 14;
            for (int L = (int)b[h].L0; L <= 4; L++) {
                const int s = 7 - L, nn = 1 << L;
                rankSmall[size_t(c) * kSmall + lvl_off(L) + (X >> s) + nn * (Y >> s) + nn * nn * (Z >> s)] = base + (L - b[h].L0);
            }
            excl += uint64_t(7 - b[h].L0) << (16 * b[h].cls);  // the lane's second block follows its first in the same row class
        }
    }
}
// SubChunkGpu records (chunk_manager.cpp:255-303) for the non-empty 16^3 regions, in (chunk, sz, sy, sx) order.
__global__ void k_sub_records(uint32_t nSubDense, const uint32_t* __restrict__ flag, const uint32_t* __restrict__ primIdx,
                              const uint64_t* __restrict__ chunkKey, const uint32_t* __restrict__ rowBase, const uint32_t* __restrict__ rankSmall,
                              uint32_t nChunks, uint32_t totalInterior, blokb_subchunk* __restrict__ subs, const uint32_t* __restrict__ extra) {
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= nSubDense || !flag[g]) return;
    const uint32_t c = g >> 9, s = g & 511u, sx = s & 7u, sy = (s >> 3) & 7u, sz = s >> 6;
    const uint32_t chunkBase = rowBase[size_t(c) * 16384];
    const uint32_t nextBase = (c + 1 < nChunks) ? rowBase[size_t(c + 1) * 16384] : totalInterior;
    const uint32_t parent = rankSmall[size_t(c) * kSmall + lvl_off(2) + (sx >> 1) + 4 * (sy >> 1) + 16 * (sz >> 1)];
    const uint64_t ck = chunkKey[c];
    const int cx = int(ck & 16383u) - kBias, cy = int((ck >> 14) & 16383u) - kBias, cz = int((ck >> 28) & 16383u) - kBias;
    blokb_subchunk o;
    o.nodeOffset = c + (extra ? extra[c] : 0u) + 8u * chunkBase;
    o.rootNodeIndex = 1u + 8u * parent + ((sx & 1) | (sy & 1) << 1 | (sz & 1) << 2);
    o.nodeCount = 1u + 8u * (nextBase - chunkBase);
    o.startDepth = 3u;
    const float ss = 128.0f / 8.0f;
    o.worldMin[0] = float(cx * 128) + float(sx) * ss; o.worldMin[1] = float(cy * 128) + float(sy) * ss; o.worldMin[2] = float(cz * 128) + float(sz) * ss;
    o.subChunkSize = ss;
    for (int a = 0; a < 3; a++) o.worldMax[a] = o.worldMin[a] + ss;
    o.pad0 = 0.0f;
    subs[primIdx[g]] = o;
}
}  // namespace

namespace {
struct Arena {  // scratch allocations freed together
    std::vector<void*> ptrs;
    cudaError_t last = cudaSuccess;
    template <class T> T* get(size_t count) {
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, std::max<size_t>(count * sizeof(T), 16));
        if (e != cudaSuccess) { last = e; return nullptr; }
        ptrs.push_back(p);
        return static_cast<T*>(p);
    }
    ~Arena() { for (void* p : ptrs) cudaFree(p); }
};
inline unsigned blocks_for(size_t n, unsigned bs) { return (unsigned)((n + bs - 1) / bs); }
}  // namespace

void release(DeviceBuild& b) {
    if (b.bricks) cudaFree(b.bricks);
    if (b.blockFirstLeaf) cudaFree(b.blockFirstLeaf);
    if (b.voxelMat) cudaFree(b.voxelMat);
    if (b.dSubs) cudaFree(b.dSubs);
    b = DeviceBuild{};
}

#define BCK(expr)                                                                                                   \
    do {                                                                                                            \
        cudaError_t e_ = (expr);                                                                                    \
        if (e_ != cudaSuccess) { err = std::string(#expr) + ": " + cudaGetErrorString(e_); release(out); return BLOKB_E_CUDA; } \
    } while (0)
#define BNULL(p) \
    if (!(p)) { err = std::string("cudaMalloc: ") + cudaGetErrorString(A.last); release(out); return BLOKB_E_CUDA; }

int from_device_voxels(const int32_t* dXyz, const uint32_t* dMat, size_t n, cudaStream_t st, DeviceBuild& out, std::string& err,
                       const std::vector<uint64_t>* emptyChunkKeys) {
    out = DeviceBuild{};
    if (n == 0) return 0;
    if (n > 0xFFFFFFF0ull) { err = "too many voxels"; return BLOKB_E_UNSUPPORTED; }
    Arena A;
    cudaEvent_t ev[3];
    for (auto& e : ev) BCK(cudaEventCreate(&e));
    struct EvGuard { cudaEvent_t* e; ~EvGuard() { for (int i = 0; i < 3; i++) cudaEventDestroy(e[i]); } } evg{ev};
    constexpr unsigned BS = 256;

    // ---- keys, stable sort, last-write-wins dedupe -----------------------------------------------------------------
    uint64_t* keysA = A.get<uint64_t>(n); BNULL(keysA)
    uint64_t* keysB = A.get<uint64_t>(n); BNULL(keysB)
    uint32_t* idxA = A.get<uint32_t>(n); BNULL(idxA)
    uint32_t* idxB = A.get<uint32_t>(n); BNULL(idxB)
    uint8_t* keep = A.get<uint8_t>(n); BNULL(keep)
    int* dScalars = A.get<int>(8); BNULL(dScalars)  // [0] range error, [1] max material, [2] unique count
    BCK(cudaMemsetAsync(dScalars, 0, 8 * sizeof(int), st));
    BCK(cudaEventRecord(ev[0], st));
    k_make_keys<<<blocks_for(n, BS), BS, 0, st>>>(dXyz, n, keysA, idxA, dScalars);
    size_t tmpBytes = 0, t2 = 0;
    cub::DoubleBuffer<uint64_t> kb(keysA, keysB);
    cub::DoubleBuffer<uint32_t> ib(idxA, idxB);
    BCK(cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, kb, ib, (int)n, 0, 63, st));
    BCK(cub::DeviceSelect::Flagged(nullptr, t2, keysA, keep, keysB, dScalars + 2, (int)n, st));
    tmpBytes = std::max(tmpBytes, t2);
    BCK(cub::DeviceScan::InclusiveSum(nullptr, t2, idxA, idxB, (int)n, st));
    tmpBytes = std::max(tmpBytes, t2);
    // later scans run over nChunks * 16384 rows; sized once the chunk count is known (see below)
    void* dTmp = A.get<uint8_t>(tmpBytes); BNULL(dTmp)
    BCK(cub::DeviceRadixSort::SortPairs(dTmp, tmpBytes, kb, ib, (int)n, 0, 63, st));
    uint64_t* sk = kb.Current(); uint64_t* skAlt = kb.Alternate();
    uint32_t* si = ib.Current(); uint32_t* siAlt = ib.Alternate();
    k_flag_last<<<blocks_for(n, BS), BS, 0, st>>>(sk, n, keep);
    BCK(cub::DeviceSelect::Flagged(dTmp, tmpBytes, sk, keep, skAlt, dScalars + 2, (int)n, st));
    BCK(cub::DeviceSelect::Flagged(dTmp, tmpBytes, si, keep, siAlt, dScalars + 2, (int)n, st));
    int hs[8];
    BCK(cudaMemcpyAsync(hs, dScalars, sizeof hs, cudaMemcpyDeviceToHost, st));
    BCK(cudaStreamSynchronize(st));
    if (hs[0]) { err = "voxel coordinate outside +-2^20 (chunk index outside [-8192, 8191])"; return BLOKB_E_UNSUPPORTED; }
    const size_t m = (size_t)(uint32_t)hs[2];
    const uint64_t* keys = skAlt;      // unique, sorted by (cz, cy, cx, z, y, x)
    const uint32_t* idx = siAlt;       // original index of the surviving write
    // ---- chunk table -----------------------------------------------------------------------------------------------
    uint32_t* head = si;               // reuse: n x u32
    uint32_t* chunkIdx = A.get<uint32_t>(m); BNULL(chunkIdx)
    k_chunk_heads<<<blocks_for(m, BS), BS, 0, st>>>(keys, m, head);
    BCK(cub::DeviceScan::InclusiveSum(dTmp, tmpBytes, head, chunkIdx, (int)m, st));
    uint32_t nChunks = 0;
    BCK(cudaMemcpyAsync(&nChunks, chunkIdx + (m - 1), 4, cudaMemcpyDeviceToHost, st));
    BCK(cudaStreamSynchronize(st));
    if (uint64_t(nChunks) * 16384ull > 0x7FFFFFFFull) { err = "too many chunks for the device builder (use blokb_upload_world)"; return BLOKB_E_UNSUPPORTED; }
    uint64_t* chunkKey = A.get<uint64_t>(nChunks); BNULL(chunkKey)
    k_chunk_table<<<blocks_for(m, BS), BS, 0, st>>>(keys, m, chunkIdx, head, chunkKey);
    BCK(cudaEventRecord(ev[1], st));
    // ---- per-chunk dense scratch --------------------------------------------------------------------------------------
    const size_t nSubDense = size_t(nChunks) * 512, nRowsW = size_t(nChunks) * 4096, nRows = size_t(nChunks) * 16384;
    unsigned long long* brickMask = A.get<unsigned long long>(size_t(nChunks) * 32768); BNULL(brickMask)
    uint32_t* brickIdx = A.get<uint32_t>(size_t(nChunks) * 32768); BNULL(brickIdx)
    uint32_t* first = A.get<uint32_t>(size_t(nChunks) * kSmall); BNULL(first)
    uint32_t* rankSmall = A.get<uint32_t>(size_t(nChunks) * kSmall); BNULL(rankSmall)
    uint32_t* rowCnt = A.get<uint32_t>(nRows); BNULL(rowCnt)
    uint32_t* rowBase = A.get<uint32_t>(nRows); BNULL(rowBase)
    uint32_t* cnt4 = A.get<uint32_t>(4 * nSubDense); BNULL(cnt4)   // flag, nBricks, nBlocks, nVox
    uint32_t* off4 = A.get<uint32_t>(4 * nSubDense); BNULL(off4)   // their exclusive scans
    size_t t3 = 0;
    BCK(cub::DeviceScan::ExclusiveSum(nullptr, t3, rowCnt, rowBase, (int)nRows, st));
    void* dTmp2 = dTmp; size_t tmp2Bytes = tmpBytes;
    if (t3 > tmpBytes) { dTmp2 = A.get<uint8_t>(t3); BNULL(dTmp2) tmp2Bytes = t3; }
    BCK(cudaMemsetAsync(brickMask, 0, size_t(nChunks) * 32768 * 8, st));
    BCK(cudaMemsetAsync(first, 0xFF, size_t(nChunks) * kSmall * 4, st));
    BCK(cudaMemsetAsync(rankSmall, 0xFF, size_t(nChunks) * kSmall * 4, st));
    k_scatter<<<blocks_for(m, BS), BS, 0, st>>>(keys, chunkIdx, idx, dMat, m, brickMask, first, reinterpret_cast<uint32_t*>(dScalars + 1));
    // ---- sub-chunk counts and offsets ---------------------------------------------------------------------------------
    uint32_t *flag = cnt4, *nBr = cnt4 + nSubDense, *nBl = cnt4 + 2 * nSubDense, *nVx = cnt4 + 3 * nSubDense;
    uint32_t *primIdx = off4, *brickOff = off4 + nSubDense, *blockOff = off4 + 2 * nSubDense, *voxOff = off4 + 3 * nSubDense;
    k_sub_counts<<<blocks_for(nSubDense * 32, BS), BS, 0, st>>>(brickMask, (uint32_t)nSubDense, flag, nBr, nBl, nVx);
    for (int k = 0; k < 4; k++) BCK(cub::DeviceScan::ExclusiveSum(dTmp2, tmp2Bytes, cnt4 + k * nSubDense, off4 + k * nSubDense, (int)nSubDense, st));
    uint32_t lastCnt[4], lastOff[4], maxMat = 0;
    for (int k = 0; k < 4; k++) {
        BCK(cudaMemcpyAsync(&lastCnt[k], cnt4 + k * nSubDense + (nSubDense - 1), 4, cudaMemcpyDeviceToHost, st));
        BCK(cudaMemcpyAsync(&lastOff[k], off4 + k * nSubDense + (nSubDense - 1), 4, cudaMemcpyDeviceToHost, st));
    }
    BCK(cudaMemcpyAsync(&maxMat, dScalars + 1, 4, cudaMemcpyDeviceToHost, st));
    BCK(cudaStreamSynchronize(st));
    out.nSubs = uint64_t(lastCnt[0]) + lastOff[0]; out.nBricks = uint64_t(lastCnt[1]) + lastOff[1];
    out.nBlocks = uint64_t(lastCnt[2]) + lastOff[2]; out.nVox = uint64_t(lastCnt[3]) + lastOff[3];
    out.nChunks = nChunks; out.matWide = maxMat > 0xFFFFu;
    if (out.nVox != m) { err = "internal: voxel count mismatch in the device builder"; release(out); return BLOKB_E_CUDA; }
    // ---- outputs --------------------------------------------------------------------------------------------------------
    BCK(cudaMalloc((void**)&out.bricks, std::max<size_t>(out.nBricks * sizeof(Node), 16)));
    BCK(cudaMalloc((void**)&out.blockFirstLeaf, std::max<size_t>(out.nBlocks * 4, 16)));
    BCK(cudaMalloc(&out.voxelMat, std::max<size_t>(out.nVox * (out.matWide ? 4 : 2), 16)));
    BCK(cudaMalloc((void**)&out.dSubs, std::max<size_t>(out.nSubs * sizeof(blokb_subchunk), 16)));
    SubRec* dRecs = A.get<SubRec>(out.nSubs); BNULL(dRecs)
    k_brick_fill<<<blocks_for(nSubDense * 32, BS), BS, 0, st>>>(brickMask, (uint32_t)nSubDense, flag, primIdx, brickOff, blockOff, voxOff, out.bricks,
                                                                brickIdx, dRecs);
    if (out.matWide) k_voxel_mat<uint32_t><<<blocks_for(m, BS), BS, 0, st>>>(keys, chunkIdx, idx, dMat, m, brickIdx, out.bricks, static_cast<uint32_t*>(out.voxelMat));
    else k_voxel_mat<uint16_t><<<blocks_for(m, BS), BS, 0, st>>>(keys, chunkIdx, idx, dMat, m, brickIdx, out.bricks, static_cast<uint16_t*>(out.voxelMat));
    // ---- the reference's node numbering ---------------------------------------------------------------------------------
    k_row_counts<<<blocks_for(nRowsW * 32, BS), BS, 0, st>>>(brickMask, first, (uint32_t)nRowsW, rowCnt);
    BCK(cub::DeviceScan::ExclusiveSum(dTmp2, tmp2Bytes, rowCnt, rowBase, (int)nRows, st));
    uint32_t lastRowCnt = 0, lastRowBase = 0;
    BCK(cudaMemcpyAsync(&lastRowCnt, rowCnt + (nRows - 1), 4, cudaMemcpyDeviceToHost, st));
    BCK(cudaMemcpyAsync(&lastRowBase, rowBase + (nRows - 1), 4, cudaMemcpyDeviceToHost, st));
    BCK(cudaStreamSynchronize(st));
    const uint64_t totalInterior = uint64_t(lastRowCnt) + lastRowBase;
    uint32_t* dExtra = nullptr;
    const size_t nEmpty = emptyChunkKeys ? emptyChunkKeys->size() : 0;
    if (nEmpty) {  // the reference packs an all-empty chunk as a single root node: later chunks' node offsets move up by one each
        std::vector<uint64_t> ck(nChunks);
        BCK(cudaMemcpyAsync(ck.data(), chunkKey, size_t(nChunks) * 8, cudaMemcpyDeviceToHost, st));
        BCK(cudaStreamSynchronize(st));
        std::vector<uint32_t> extra(nChunks);
        for (uint32_t c = 0; c < nChunks; c++)
            extra[c] = uint32_t(std::lower_bound(emptyChunkKeys->begin(), emptyChunkKeys->end(), ck[c]) - emptyChunkKeys->begin());
        dExtra = A.get<uint32_t>(nChunks); BNULL(dExtra)
        BCK(cudaMemcpyAsync(dExtra, extra.data(), size_t(nChunks) * 4, cudaMemcpyHostToDevice, st));
        BCK(cudaStreamSynchronize(st));
    }
    out.refNodes = uint64_t(nChunks) + nEmpty + 8ull * totalInterior;
    if (out.refNodes > 0xFFFFFFF0ull) { err = "world exceeds the reference's 32-bit node indices"; release(out); return BLOKB_E_UNSUPPORTED; }
    k_ranks<<<blocks_for(nRowsW * 32, BS), BS, 0, st>>>(brickMask, first, (uint32_t)nRowsW, rowBase, brickIdx, out.bricks, out.blockFirstLeaf, rankSmall, dExtra);
    k_sub_records<<<blocks_for(nSubDense, BS), BS, 0, st>>>((uint32_t)nSubDense, flag, primIdx, chunkKey, rowBase, rankSmall, nChunks,
                                                            (uint32_t)totalInterior, out.dSubs, dExtra);
    BCK(cudaEventRecord(ev[2], st));
    out.subs.resize(out.nSubs);
    out.recs.resize(out.nSubs);
    if (out.nSubs) {
        BCK(cudaMemcpyAsync(out.subs.data(), out.dSubs, out.nSubs * sizeof(blokb_subchunk), cudaMemcpyDeviceToHost, st));
        BCK(cudaMemcpyAsync(out.recs.data(), dRecs, out.nSubs * sizeof(SubRec), cudaMemcpyDeviceToHost, st));
    }
    BCK(cudaStreamSynchronize(st));
    BCK(cudaGetLastError());
    cudaEventElapsedTime(&out.sortMs, ev[0], ev[1]);
    cudaEventElapsedTime(&out.buildMs, ev[1], ev[2]);
    return 0;
}

int from_voxels(const int32_t* hXyz, const uint32_t* hMat, size_t n, cudaStream_t st, DeviceBuild& out, std::string& err) {
    out = DeviceBuild{};
    if (n == 0) return 0;
    Arena A;
    int32_t* dXyz = A.get<int32_t>(3 * n); BNULL(dXyz)
    uint32_t* dMat = A.get<uint32_t>(n); BNULL(dMat)
    BCK(cudaMemcpyAsync(dXyz, hXyz, 3 * n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    BCK(cudaMemcpyAsync(dMat, hMat, n * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    return from_device_voxels(dXyz, dMat, n, st, out, err, nullptr);  // synchronises the stream before it returns
}

namespace {
// Dense chunk (Chunk::density / materialIds, x + y*C + z*C*C, chunk.hpp:33-42) -> voxel list entries for density > 0
// (buildSvoFromDensity's test, chunk_manager.cpp:112-116). Order is irrelevant: the builder sorts.
__global__ void k_dense_emit(const float* __restrict__ density, const uint32_t* __restrict__ ids, int cx, int cy, int cz, int32_t* __restrict__ xyz,
                             uint32_t* __restrict__ mat, unsigned long long* __restrict__ counter) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;  // 128^3 cells
    const bool filled = density[t] > 0.0f;
    const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, filled);
    if (!ballot) return;
    const int lane = threadIdx.x & 31;
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(counter, (unsigned long long)__popc(ballot));
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
    if (filled) {
        const size_t o = size_t(base) + __popc(ballot & ((1u << lane) - 1u));
        xyz[3 * o] = cx * 128 + int(t & 127u); xyz[3 * o + 1] = cy * 128 + int((t >> 7) & 127u); xyz[3 * o + 2] = cz * 128 + int(t >> 14);
        mat[o] = ids[t];
    }
}
// keep[i] = 1 unless voxel i lies in one of the listed chunks (sorted keys, chunk key = voxel key >> 21)
__global__ void k_flag_outside(const int32_t* __restrict__ xyz, size_t n, const uint64_t* __restrict__ chunkKeys, uint32_t nKeys, uint8_t* __restrict__ keep) {
    const size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int cx = floordiv128(xyz[3 * i]), cy = floordiv128(xyz[3 * i + 1]), cz = floordiv128(xyz[3 * i + 2]);
    const uint64_t key = uint64_t(cz + kBias) << 28 | uint64_t(cy + kBias) << 14 | uint64_t(cx + kBias);
    uint32_t lo = 0, hi = nKeys;
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (chunkKeys[mid] < key) lo = mid + 1; else hi = mid;
    }
    keep[i] = (lo < nKeys && chunkKeys[lo] == key) ? 0 : 1;
}
struct V3 { int32_t x, y, z; };
}  // namespace

void VoxelList::release() {
    if (xyz) cudaFree(xyz);
    if (mat) cudaFree(mat);
    xyz = nullptr; mat = nullptr; n = 0;
    emptyKeys.clear();
}

#undef BCK
#undef BNULL
#define LCK(expr)                                                                                          \
    do {                                                                                                   \
        cudaError_t e_ = (expr);                                                                           \
        if (e_ != cudaSuccess) { err = std::string(#expr) + ": " + cudaGetErrorString(e_); return BLOKB_E_CUDA; } \
    } while (0)
#define LNULL(p) \
    if (!(p)) { err = std::string("cudaMalloc: ") + cudaGetErrorString(A.last); return BLOKB_E_CUDA; }

int replace_chunks(VoxelList& L, const blokb_dense_chunk* chunks, size_t nChunks, cudaStream_t st, std::string& err) {
    constexpr size_t kCells = size_t(128) * 128 * 128;
    size_t nNew = 0;
    std::vector<uint64_t> keys(nChunks), nowEmpty;
    for (size_t c = 0; c < nChunks; c++) {
        if (!chunks[c].density || !chunks[c].materialIds) { err = "dense chunk " + std::to_string(c) + ": null array"; return BLOKB_E_ARG; }
        const int cx = chunks[c].cx, cy = chunks[c].cy, cz = chunks[c].cz;
        if (cx < -kBias || cx >= kBias || cy < -kBias || cy >= kBias || cz < -kBias || cz >= kBias) { err = "chunk index outside [-8192, 8191]"; return BLOKB_E_UNSUPPORTED; }
        keys[c] = uint64_t(cz + kBias) << 28 | uint64_t(cy + kBias) << 14 | uint64_t(cx + kBias);
        const float* dn = chunks[c].density;
        size_t k = 0;
        for (size_t i = 0; i < kCells; i++) k += dn[i] > 0.0f;
        nNew += k;
        if (k == 0) nowEmpty.push_back(keys[c]);
    }
    std::vector<uint64_t> sortedKeys = keys;
    std::sort(sortedKeys.begin(), sortedKeys.end());
    if (std::adjacent_find(sortedKeys.begin(), sortedKeys.end()) != sortedKeys.end()) { err = "dense chunk listed twice"; return BLOKB_E_ARG; }
    if (L.n + nNew > 0xFFFFFFF0ull) { err = "too many voxels"; return BLOKB_E_UNSUPPORTED; }

    Arena A;  // temporaries; the new list arrays are taken out of it on success
    const size_t cap = L.n + nNew;
    int32_t* nx = A.get<int32_t>(3 * std::max<size_t>(cap, 1)); LNULL(nx)
    uint32_t* nm = A.get<uint32_t>(std::max<size_t>(cap, 1)); LNULL(nm)
    unsigned long long* dCount = A.get<unsigned long long>(1); LNULL(dCount)
    unsigned long long kept = 0;
    if (L.n) {
        // stable compaction of the voxels that stay (order matters: a voxel listed twice keeps its last material)
        uint64_t* dKeys = A.get<uint64_t>(std::max<size_t>(nChunks, 1)); LNULL(dKeys)
        uint8_t* keep = A.get<uint8_t>(L.n); LNULL(keep)
        if (nChunks) LCK(cudaMemcpyAsync(dKeys, sortedKeys.data(), nChunks * 8, cudaMemcpyHostToDevice, st));
        k_flag_outside<<<blocks_for(L.n, 256), 256, 0, st>>>(L.xyz, L.n, dKeys, (uint32_t)nChunks, keep);
        size_t t1 = 0, t2 = 0;
        LCK(cub::DeviceSelect::Flagged(nullptr, t1, reinterpret_cast<const V3*>(L.xyz), keep, reinterpret_cast<V3*>(nx), dCount, (int64_t)L.n, st));
        LCK(cub::DeviceSelect::Flagged(nullptr, t2, L.mat, keep, nm, dCount, (int64_t)L.n, st));
        void* tmp = A.get<uint8_t>(std::max(t1, t2)); LNULL(tmp)
        LCK(cub::DeviceSelect::Flagged(tmp, t1, reinterpret_cast<const V3*>(L.xyz), keep, reinterpret_cast<V3*>(nx), dCount, (int64_t)L.n, st));
        LCK(cub::DeviceSelect::Flagged(tmp, t2, L.mat, keep, nm, dCount, (int64_t)L.n, st));
        LCK(cudaMemcpyAsync(&kept, dCount, 8, cudaMemcpyDeviceToHost, st));
        LCK(cudaStreamSynchronize(st));
    }
    // dCount now doubles as the append cursor of k_dense_emit (64-bit; DeviceSelect wrote the kept count into it)
    LCK(cudaMemcpyAsync(dCount, &kept, 8, cudaMemcpyHostToDevice, st));
    if (nNew) {
        float* dDen = A.get<float>(kCells); LNULL(dDen)
        uint32_t* dIds = A.get<uint32_t>(kCells); LNULL(dIds)
        for (size_t c = 0; c < nChunks; c++) {
            LCK(cudaMemcpyAsync(dDen, chunks[c].density, kCells * 4, cudaMemcpyHostToDevice, st));
            LCK(cudaMemcpyAsync(dIds, chunks[c].materialIds, kCells * 4, cudaMemcpyHostToDevice, st));
            k_dense_emit<<<(unsigned)(kCells / 256), 256, 0, st>>>(dDen, dIds, chunks[c].cx, chunks[c].cy, chunks[c].cz, nx, nm, dCount);
        }
    }
    unsigned long long got = 0;
    LCK(cudaMemcpyAsync(&got, dCount, 8, cudaMemcpyDeviceToHost, st));
    LCK(cudaStreamSynchronize(st));
    LCK(cudaGetLastError());
    if (got != kept + nNew) { err = "dense chunk arrays changed during the upload"; return BLOKB_E_ARG; }
    // commit: the list takes the new arrays, the old ones go
    for (auto& p : A.ptrs) if (p == nx || p == nm) p = nullptr;
    if (L.xyz) cudaFree(L.xyz);
    if (L.mat) cudaFree(L.mat);
    L.xyz = nx; L.mat = nm; L.n = size_t(got);
    std::vector<uint64_t> e;
    std::set_difference(L.emptyKeys.begin(), L.emptyKeys.end(), sortedKeys.begin(), sortedKeys.end(), std::back_inserter(e));
    std::sort(nowEmpty.begin(), nowEmpty.end());
    L.emptyKeys.clear();
    std::merge(e.begin(), e.end(), nowEmpty.begin(), nowEmpty.end(), std::back_inserter(L.emptyKeys));
    return 0;
}

int from_list(const VoxelList& L, cudaStream_t st, DeviceBuild& out, std::string& err) {
    out = DeviceBuild{};
    if (L.n == 0) { out.refNodes = L.emptyKeys.size(); return 0; }
    return from_device_voxels(L.xyz, L.mat, L.n, st, out, err, &L.emptyKeys);
}

int from_dense(const blokb_dense_chunk* chunks, size_t nChunks, cudaStream_t st, DeviceBuild& out, std::string& err, VoxelList* keep) {
    out = DeviceBuild{};
    VoxelList L;
    int rc = replace_chunks(L, chunks, nChunks, st, err);
    if (!rc) rc = from_list(L, st, out, err);
    if (!rc && keep) { keep->release(); *keep = std::move(L); L = VoxelList{}; }
    L.release();
    return rc;
}

}  // namespace build
}  // namespace blokb
