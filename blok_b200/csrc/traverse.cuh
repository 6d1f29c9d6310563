// Hierarchical 3-D DDA over the compact 64-tree (scene.h) — the B200-side replacement for the reference's
// driver BVH + per-AABB stack traversal (renderer_raytracing.cpp:15-254; intersect.rint:70-205).
//
// Why a DDA reproduces the reference bit for bit (DESIGN.md §3 has the full argument):
//   * every box the reference tests (sub-chunk root, octree children, leaves) has integer bounding planes, and
//     each slab value is the fp32 expression  t(p) = (float(p) - origin) * invDir  (intersect.rint:47-55, 179-186)
//     with invDir = 1 / (|d| < 1e-6 ? 1e-6 : d) (:79);  t(p) is monotone in p, so a child's interval nests inside
//     its parent's in floating point too, and the leaf the reference reports is simply the filled voxel with the
//     smallest  T = max(tmin, entry)  among voxels whose clipped interval is strictly non-empty (:193) — no
//     dependence on traversal order, stack depth or BVH candidate order;
//   * this DDA takes every stepping decision by comparing those same fp32 t(p) values (never an accumulated
//     tMax += tDelta), so it visits exactly the voxels with a strictly non-empty interval, front to back;
//   * coarse cells (4^k) use the same planes, so skipping an empty coarse cell skips only voxels the reference
//     would not have reached either.
// Compiled with --fmad=false: every * and + below rounds once, like the oracle's -ffp-contract=off build.
#pragma once
#include <cstdint>

#include "hd.h"
#include "scene.h"

namespace blokb {

struct HitRec {
    uint32_t node, material;
    float t;
    uint32_t facePrim;
};

BLOKB_HD float plane_t(int planeWorld, float o, float inv) { return (float(planeWorld) - o) * inv; }

// intersect.rint:58-68
BLOKB_HD uint32_t hit_face(float hx, float hy, float hz, float cx, float cy, float cz) {
    const float dx = hx - cx, dy = hy - cy, dz = hz - cz;
    const float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
    if (ax >= ay && ax >= az) return dx > 0.0f ? 0u : 1u;
    if (ay >= az) return dy > 0.0f ? 2u : 3u;
    return dz > 0.0f ? 4u : 5u;
}

BLOKB_HD uint64_t block_bits(uint32_t b) {
    return 0x0000000000330033ull << (2u * (b & 1u) + 8u * ((b >> 1) & 1u) + 32u * ((b >> 2) & 1u));
}

// ---- mirrored frame ---------------------------------------------------------------------------------
// Per axis the walk uses a MIRRORED integer coordinate m in [0, cap), cap = 4^levels, that always grows along the
// ray: m = voxel - origin for a positive direction component, cap - (voxel - origin) - size for a negative one.
// The world coordinate of mirrored plane m is  P(m) = base + sgn*m  (base = origin or origin + cap), so its slab
// value is still the reference's expression (float(P) - o) * inv with the ORIGINAL plane coordinate; in the mirrored
// frame that value is strictly increasing in m on every axis. Consequences:
//   * a cell's exit plane is always m + size; a step is m += size on every axis whose exit value equals the minimum;
//   * the child (of 4 along an axis) holding parameter t is  #{inner planes with value <= t}  — the same rule
//     settle() applied at voxel granularity (value(k) <= t < value(k+1)), evaluated with two plane tests
//     (binary search) at the granularity actually needed. Positions finer than the current cell are never
//     computed: the walk carries only the cell corner, and refines it when (and only when) it descends;
//   * actual child / cell indices are the mirrored ones XOR a per-axis constant (cap - 1 = all ones).
// w* hold float(P(m*)) of the current cell's near planes (exact: |P| < 2^24), so a plane value costs FFMA+FADD+FMUL.

BLOKB_HD float pow2f(int e) { return __uint_as_float(uint32_t(127 + e) << 23); }

// Refine (m, w) from a cell of size 4*s' to the child of size s' = 1 << sh that holds parameter t; adds the child's
// mirrored index (0..3) << axisShift to `bit`.
BLOKB_HD void select_child(uint32_t& m, float& w, uint32_t& bit, int axisShift, uint32_t s, float fs, float fs2, float sgn, float o,
                           float inv, float t) {
    const float p2 = __fmaf_rn(sgn, fs2, w);
    const bool hi = t >= (p2 - o) * inv;
    const float w1 = hi ? p2 : w;
    const float p1 = __fmaf_rn(sgn, fs, w1);
    const bool lo = t >= (p1 - o) * inv;
    w = lo ? p1 : w1;
    if (hi) { m += 2u * s; bit += 2u << axisShift; }
    if (lo) { m += s; bit += 1u << axisShift; }
}

enum : int { WALK_CONTINUE = 0, WALK_HIT = 1, WALK_MISS = 2 };

// One ray's walk as an explicit state machine, so that kernels can interleave / refill rays per lane
// (k_primary's persistent warps) while traverse() below simply runs it to completion.
struct Walker {
    float ox, oy, oz, dx, dy, dz, ix, iy, iz;  // origin, direction, 1 / substituted direction
    float gsx, gsy, gsz;                       // +1 / -1: sign of the mirrored frame per axis
    uint32_t fx, fy, fz;                       // mirrored <-> actual coordinate XOR constants (0 or cap - 1)
    float t, tEnd;
    uint32_t mx, my, mz;                       // mirrored corner of the current child cell
    float wx, wy, wz;                          // float(P) of its near planes
    int sh;                                    // child cell size = 1 << sh; current cell = 4x that
    float fs;                                  // float(1 << sh)
    uint32_t bit;                              // mirrored child index ix + 4 iy + 16 iz inside the current cell
    uint64_t mask;                             // occupancy of the current cell's 4^3 children (actual orientation)
    uint64_t subMask;                          // cached sub-chunk record
    uint32_t subBrickBase, prim;
    uint32_t blockBase, matBase;               // cached brick record (its voxelMask is `mask` while sh == 0)

    BLOKB_HD uint32_t flip() const { return (fx & 3u) | (fy & 12u) | (fz & 48u); }

    BLOKB_HD void select(const DeviceWorld&) {
        const uint32_t s = 1u << sh;
        const float fs2 = fs + fs;
        bit = 0;
        select_child(mx, wx, bit, 0, s, fs, fs2, gsx, ox, ix, t);
        select_child(my, wy, bit, 2, s, fs, fs2, gsy, oy, iy, t);
        select_child(mz, wz, bit, 4, s, fs, fs2, gsz, oz, iz, t);
    }

    // Returns false when the ray misses the grid box altogether.
    BLOKB_HD bool init(const DeviceWorld& W, float ox_, float oy_, float oz_, float dx_, float dy_, float dz_, float tmin, float tmax) {
        ox = ox_; oy = oy_; oz = oz_; dx = dx_; dy = dy_; dz = dz_;
        const float sx = fabsf(dx) < 1e-6f ? 1e-6f : dx, sy = fabsf(dy) < 1e-6f ? 1e-6f : dy, sz = fabsf(dz) < 1e-6f ? 1e-6f : dz;
        ix = 1.0f / sx; iy = 1.0f / sy; iz = 1.0f / sz;
        const bool px = ix > 0.0f, py = iy > 0.0f, pz = iz > 0.0f;
        const int gx = W.origin[0], gy = W.origin[1], gz = W.origin[2];
        {   // clip to the grid box (same plane formula)
            const float ax = plane_t(gx, ox, ix), bx = plane_t(gx + W.dim[0], ox, ix);
            const float ay = plane_t(gy, oy, iy), by = plane_t(gy + W.dim[1], oy, iy);
            const float az = plane_t(gz, oz, iz), bz = plane_t(gz + W.dim[2], oz, iz);
            const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
            const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
            t = fmaxf(tn, tmin);
            tEnd = fminf(tf, tmax);
            if (!(t < tEnd)) return false;
        }
        const int L = W.levels;
        const uint32_t cap = 1u << (2 * L);
        fx = px ? 0u : cap - 1u; fy = py ? 0u : cap - 1u; fz = pz ? 0u : cap - 1u;
        gsx = px ? 1.0f : -1.0f; gsy = py ? 1.0f : -1.0f; gsz = pz ? 1.0f : -1.0f;
        mx = my = mz = 0;
        wx = float(px ? gx : gx + int(cap)); wy = float(py ? gy : gy + int(cap)); wz = float(pz ? gz : gz + int(cap));
        sh = 2 * (L - 1);
        fs = pow2f(sh);
        mask = W.dense[L][0];
        subMask = 0ull; subBrickBase = 0; prim = 0; blockBase = 0; matBase = 0;
        select(W);
        return true;
    }

    // One iteration: descend into the occupied child holding the ray, or leave the empty one. Every fetch below is
    // in bounds by construction: the walk only ever selects cells that contain a point of the ray with
    // t in [grid entry, grid exit) under the plane-compare semantics, i.e. cells that intersect the grid box.
    template <bool kStats>
    BLOKB_HD int advance(const DeviceWorld& W, HitRec& out, uint32_t& steps) {
        const uint32_t abit = bit ^ flip();  // actual-orientation child index: x + 4y + 16z
        if ((mask >> abit) & 1ull) {
            // ---- the child holding the ray is occupied: it becomes the current cell
            if (sh == 0) {
                // ---- hit: the child is a filled voxel; t is its slab entry (or tmin / grid entry for the first cell)
                const uint32_t x = mx ^ fx, y = my ^ fy, z = mz ^ fz;
                const float hx = ox + dx * t, hy = oy + dy * t, hz = oz + dz * t;  // intersect.rint:138
                const uint32_t face = hit_face(hx, hy, hz, float(W.origin[0] + int(x)) + 0.5f, float(W.origin[1] + int(y)) + 0.5f,
                                               float(W.origin[2] + int(z)) + 0.5f);
                const uint64_t vm = mask;
                const uint32_t b = ((x >> 1) & 1) | (((y >> 1) & 1) << 1) | (((z >> 1) & 1) << 2);
                uint32_t rank = 0;
                for (uint32_t j = 0; j < 7; j++) rank += (j < b && (vm & block_bits(j))) ? 1u : 0u;
                const uint32_t leaf = W.blockFirstLeaf[blockBase + rank] + ((x & 1) | ((y & 1) << 1) | ((z & 1) << 2));
                const uint32_t mi = matBase + __popcll(vm & ((1ull << abit) - 1ull));
                const uint32_t mat = W.matWide ? reinterpret_cast<const uint32_t*>(W.voxelMat)[mi]
                                               : (uint32_t) reinterpret_cast<const uint16_t*>(W.voxelMat)[mi];
                out.node = leaf; out.material = mat; out.t = t; out.facePrim = face | (prim << 3);
                return WALK_HIT;
            }
            if (sh >= 6) {  // child is a level >= 3 cell: dense mask arrays
                const int k = sh >> 1;
                const uint32_t cx = (mx ^ fx) >> sh, cy = (my ^ fy) >> sh, cz = (mz ^ fz) >> sh;
                BLOKB_ASSERT(cx < (uint32_t)W.denseDim[k][0] && cy < (uint32_t)W.denseDim[k][1] && cz < (uint32_t)W.denseDim[k][2]);
                mask = W.dense[k][(cz * uint32_t(W.denseDim[k][1]) + cy) * uint32_t(W.denseDim[k][0]) + cx];
            } else if (sh == 4) {  // child is a 16^3 sub-chunk
                const uint32_t cx = (mx ^ fx) >> 4, cy = (my ^ fy) >> 4, cz = (mz ^ fz) >> 4;
                BLOKB_ASSERT(cx < (uint32_t(W.dim[0]) >> 4) && cy < (uint32_t(W.dim[1]) >> 4) && cz < (uint32_t(W.dim[2]) >> 4));
                prim = W.topGrid[(cz * (uint32_t(W.dim[1]) >> 4) + cy) * (uint32_t(W.dim[0]) >> 4) + cx];
                const SubRec sr = W.subs[prim];
                subMask = sr.brickMask; subBrickBase = sr.brickBase;
                mask = subMask;
            } else {  // sh == 2: child is a 4^3 brick of the cached sub-chunk
                const Brick br = W.bricks[subBrickBase + __popcll(subMask & ((1ull << abit) - 1ull))];
                mask = br.voxelMask; blockBase = br.blockBase; matBase = br.matBase;
            }
            sh -= 2;
            fs = pow2f(sh);
            select(W);
            return WALK_CONTINUE;
        }
        // ---- the child cell (size 1 << sh) holding the ray is empty: leave it through its nearest exit plane(s)
        if (kStats) steps++;
        const float qx = __fmaf_rn(gsx, fs, wx), qy = __fmaf_rn(gsy, fs, wy), qz = __fmaf_rn(gsz, fs, wz);
        const float tx = (qx - ox) * ix, ty = (qy - oy) * iy, tz = (qz - oz) * iz;
        const float te = fminf(fminf(tx, ty), tz);
        if (!(te < tEnd)) return WALK_MISS;
        const uint32_t s = 1u << sh;
        uint32_t diff = 0;
        if (tx <= te) { diff |= mx ^ (mx + s); mx += s; wx = qx; bit += 1u; }
        if (ty <= te) { diff |= my ^ (my + s); my += s; wy = qy; bit += 4u; }
        if (tz <= te) { diff |= mz ^ (mz + s); mz += s; wz = qz; bit += 16u; }
        t = te;
        const int h = 31 - __clz(diff);
        if (h >= sh + 2) {
            // ---- left the current cell too: the new current cell is the smallest aligned one holding both positions.
            // (Leaving the root cell is impossible here: its exit planes lie at or beyond the grid's, so tEnd ended the walk.)
            sh = h & ~1;
            BLOKB_ASSERT(sh < 2 * W.levels);
            fs = pow2f(sh);
            const uint32_t keep = ~((1u << sh) - 1u);
            mx &= keep; my &= keep; mz &= keep;
            const int cap = 1 << (2 * W.levels);
            wx = float(gsx > 0.0f ? W.origin[0] + int(mx) : W.origin[0] + cap - int(mx));
            wy = float(gsy > 0.0f ? W.origin[1] + int(my) : W.origin[1] + cap - int(my));
            wz = float(gsz > 0.0f ? W.origin[2] + int(mz) : W.origin[2] + cap - int(mz));
            if (sh >= 4) {  // current cell is a level >= 3 cell
                const int k = (sh >> 1) + 1, s2 = sh + 2;
                const uint32_t cx = (mx ^ fx) >> s2, cy = (my ^ fy) >> s2, cz = (mz ^ fz) >> s2;
                BLOKB_ASSERT(cx < (uint32_t)W.denseDim[k][0] && cy < (uint32_t)W.denseDim[k][1] && cz < (uint32_t)W.denseDim[k][2]);
                mask = W.dense[k][(cz * uint32_t(W.denseDim[k][1]) + cy) * uint32_t(W.denseDim[k][0]) + cx];
            } else {
                mask = subMask;  // sh == 2: same sub-chunk, record still cached
            }
            bit = ((mx >> sh) & 3u) | (((my >> sh) & 3u) << 2) | (((mz >> sh) & 3u) << 4);
        }
        return WALK_CONTINUE;
    }
};

// Closest hit of one ray. Because the walk is strictly front to back, the first filled voxel is both the
// closest-hit and the any-hit (shadow) answer. Returns false on miss. `steps` counts DDA steps (profiling).
template <bool kStats>
BLOKB_HD bool traverse(const DeviceWorld& W, float ox, float oy, float oz, float dx, float dy, float dz,
                                         float tmin, float tmax, HitRec& out, uint32_t& steps) {
    Walker wk;
    if (!wk.init(W, ox, oy, oz, dx, dy, dz, tmin, tmax)) return false;
    for (;;) {
        const int r = wk.advance<kStats>(W, out, steps);
        if (r != WALK_CONTINUE) return r == WALK_HIT;
    }
}

}  // namespace blokb
