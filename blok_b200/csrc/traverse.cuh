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

// Cell index along one axis at parameter t: the k in [lo, hi] with t(k) <= t < t(k+1) (mirrored for inv < 0).
BLOKB_HD int settle_axis(int g, int lo, int hi, int org, float o, float inv, float t) {
    g = g < lo ? lo : (g > hi ? hi : g);
    if (inv > 0.0f) {
        while (g < hi && plane_t(org + g + 1, o, inv) <= t) g++;
        while (g > lo && plane_t(org + g, o, inv) > t) g--;
    } else {
        while (g > lo && plane_t(org + g, o, inv) <= t) g--;
        while (g < hi && plane_t(org + g + 1, o, inv) > t) g++;
    }
    return g;
}

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

// Closest hit of one ray. Because the walk is strictly front to back, the first filled voxel is both the
// closest-hit and the any-hit (shadow) answer. Returns false on miss. `steps` counts DDA steps (profiling).
template <bool kStats>
BLOKB_HD bool traverse(const DeviceWorld& W, float ox, float oy, float oz, float dx, float dy, float dz,
                                         float tmin, float tmax, HitRec& out, uint32_t& steps) {
    const float sx = fabsf(dx) < 1e-6f ? 1e-6f : dx, sy = fabsf(dy) < 1e-6f ? 1e-6f : dy, sz = fabsf(dz) < 1e-6f ? 1e-6f : dz;
    const float ix = 1.0f / sx, iy = 1.0f / sy, iz = 1.0f / sz;
    const bool px = ix > 0.0f, py = iy > 0.0f, pz = iz > 0.0f;
    const int gx = W.origin[0], gy = W.origin[1], gz = W.origin[2];
    const int nx = W.dim[0], ny = W.dim[1], nz = W.dim[2];

    // clip to the grid box (same plane formula)
    float t, tEnd;
    {
        const float ax = plane_t(gx, ox, ix), bx = plane_t(gx + nx, ox, ix);
        const float ay = plane_t(gy, oy, iy), by = plane_t(gy + ny, oy, iy);
        const float az = plane_t(gz, oz, iz), bz = plane_t(gz + nz, oz, iz);
        const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
        const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
        t = fmaxf(tn, tmin);
        tEnd = fminf(tf, tmax);
        if (!(t < tEnd)) return false;
    }
    int x = settle_axis(__float2int_rd(ox + sx * t) - gx, 0, nx - 1, gx, ox, ix, t);
    int y = settle_axis(__float2int_rd(oy + sy * t) - gy, 0, ny - 1, gy, oy, iy, t);
    int z = settle_axis(__float2int_rd(oz + sz * t) - gz, 0, nz - 1, gz, oz, iz, t);

    int k = W.levels;
    uint64_t mask = W.dense[k][0];
    SubRec sub = {0ull, 0u, 0u};
    Brick brick = {0ull, 0u, 0u};
    uint32_t prim = 0;

    for (uint32_t it = 0; it < (1u << 20); it++) {
        // ---- descend while the child cell holding (x,y,z) is occupied
        int sh;
        uint32_t bit;
        for (;;) {
            sh = 2 * (k - 1);
            bit = ((x >> sh) & 3) | (((y >> sh) & 3) << 2) | (((z >> sh) & 3) << 4);
            if (!((mask >> bit) & 1ull)) break;
            k--;
            if (k == 0) {
                // ---- hit: voxel (x,y,z); t is its slab entry (or tmin / grid entry for the first cell)
                const float hx = ox + dx * t, hy = oy + dy * t, hz = oz + dz * t;  // intersect.rint:138
                const uint32_t face = hit_face(hx, hy, hz, float(gx + x) + 0.5f, float(gy + y) + 0.5f, float(gz + z) + 0.5f);
                const uint64_t vm = brick.voxelMask;
                const uint32_t b = ((x >> 1) & 1) | (((y >> 1) & 1) << 1) | (((z >> 1) & 1) << 2);
                uint32_t rank = 0;
                for (uint32_t j = 0; j < 7; j++) rank += (j < b && (vm & block_bits(j))) ? 1u : 0u;
                const uint32_t leaf = W.blockFirstLeaf[brick.blockBase + rank] + ((x & 1) | ((y & 1) << 1) | ((z & 1) << 2));
                const uint32_t mi = brick.matBase + __popcll(vm & ((1ull << bit) - 1ull));
                const uint32_t mat = W.matWide ? reinterpret_cast<const uint32_t*>(W.voxelMat)[mi]
                                               : (uint32_t) reinterpret_cast<const uint16_t*>(W.voxelMat)[mi];
                out.node = leaf; out.material = mat; out.t = t; out.facePrim = face | (prim << 3);
                return true;
            }
            if (k >= 3) {
                const int s2 = 2 * k;
                mask = W.dense[k][(size_t(z >> s2) * W.denseDim[k][1] + (y >> s2)) * W.denseDim[k][0] + (x >> s2)];
            } else if (k == 2) {
                prim = W.topGrid[(size_t(z >> 4) * (ny >> 4) + (y >> 4)) * (nx >> 4) + (x >> 4)];
                sub = W.subs[prim];
                mask = sub.brickMask;
            } else {
                brick = W.bricks[sub.brickBase + __popcll(sub.brickMask & ((1ull << bit) - 1ull))];
                mask = brick.voxelMask;
            }
        }
        // ---- the level-(k-1) cell of size s holding (x,y,z) is empty: leave it through its nearest exit plane(s)
        if (kStats) steps++;
        const int s = 1 << sh;
        const int cx = x & ~(s - 1), cy = y & ~(s - 1), cz = z & ~(s - 1);
        const float tx = plane_t(gx + cx + (px ? s : 0), ox, ix);
        const float ty = plane_t(gy + cy + (py ? s : 0), oy, iy);
        const float tz = plane_t(gz + cz + (pz ? s : 0), oz, iz);
        const float te = fminf(fminf(tx, ty), tz);
        if (!(te < tEnd)) return false;
        int qx, qy, qz;
        if (tx <= te) qx = px ? cx + s : cx - 1;
        else qx = (s == 1) ? x : settle_axis(__float2int_rd(ox + sx * te) - gx, cx, cx + s - 1, gx, ox, ix, te);
        if (ty <= te) qy = py ? cy + s : cy - 1;
        else qy = (s == 1) ? y : settle_axis(__float2int_rd(oy + sy * te) - gy, cy, cy + s - 1, gy, oy, iy, te);
        if (tz <= te) qz = pz ? cz + s : cz - 1;
        else qz = (s == 1) ? z : settle_axis(__float2int_rd(oz + sz * te) - gz, cz, cz + s - 1, gz, oz, iz, te);
        if ((unsigned)qx >= (unsigned)nx || (unsigned)qy >= (unsigned)ny || (unsigned)qz >= (unsigned)nz) return false;
        const uint32_t diff = uint32_t((x ^ qx) | (y ^ qy) | (z ^ qz));
        x = qx; y = qy; z = qz; t = te;
        // smallest aligned cell holding both the old and the new position
        const int lev = ((31 - __clz(diff)) >> 1) + 1;
        if (lev > k) {
            k = lev;
            if (k >= 3) {
                const int s2 = 2 * k;
                mask = W.dense[k][(size_t(z >> s2) * W.denseDim[k][1] + (y >> s2)) * W.denseDim[k][0] + (x >> s2)];
            } else {
                mask = sub.brickMask;  // k == 2: same sub-chunk, record still cached
            }
        }
    }
    return false;
}

}  // namespace blokb
