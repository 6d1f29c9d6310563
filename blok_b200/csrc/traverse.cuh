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

#ifndef BLOKB_SELECT_V2
#define BLOKB_SELECT_V2 1
#endif
// Two north-star items were built as variants of descend(), measured slower on c3 and removed in round 3 (numbers and the code's
// shape are in DESIGN.md 5; the code is in history at 312663e): top levels staged in shared memory, ballot / shuffle node sharing.
#ifndef BLOKB_OCC_V2
#define BLOKB_OCC_V2 1
#endif

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
#if BLOKB_SELECT_V2
    // child index first, then one multiply-add each for the corner and the 6-bit index (IMAD runs on the FMA pipe; the ALU pipe —
    // compares, selects, logic — is the one this kernel saturates)
    const uint32_t c = (hi ? 2u : 0u) + (lo ? 1u : 0u);
    m = c * s + m;
    bit = c * (1u << axisShift) + bit;
#else
    if (hi) { m += 2u * s; bit += 2u << axisShift; }
    if (lo) { m += s; bit += 1u << axisShift; }
#endif
}

// Mirrored voxel coordinate m (and float(P(m)) of its near plane) of the voxel holding parameter t on one axis:
// the k with T(k) <= t < T(k + 1). The guess floor(o + d*t) is off by at most a cell or two near planes; the loops
// settle it with the exact slab values. [lo, hi] = the grid's extent in the mirrored frame (the clip in init()
// guarantees the answer lies inside it; the bounds only make termination unconditional).
BLOKB_HD void locate_axis(uint32_t& m, float& w, int g, int dim, int cap, bool pos, float o, float s, float inv, float t) {
    int x = __float2int_rd(o + s * t) - g;
    x = x < 0 ? 0 : (x > dim - 1 ? dim - 1 : x);
    int k = pos ? x : cap - 1 - x;
    const int lo = pos ? 0 : cap - dim, hi = pos ? dim - 1 : cap - 1;
    const int base = pos ? g : g + cap, sg = pos ? 1 : -1;  // P(k) = base + sg * k
    while (k < hi && plane_t(base + sg * (k + 1), o, inv) <= t) k++;
    while (k > lo && plane_t(base + sg * k, o, inv) > t) k--;
    m = uint32_t(k);
    w = float(base + sg * k);
}

// Per-ray stack of the nodes on the current root-to-cell path, one 16-byte entry per level (entry k-1 = the level-k
// cell). In kernels it lives in shared memory, laid out [level][thread] so that a warp's accesses are conflict-free
// 16-byte rows (addressed through a 32-bit shared-window address); the host emulation passes a plain array
// (stride 16 bytes). Written when descending, read when the walk leaves a cell and returns to an ancestor.
struct NodeStack {
#if defined(__CUDACC__)
    uint32_t base;    // shared-window byte address of this thread's level-1 entry
    uint32_t stride;  // bytes between consecutive levels
    BLOKB_HD void put(int level, uint32_t mlo, uint32_t mhi, uint32_t b, uint32_t aux) const {
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(base + uint32_t(level - 1) * stride), "r"(mlo), "r"(mhi), "r"(b), "r"(aux) : "memory");
    }
    BLOKB_HD uint4 get(int level) const {
        uint4 v;
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(base + uint32_t(level - 1) * stride) : "memory");
        return v;
    }
#else
    uint4* base;
    uint32_t stride;  // entries between consecutive levels
    BLOKB_HD void put(int level, uint32_t mlo, uint32_t mhi, uint32_t b, uint32_t aux) const { base[(level - 1) * stride] = make_uint4(mlo, mhi, b, aux); }
    BLOKB_HD uint4 get(int level) const { return base[(level - 1) * stride]; }
#endif
};

// Where a finished walk left the node stack, for the next ray of the same thread (sun shadow ray, bounce ray: both start one
// thousandth of a voxel off the voxel just hit): (x, y, z) = grid coordinates of the cell the walk ended in, level = the lowest stack
// entry that is still the ancestor chain of that cell (entries level..L are valid; 0 = nothing usable, start at the root). A ray
// whose start voxel shares its level-k ancestor with (x, y, z) begins in that ancestor — one LDS.128 — instead of descending from the
// root through L - k nodes it would find there anyway (shadow rays: 7.0 descents per ray, 4.5 of them this prefix).
struct WalkResume {
    uint32_t x = 0, y = 0, z = 0;
    int level = 0;
};

// One ray's walk as an explicit state machine: init(), then while the child holding the ray is occupied descend()
// (or report a hit at voxel level), else step(). traverse() below runs it to completion.
//
// Known resolution: (m, w) always describe the smallest cell, of size 1 << rs, that is KNOWN to hold the ray at
// parameter t. A step at cell size s leaves the position known at granularity s (the crossed axes sit exactly on
// the entered plane, the others are still inside the same s-slab), so when the walk then climbs to an ancestor and
// descends again through occupied cells — a ray skimming a surface from brick to brick — the children are read
// off m's bits; plane tests are spent only when descending BELOW the known resolution, and the float planes are
// recomputed only when a step is taken at a size ABOVE it.
struct Walker {
    float ox, oy, oz, ix, iy, iz;  // origin, 1 / substituted direction
    float gsx, gsy, gsz;           // +1 / -1: sign of the mirrored frame per axis
    uint32_t flip;                 // mirrored <-> actual child index XOR constant (3 | 12 | 48 per negative axis)
    float t, tEnd;
    uint32_t mx, my, mz;           // mirrored corner of the known cell (size 1 << rs)
    float wx, wy, wz;              // float(P) of its near planes
    int sh;                        // child cell size = 1 << sh; current cell = 4x that (level sh / 2 + 1)
    int rs;                        // known resolution (see above); rs <= sh whenever a child index is needed
    float fs;                      // float(1 << sh)
    uint32_t bit;                  // mirrored child index ix + 4 iy + 16 iz inside the current cell
    uint32_t mlo, mhi;             // current cell's node: occupancy of its 4^3 children (actual orientation),
    uint32_t base, aux;            //   first-child index, level-specific word (scene.h)

    BLOKB_HD void select() {
        const uint32_t s = 1u << sh;
        const float fs2 = fs + fs;
        bit = 0;
        select_child(mx, wx, bit, 0, s, fs, fs2, gsx, ox, ix, t);
        select_child(my, wy, bit, 2, s, fs, fs2, gsy, oy, iy, t);
        select_child(mz, wz, bit, 4, s, fs, fs2, gsz, oz, iz, t);
        rs = sh;
    }
    BLOKB_HD void bit_from_m() { bit = ((mx >> sh) & 3u) | (((my >> sh) & 3u) << 2) | (((mz >> sh) & 3u) << 4); }

    // Returns false when the ray misses the grid box altogether. `locate` picks how the start cell is found (below).
    BLOKB_HD bool init(const DeviceWorld& W, float ox_, float oy_, float oz_, float dx, float dy, float dz, float tmin, float tmax, bool locate,
                       const NodeStack* stk = nullptr, const WalkResume* from = nullptr) {
        ox = ox_; oy = oy_; oz = oz_;
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
        const int cap = 1 << (2 * L);
        flip = (px ? 0u : 3u) | (py ? 0u : 12u) | (pz ? 0u : 48u);
        gsx = px ? 1.0f : -1.0f; gsy = py ? 1.0f : -1.0f; gsz = pz ? 1.0f : -1.0f;
        sh = 2 * (L - 1);
        if (locate && from && from->level > 0) {
            // start voxel first, then the deepest cell on the previous walk's stack that holds it (see WalkResume)
            locate_axis(mx, wx, gx, W.dim[0], cap, px, ox, sx, ix, t);
            locate_axis(my, wy, gy, W.dim[1], cap, py, oy, sy, iy, t);
            locate_axis(mz, wz, gz, W.dim[2], cap, pz, oz, sz, iz, t);
            rs = 0;
            const uint32_t capm = uint32_t(cap - 1);
            const uint32_t ax = px ? mx : capm - mx, ay = py ? my : capm - my, az = pz ? mz : capm - mz;   // grid coordinates of the start voxel
            const uint32_t diff = (ax ^ from->x) | (ay ^ from->y) | (az ^ from->z) | 1u;
            int k = ((31 - __clz(diff)) >> 1) + 1;   // smallest level whose cell holds both voxels
            if (k < from->level) k = from->level;
            if (k < L) {
                sh = 2 * (k - 1);
                fs = pow2f(sh);
                const uint4 c = stk->get(k);
                mlo = c.x; mhi = c.y; base = c.z; aux = c.w;
                bit_from_m();
                return true;
            }
            fs = pow2f(sh);
            const uint4 r = __ldg(W.nodes);
            mlo = r.x; mhi = r.y; base = r.z; aux = r.w;
            bit_from_m();
            return true;
        }
        fs = pow2f(sh);
        const uint4 r = __ldg(W.nodes);
        mlo = r.x; mhi = r.y; base = r.z; aux = r.w;
        if (locate) {
            // Rays that start ON geometry (shadow, bounce): their first descent runs all the way to voxel level, so find
            // the start voxel per axis once — an arithmetic guess settled with the same plane tests every other decision
            // uses (T(m) <= t < T(m + 1) in the mirrored frame) — and let the whole descent read child indices off m.
            locate_axis(mx, wx, gx, W.dim[0], cap, px, ox, sx, ix, t);
            locate_axis(my, wy, gy, W.dim[1], cap, py, oy, sy, iy, t);
            locate_axis(mz, wz, gz, W.dim[2], cap, pz, oz, sz, iz, t);
            rs = 0;
            bit_from_m();
        } else {
            // Rays that enter the grid from outside (primary): the first descent stops at the first empty cell, usually a
            // coarse one, so refine level by level with plane tests only as deep as it goes.
            mx = my = mz = 0;
            wx = float(px ? gx : gx + cap); wy = float(py ? gy : gy + cap); wz = float(pz ? gz : gz + cap);
            select();
        }
        return true;
    }

    BLOKB_HD uint32_t abit() const { return bit ^ flip; }  // actual-orientation child index: x + 4y + 16z
    BLOKB_HD bool occupied() const {
        const uint32_t ab = abit();
#if BLOKB_OCC_V2
        return int32_t(((ab & 32u) ? mhi : mlo) << (~ab & 31u)) < 0;   // the child's bit moved into the sign position
#else
        return (((ab & 32u) ? mhi : mlo) >> (ab & 31u)) & 1u;
#endif
    }
    // number of occupied children before child `ab` = its index among the node's children
    BLOKB_HD uint32_t child_rank(uint32_t ab) const {
#if BLOKB_OCC_V2
        const uint32_t below = 0x7FFFFFFFu >> (~ab & 31u);   // (1 << n) - 1 from the shift amount occupied() already formed
#else
        const uint32_t below = (1u << (ab & 31u)) - 1u;
#endif
        return (ab & 32u) ? __popc(mlo) + __popc(mhi & below) : __popc(mlo & below);
    }

    // The occupied child holding the ray becomes the current cell (sh >= 2). Same code at every level.
    BLOKB_HD void descend(const DeviceWorld& W, const NodeStack& stk) {
        stk.put((sh >> 1) + 1, mlo, mhi, base, aux);
        const uint32_t idx = base + child_rank(abit());
        const uint4 c = __ldg(W.nodes + idx);
        mlo = c.x; mhi = c.y; base = c.z; aux = c.w;
        sh -= 2;
        fs *= 0.25f;
        if (sh < rs) select();  // below the known resolution: two plane tests per axis
        else bit_from_m();      // still inside the known cell: its corner's bits are the child indices
    }

    // Leave the empty child cell (size 1 << sh) through its nearest exit plane(s). Returns false when the ray ends.
    BLOKB_HD bool step(const DeviceWorld& W, const NodeStack& stk) {
        if (sh > rs) {
            // stepping at a size above the known resolution: widen the known cell to the child cell being left
            const uint32_t keep = ~((1u << sh) - 1u);
            mx &= keep; my &= keep; mz &= keep;
            const int cap = 1 << (2 * W.levels);
            wx = float(gsx > 0.0f ? W.origin[0] + int(mx) : W.origin[0] + cap - int(mx));
            wy = float(gsy > 0.0f ? W.origin[1] + int(my) : W.origin[1] + cap - int(my));
            wz = float(gsz > 0.0f ? W.origin[2] + int(mz) : W.origin[2] + cap - int(mz));
            rs = sh;
        }
        const float qx = __fmaf_rn(gsx, fs, wx), qy = __fmaf_rn(gsy, fs, wy), qz = __fmaf_rn(gsz, fs, wz);
        const float tx = (qx - ox) * ix, ty = (qy - oy) * iy, tz = (qz - oz) * iz;
        const float te = fminf(fminf(tx, ty), tz);
        if (!(te < tEnd)) return false;
        const uint32_t s = 1u << sh;
        uint32_t diff = 0;
        if (tx <= te) { diff |= mx ^ (mx + s); mx += s; wx = qx; bit += 1u; }
        if (ty <= te) { diff |= my ^ (my + s); my += s; wy = qy; bit += 4u; }
        if (tz <= te) { diff |= mz ^ (mz + s); mz += s; wz = qz; bit += 16u; }
        t = te;
        const int h = 31 - __clz(diff);
        if (h >= sh + 2) {
            // ---- left the current cell too: the new current cell is the smallest aligned one holding both positions.
            // (Leaving the root cell is impossible here: its exit planes lie at or beyond the grid's, so tEnd ended the
            // walk.) The position stays known at the size just stepped: rs is unchanged.
            sh = h & ~1;
            BLOKB_ASSERT(sh < 2 * W.levels);
            fs = pow2f(sh);
            const uint4 c = stk.get((sh >> 1) + 1);
            mlo = c.x; mhi = c.y; base = c.z; aux = c.w;
            bit_from_m();
        }
        return true;
    }

    // Material id of the filled voxel the walk stands on (sh == 0, occupied()): all a bounce ray needs of its hit.
    BLOKB_HD uint32_t hit_material(const DeviceWorld& W) const {
        const uint32_t mi = base + child_rank(abit());
        return W.matWide ? __ldg(reinterpret_cast<const uint32_t*>(W.voxelMat) + mi)
                         : (uint32_t)__ldg(reinterpret_cast<const uint16_t*>(W.voxelMat) + mi);
    }

    // The child holding the ray is a filled voxel (sh == 0, occupied(); then rs == 0 too): t is its slab entry (or
    // tmin / grid entry for the first cell). (dx, dy, dz) is the ray's direction as given (intersect.rint:138).
    BLOKB_HD void finish_hit(const DeviceWorld& W, const NodeStack& stk, float dx, float dy, float dz, HitRec& out) const {
        BLOKB_ASSERT(rs == 0 && sh == 0);
        const uint32_t ab = abit();
        const float hx = ox + dx * t, hy = oy + dy * t, hz = oz + dz * t;
        // voxel centre = near plane + half a voxel along the ray's sign = float(voxel coordinate) + 0.5 exactly
        const uint32_t face = hit_face(hx, hy, hz, wx + 0.5f * gsx, wy + 0.5f * gsy, wz + 0.5f * gsz);
        const uint64_t vm = uint64_t(mlo) | (uint64_t(mhi) << 32);
        const uint32_t b = ((ab >> 1) & 1u) | (((ab >> 3) & 1u) << 1) | (((ab >> 5) & 1u) << 2);  // 2^3 block inside the brick
        uint32_t rank = 0;
        for (uint32_t j = 0; j < 7; j++) rank += (j < b && (vm & block_bits(j))) ? 1u : 0u;
        const uint32_t leaf = __ldg(W.blockFirstLeaf + aux + rank) + ((ab & 1u) | (((ab >> 2) & 1u) << 1) | (((ab >> 4) & 1u) << 2));
        const uint32_t mi = base + child_rank(ab);
        const uint32_t mat = W.matWide ? __ldg(reinterpret_cast<const uint32_t*>(W.voxelMat) + mi)
                                       : (uint32_t)__ldg(reinterpret_cast<const uint16_t*>(W.voxelMat) + mi);
        const uint32_t prim = stk.get(2).w;  // the enclosing sub-chunk's node: aux = gl_PrimitiveID
        out.node = leaf; out.material = mat; out.t = t; out.facePrim = face | (prim << 3);
    }
};

// Closest hit of one ray. Because the walk is strictly front to back, the first filled voxel is both the
// closest-hit and the any-hit (shadow) answer. Returns false on miss. `steps` counts DDA steps (profiling).
// Loop shape: all lanes of a warp first finish descending, then step together, and the hit record is assembled
// after the loop — lanes re-converge at every step instead of drifting apart between descend and step phases.
// `resume` (optional, in and out): on entry, where this thread's previous walk left the stack (rays with locate = true start from
// there when they can); on return, where THIS walk left it — the hit voxel and level 2 after a hit, nothing after a miss.
template <bool kStats>
BLOKB_HD bool traverse(const DeviceWorld& W, const NodeStack& stk, float ox, float oy, float oz, float dx, float dy, float dz,
                       float tmin, float tmax, HitRec& out, uint32_t& steps, bool locate = true, WalkResume* resume = nullptr) {
    Walker wk;
#if defined(BLOKB_NO_RESUME)
    const bool inside = wk.init(W, ox, oy, oz, dx, dy, dz, tmin, tmax, locate);   // measurement variant: every walk starts at the root
#else
    const bool inside = wk.init(W, ox, oy, oz, dx, dy, dz, tmin, tmax, locate, &stk, resume);
#endif
    if (resume) resume->level = 0;
    if (!inside) return false;
    bool hit = false;
    for (;;) {
        while (wk.occupied()) {
            if (wk.sh == 0) { hit = true; break; }
            wk.descend(W, stk);
        }
        if (hit) break;
        if (kStats) steps++;
        if (!wk.step(W, stk)) return false;
    }
    wk.finish_hit(W, stk, dx, dy, dz, out);
    if (resume) {
        const uint32_t capm = (1u << (2 * W.levels)) - 1u;
        resume->x = (wk.flip & 3u) ? capm - wk.mx : wk.mx;
        resume->y = (wk.flip & 12u) ? capm - wk.my : wk.my;
        resume->z = (wk.flip & 48u) ? capm - wk.mz : wk.mz;
        resume->level = 2;   // the brick (level 1) sits in registers; entries 2..L are the hit voxel's ancestors
    }
    return true;
}

}  // namespace blokb
