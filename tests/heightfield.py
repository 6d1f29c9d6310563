"""Size-independent property check for hit records on heightfield-shell scenes (scenes.terrain). TEST INFRASTRUCTURE.

At BASELINE configs[4] size (8192^3, 201 M voxels) the reference-format scene the oracle needs is ~6 GB and takes the
reference's single-threaded builder minutes, so full-size parity there is checked through what the domain offers instead:
the scene is a function (voxel (x, y, z) is filled iff h[z, x] - thickness < y <= h[z, x]), hence for any ray

  * a reported hit must lie on a face of a FILLED voxel, with that column's material, and
  * every voxel the ray crosses before the reported t (or before leaving the grid, for a reported miss) must be EMPTY.

The marcher below is a plain float64 Amanatides-Woo walk over unit voxels with the reference's direction substitution
(intersect.rint:79). It shares no code and no data structure with the CUDA traversal or the oracle; voxels the ray
crosses for less than `eps` of parameter are ignored, which makes the check independent of fp32-vs-fp64 rounding at
grazing corners (the bit-exact comparison against the oracle covers those at sizes the oracle handles).
tests/test_heightfield_check.py pins this checker against the oracle's hits at test scale.
"""
import numpy as np


def filled(h, thickness, x, y, z):
    N = h.shape[0]
    inside = (x >= 0) & (x < N) & (z >= 0) & (z < N)
    hh = h[np.clip(z, 0, N - 1), np.clip(x, 0, N - 1)]
    return inside & (y <= hh) & (y > hh - thickness)


def column_material(h, x, z):
    N = h.shape[0]
    hh = h[z, x].astype(np.int64)
    return (((hh - N // 4) * 8 // max(N // 8, 1)).clip(0, 7) + 1).astype(np.uint32)


def _subst_dir(d):
    d = d.astype(np.float64).copy()
    small = np.abs(d) < 1e-6
    d[small] = 1e-6
    return d


def check_hits(h, thickness, rays, node, material, t, face, eps=1e-3, max_steps=200000):
    """rays (n, 8) f32 {o, d, tmin, tmax}; node/material/t/face: the hit records under test (node 0xFFFFFFFF = miss).
    Returns counts of violated properties (all zero = pass) plus bookkeeping."""
    n = len(rays)
    o = rays[:, 0:3].astype(np.float64)
    dg = rays[:, 3:6].astype(np.float64)          # direction as given (hit position)
    d = _subst_dir(rays[:, 3:6])                  # substituted direction (slabs)
    inv = 1.0 / d
    N = h.shape[0]
    hit = node != 0xFFFFFFFF
    tg = t.astype(np.float64)

    # ---- 1. the reported hit lies on a face of a filled voxel of the right material
    p = o + dg * tg[:, None]
    axis = (face >> 1).astype(np.int64)
    positive = (face & 1) == 0                     # faces 0/2/4 = +x/+y/+z normals: the hit plane is the voxel's far side
    ok_hit = np.zeros(n, bool)
    ok_mat = np.zeros(n, bool)
    on_plane = np.zeros(n, bool)
    rows = np.arange(n)
    pa = p[rows, np.where(hit, axis, 0)]
    on_plane = np.abs(pa - np.rint(pa)) < 1e-2 * np.maximum(1.0, np.abs(tg) * 1e-4)
    va = np.rint(pa).astype(np.int64) - np.where(positive, 1, 0)
    nudge = np.maximum(1e-4, 4e-7 * np.abs(tg))     # fp32 t carries ~1e-7 relative error: the in-plane position is known to this
    for k in range(8):
        sgn = np.array([1.0 if k & 1 else -1.0, 1.0 if k & 2 else -1.0, 1.0 if k & 4 else -1.0])
        v = np.floor(p + sgn[None, :] * nudge[:, None]).astype(np.int64)  # candidates for the two in-plane axes (edge / corner hits)
        for a in range(3):
            sel = hit & (axis == a)
            v[sel, a] = va[sel]
        f = filled(h, thickness, v[:, 0], v[:, 1], v[:, 2]) & hit
        m = f & (column_material(h, np.clip(v[:, 0], 0, N - 1), np.clip(v[:, 2], 0, N - 1)) == material)
        ok_hit |= f
        ok_mat |= m
    bad_hit_voxel = int(np.count_nonzero(hit & ~(ok_hit & on_plane)))
    bad_material = int(np.count_nonzero(hit & ok_hit & ~ok_mat))

    # ---- 2. nothing filled before the reported t / before the ray leaves the grid
    ymin, ymax = int(h.min()) - thickness + 1, int(h.max())
    lo = np.array([0.0, float(ymin), 0.0])
    hi = np.array([float(N), float(ymax + 1), float(N)])
    ta, tb = (lo - o) * inv, (hi - o) * inv
    tn = np.maximum(np.minimum(ta, tb).max(1), rays[:, 6].astype(np.float64))
    tf = np.minimum(np.maximum(ta, tb).min(1), rays[:, 7].astype(np.float64))
    limit = np.where(hit, np.minimum(tg - eps, tf), tf)
    tc = tn.copy()
    v = np.floor(o + d * (tc + 1e-7)[:, None]).astype(np.int64)
    step = np.where(d > 0, 1, -1).astype(np.int64)
    nxt = np.where(d > 0, 1, 0).astype(np.int64)
    active = tc < limit
    blocked = np.zeros(n, bool)
    steps = 0
    while active.any() and steps < max_steps:
        idx = np.nonzero(active)[0]
        vi = v[idx]
        tmax = ((vi + nxt[idx]) - o[idx]) * inv[idx]
        te = tmax.min(1)
        span = np.minimum(te, limit[idx]) - tc[idx]
        f = filled(h, thickness, vi[:, 0], vi[:, 1], vi[:, 2]) & (span > eps)
        blocked[idx[f]] = True
        ax = tmax.argmin(1)
        v[idx, ax] += step[idx, ax]
        tc[idx] = te
        inb = (v[idx, 0] >= 0) & (v[idx, 0] < N) & (v[idx, 2] >= 0) & (v[idx, 2] < N) & (v[idx, 1] >= ymin - 1) & (v[idx, 1] <= ymax + 1)
        active[idx] = (~f) & (te < limit[idx]) & inb
        steps += 1
    return {"rays": n, "hits": int(hit.sum()), "bad_hit_voxel": bad_hit_voxel, "bad_material": bad_material,
            "filled_before_hit": int(np.count_nonzero(blocked & hit)), "filled_on_missed_ray": int(np.count_nonzero(blocked & ~hit)),
            "unfinished": int(active.sum()), "march_steps": steps, "eps": eps}


def passed(res) -> bool:
    return not (res["bad_hit_voxel"] or res["bad_material"] or res["filled_before_hit"] or res["filled_on_missed_ray"] or res["unfinished"])
