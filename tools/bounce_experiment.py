"""Experiment: how much does ray ordering buy for incoherent bounce rays? Times blokb_trace_rays (1 thread per ray,
consecutive rays per warp) on the c3 scene for: primary rays in row-major / tile order, sun-shadow rays, cosine-sampled
bounce rays in pixel order, sorted by direction octant, and by octant + Morton code of the origin."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import blok_b200  # noqa: E402
from blok_b200 import IMG_HIT, workloads  # noqa: E402

name = "c3"
wl = workloads.WORKLOADS[name]
W, H = wl["width"], wl["height"]
world = workloads.build_world(name)
r = blok_b200.Renderer(W, H, 0)
r.add_world(world)
f = workloads.camera_for(name).frame(W, H, 0, 1)
r.trace_primary(f)
hits = r.readback(IMG_HIT).reshape(-1)

# primary rays on the host (float32 arithmetic close enough for a timing experiment)
invV = np.array(f.invView, np.float32).reshape(4, 4).T
invP = np.array(f.invProj, np.float32).reshape(4, 4).T
ys, xs = np.mgrid[0:H, 0:W]
dx = ((xs + 0.5) / W * 2 - 1).astype(np.float32).ravel()
dy = ((ys + 0.5) / H * 2 - 1).astype(np.float32).ravel()
tg = (invP @ np.stack([dx, dy, np.ones_like(dx), np.ones_like(dx)])).T[:, :3]
tg /= np.linalg.norm(tg, axis=1, keepdims=True)
wd = (invV[:3, :3] @ tg.T).T
wd /= np.linalg.norm(wd, axis=1, keepdims=True)
cam = np.array(f.camPos, np.float32)
n = W * H
prim = np.concatenate([np.tile(cam, (n, 1)), wd, np.full((n, 1), 0.001), np.full((n, 1), 10000.0)], 1).astype(np.float32)


STREAM = torch.cuda.Stream()


def timed(rays, label, reps=5, reorder=False):
    d = torch.from_numpy(np.ascontiguousarray(rays)).cuda()
    out = torch.empty((len(rays), 4), dtype=torch.int32, device="cuda")
    st = STREAM.cuda_stream
    for _ in range(2):
        r.trace_rays_device(d.data_ptr(), len(rays), out.data_ptr(), stream=st, reorder=reorder)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(STREAM)
    for _ in range(reps):
        r.trace_rays_device(d.data_ptr(), len(rays), out.data_ptr(), stream=st, reorder=reorder)
    b.record(STREAM)
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    r.stats_enable(True)
    r.trace_rays_device(d.data_ptr(), len(rays), out.data_ptr(), stream=st, reorder=reorder)
    torch.cuda.synchronize()
    steps = r.stats_read()[0]
    r.stats_enable(False)
    hitfrac = float((out[:, 0] != -1).float().mean())
    print(f"{label:50s} {len(rays):9d} rays  {ms:8.3f} ms  {len(rays) / ms / 1e3:8.1f} Mrays/s  steps/ray {steps / len(rays):6.2f}  hit {hitfrac:.3f}", flush=True)


def tile_order(w, h, tw=8, th=4):
    idx = np.arange(w * h).reshape(h, w)
    return idx.reshape(h // th, th, w // tw, tw).transpose(0, 2, 1, 3).reshape(-1)


def morton3(x, y, z):
    def part(v):
        v = v.astype(np.uint64) & 0x1FFFFF
        v = (v | (v << 32)) & 0x1F00000000FFFF
        v = (v | (v << 16)) & 0x1F0000FF0000FF
        v = (v | (v << 8)) & 0x100F00F00F00F00F
        v = (v | (v << 4)) & 0x10C30C30C30C30C3
        v = (v | (v << 2)) & 0x1249249249249249
        return v
    return part(x) | (part(y) << 1) | (part(z) << 2)


timed(prim, "primary, row-major")
to = tile_order(W, H)
timed(prim[to], "primary, 8x4 tile order")

hit = hits["node"] != 0xFFFFFFFF
normals = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], np.float32)
N = normals[np.where(hit, hits["facePrim"] & 7, 0)]
flip = (N * wd).sum(1) > 0
N = np.where(flip[:, None], -N, N)
pos = cam + wd * hits["t"][:, None]
sun = np.array([0.5, 0.8, 0.3], np.float32)
sun /= np.linalg.norm(sun)
lit = hit & ((N * sun).sum(1) > 0)
sel = to[lit[to]]
sh = np.concatenate([pos[sel] + N[sel] * 0.001, np.tile(sun, (len(sel), 1)), np.full((len(sel), 1), 0.001), np.full((len(sel), 1), 1000.0)], 1).astype(np.float32)
timed(sh, "shadow rays, compacted, tile order")

rng = np.random.default_rng(1)
sel = to[hit[to]]
u1, u2 = rng.random(len(sel), dtype=np.float32), rng.random(len(sel), dtype=np.float32)
rr, phi = np.sqrt(u1), 2 * np.pi * u2
lx, ly, lz = rr * np.cos(phi), rr * np.sin(phi), np.sqrt(np.maximum(0, 1 - u1))
Ns = N[sel]
up = np.where((np.abs(Ns[:, 2]) < 0.999)[:, None], np.array([0, 0, 1], np.float32), np.array([1, 0, 0], np.float32))
T = np.cross(up, Ns)
T /= np.linalg.norm(T, axis=1, keepdims=True)
B = np.cross(Ns, T)
bd = T * lx[:, None] + B * ly[:, None] + Ns * lz[:, None]
bd /= np.linalg.norm(bd, axis=1, keepdims=True)
bo = pos[sel] + Ns * 0.002
bounce = np.concatenate([bo, bd, np.full((len(sel), 1), 0.001), np.full((len(sel), 1), 10000.0)], 1).astype(np.float32)
timed(bounce, "bounce rays, compacted, tile order")
timed(bounce, "  same list through blokb_trace_rays_reordered (incl. sort)", reorder=True)
timed(sh, "  shadow rays through blokb_trace_rays_reordered (incl. sort)", reorder=True)
octant = (bd[:, 0] < 0).astype(np.int64) | ((bd[:, 1] < 0).astype(np.int64) << 1) | ((bd[:, 2] < 0).astype(np.int64) << 2)
o1 = np.argsort(octant, kind="stable")
timed(bounce[o1], "bounce rays, sorted by octant (stable)")
mc = morton3((bo[:, 0]).astype(np.int64) >> 2, (bo[:, 1]).astype(np.int64) >> 2, (bo[:, 2]).astype(np.int64) >> 2)
key = (octant.astype(np.uint64) << np.uint64(60)) | mc
o2 = np.argsort(key, kind="stable")
timed(bounce[o2], "bounce rays, sorted by octant + Morton(origin/4)")
# direction-cell binning: octant + quantised direction (8x8 over the two minor components) + morton
q = np.clip(((bd + 1) * 2).astype(np.int64), 0, 3)
dkey = (q[:, 0] | (q[:, 1] << 2) | (q[:, 2] << 4)).astype(np.uint64)
key2 = (dkey << np.uint64(58)) | mc
o3 = np.argsort(key2, kind="stable")
timed(bounce[o3], "bounce rays, sorted by 4^3 direction cell + Morton")
for nb in (4, 8):  # direction cell alone, stable: inside a cell the list keeps its (tile = spatial) order
    qq = np.clip(((bd + 1) * (nb / 2)).astype(np.int64), 0, nb - 1)
    dk = (qq[:, 0] + nb * (qq[:, 1] + nb * qq[:, 2])).astype(np.uint64)
    timed(bounce[np.argsort(dk, kind="stable")], f"bounce rays, {nb}^3 direction cell only (stable)")
    mc5 = morton3((bo[:, 0]).astype(np.int64) >> 6, (bo[:, 1]).astype(np.int64) >> 6, (bo[:, 2]).astype(np.int64) >> 6)
    timed(bounce[np.lexsort((mc5, dk))], f"bounce rays, {nb}^3 direction cell + Morton(origin/64)")
for nb in (8, 16):
    qq = np.clip(((bd + 1) * (nb / 2)).astype(np.int64), 0, nb - 1)
    dk = (qq[:, 0] + nb * (qq[:, 1] + nb * qq[:, 2])).astype(np.uint64)
    for shift, lab in ((2, "origin/4"), (4, "origin/16"), (0, "origin/1")):
        mc2 = morton3((bo[:, 0]).astype(np.int64) >> shift, (bo[:, 1]).astype(np.int64) >> shift, (bo[:, 2]).astype(np.int64) >> shift)
        oo = np.lexsort((mc2, dk))
        timed(bounce[oo], f"bounce rays, {nb}^3 direction cell + Morton({lab})")
# Morton first (spatial blocks of 16^3), direction cell inside
mc16 = morton3((bo[:, 0]).astype(np.int64) >> 4, (bo[:, 1]).astype(np.int64) >> 4, (bo[:, 2]).astype(np.int64) >> 4)
oo = np.lexsort((dkey, mc16))
timed(bounce[oo], "bounce rays, Morton(origin/16) then 4^3 direction cell")
mc64 = morton3((bo[:, 0]).astype(np.int64) >> 6, (bo[:, 1]).astype(np.int64) >> 6, (bo[:, 2]).astype(np.int64) >> 6)
oo = np.lexsort((mc, dkey, mc64))
timed(bounce[oo], "bounce rays, Morton(origin/64), 4^3 dir cell, Morton(origin/4)")
up_frac = (bd[:, 1] > 0).mean()
print("bounce rays going up:", up_frac)
r.close()
