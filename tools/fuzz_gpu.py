"""GPU fuzz: random worlds x random + degenerate rays, CUDA path (through the C-ABI) vs the CPU checker, bit for bit.
  python tools/fuzz_gpu.py [seconds=60] [seed=1]
Per round: a random voxel world (clustered blobs + noise, random offset incl. negative coordinates, 1-40 k voxels, material ids
up to 70 000 now and then), uploaded twice — reference-format arrays through blokb_upload_world, voxel list through
blokb_upload_voxels — whose compact arrays must be byte-identical; then ~200 k rays (tests/conftest.make_rays: random,
axis-aligned, near-zero and zero components, lattice-plane origins; a third starting inside the grid) traced by the ray-list entry
points (plain and reordered) and by the checker. Exits non-zero on the first mismatch and prints the seed that reproduces it.
The checker is oracle/_ref (the reference's sources) when present, else the restatement — like bench.py's cpu leg, this tool is
checker-side infrastructure, not part of the product path."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import blok_b200  # noqa: E402
from blok_b200 import WORLD_BLOCKFIRSTLEAF, WORLD_NODES, WORLD_VOXELMAT, WorldSvoGpu, scenes  # noqa: E402
from oracle import pyoracle  # noqa: E402
from tests.conftest import make_rays, to_scene  # noqa: E402

from bench import shadow_rays  # noqa: E402  (raygen.rgen:280-298 restated in numpy for the checker side)


def _lit_mask(rays, h, t):
    """Which primary hits cast a sun shadow ray (the selection inside bench.shadow_rays, exposed as a mask)."""
    normals = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], np.float32)
    hit = t >= 0
    N = normals[np.where(hit, h[:, 3], 0)]
    d = rays[:, 3:6]
    flip = ((N[:, 0] * d[:, 0] + N[:, 1] * d[:, 1]) + N[:, 2] * d[:, 2]) > 0
    N = np.where(flip[:, None], -N, N)
    s = np.array([0.5, 0.8, 0.3], np.float32)
    sun = (s * (np.float32(1.0) / np.sqrt((s[0] * s[0] + s[1] * s[1]) + s[2] * s[2], dtype=np.float32))).astype(np.float32)
    return ((N[:, 0] * sun[0] + N[:, 1] * sun[1]) + N[:, 2] * sun[2]) > 0


FRAME = {}
frames_total = 0
renders_total = 0
budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1
O = pyoracle.RefLib() if os.path.exists(pyoracle.REF_SO) else pyoracle.OracleLib()
mats = scenes.default_materials(16)
a, b = blok_b200.Renderer(64, 64, 0), blok_b200.Renderer(64, 64, 0)
t_end = time.time() + budget
rounds = rays_total = 0
seed = seed0
while time.time() < t_end:
    rng = np.random.default_rng(seed)
    extent = int(rng.choice([20, 60, 130, 300, 700]))
    off = rng.integers(-400, 400, 3)
    nblobs = int(rng.integers(1, 6))
    pts = []
    for _ in range(nblobs):
        c = rng.uniform(0, extent, 3)
        rad = rng.uniform(2, max(3, extent / 6))
        k = int(rng.integers(200, 8000))
        p = c + rng.normal(size=(k, 3)) * rad / 2
        pts.append(p)
    pts.append(rng.uniform(0, extent, (int(rng.integers(0, 3000)), 3)))
    xyz = (np.floor(np.concatenate(pts)).astype(np.int64).clip(0, extent - 1) + off).astype(np.int32)
    mid = rng.integers(1, 17, len(xyz)).astype(np.uint32)
    if rng.random() < 0.2:
        mid = (mid + 70000).astype(np.uint32)
    world = WorldSvoGpu.from_voxels(xyz, mid, mats)
    a.add_world(world)
    b.add_voxels(xyz, mid, mats)
    for k in (WORLD_NODES, WORLD_BLOCKFIRSTLEAF, WORLD_VOXELMAT):
        if a.world_export(k).tobytes() != b.world_export(k).tobytes():
            print(f"MISMATCH seed {seed}: device-built world differs from the host path in array {k}")
            sys.exit(1)
    lo, hi = float(off.min()) - 30, float(off.max()) + extent + 30
    n = 200_000
    rays = make_rays(rng, n, lo, hi)
    k = n // 3
    rays[:k, 0:3] = (off + rng.uniform(0, extent, (k, 3))).astype(np.float32)
    rays[: k // 2, 0:3] = np.round(rays[: k // 2, 0:3])
    h, t, _ = O.trace(to_scene(world), rays, threads=os.cpu_count() or 1)
    hit = t >= 0
    import torch
    d = torch.from_numpy(rays).cuda()
    for r, label in ((a, "host-built world"), (b, "device-built world")):
        for reorder in (False, True):
            o = torch.empty((n, 4), dtype=torch.int32, device="cuda")
            r.trace_rays_device(d.data_ptr(), n, o.data_ptr(), reorder=reorder)
            r.sync()
            out = o.cpu().numpy().view(blok_b200.HIT_DT).reshape(-1)
            mm = [int((out["node"] != h[:, 1]).sum()), int((out["t"].view(np.uint32) != t.view(np.uint32)).sum()),
                  int((out["material"] != h[:, 2]).sum()), int(((out["facePrim"] >> 3)[hit] != h[hit, 0]).sum()),
                  int(((out["facePrim"] & 7)[hit] != h[hit, 3]).sum())]
            if any(mm):
                print(f"MISMATCH seed {seed} ({label}, reorder={reorder}): node/t/material/prim/face = {mm}; voxels {len(xyz)} extent {extent} offset {off}")
                sys.exit(1)
    # ---- frame kernels: a random camera, primary + sun-shadow pass (k_primary<true>) vs the checker's primary rays and shadow rays
    centre = off + extent / 2.0
    dist = rng.uniform(0.2, 2.5) * extent
    ang, elev = rng.uniform(0, 2 * np.pi), rng.uniform(-0.3, 1.2)
    pos = centre + dist * np.array([np.cos(ang) * np.cos(elev), np.sin(elev), np.sin(ang) * np.cos(elev)])
    look = centre + rng.normal(size=3) * extent * 0.2 - pos
    yaw = float(np.degrees(np.arctan2(look[2], look[0])))
    pitch = float(np.clip(np.degrees(np.arctan2(look[1], np.hypot(look[0], look[2]))), -89, 89))
    cam = blok_b200.Camera(tuple(float(v) for v in pos), yaw, pitch, fov=float(rng.choice([30.0, 60.0, 90.0, 110.0])))
    fw, fh = (int(v) for v in rng.choice([[64, 64], [200, 120], [133, 77], [256, 144]]))
    ubo = cam.frame(fw, fh, int(rng.integers(0, 1000)), 1)
    prays = O.primary_rays(ubo.to_numpy())
    ph, pt, _ = O.trace(to_scene(world), prays, threads=os.cpu_count() or 1)
    srays = shadow_rays(prays, ph, pt)
    lit_idx = np.nonzero((pt >= 0) & _lit_mask(prays, ph, pt))[0]
    _, st, _ = O.trace(to_scene(world), srays, any_hit=True, threads=os.cpu_count() or 1) if len(srays) else (None, np.zeros(0, np.float32), None)
    want_shadow = np.zeros(fw * fh, np.uint8)
    want_shadow[lit_idx] = np.where(st >= 0, 2, 1)
    for r, label in ((a, "host-built world"), (b, "device-built world")):
        if (label, fw, fh) not in FRAME:
            FRAME[(label, fw, fh)] = blok_b200.Renderer(fw, fh, 0)
        r2 = FRAME[(label, fw, fh)]
        (r2.add_world(world) if label.startswith("host") else r2.add_voxels(xyz, mid, mats))
        r2.trace_primary(ubo, shadow=True)
        out = r2.readback(blok_b200.IMG_HIT).reshape(-1)
        sh = r2.readback(blok_b200.IMG_SHADOW).reshape(-1)
        phit = pt >= 0
        mm = [int((out["node"] != ph[:, 1]).sum()), int((out["t"].view(np.uint32) != pt.view(np.uint32)).sum()),
              int((out["material"] != ph[:, 2]).sum()), int(((out["facePrim"] & 7)[phit] != ph[phit, 3]).sum()), int((sh != want_shadow).sum())]
        if any(mm):
            print(f"MISMATCH seed {seed} frame ({label}): node/t/material/face/shadow = {mm}; camera {pos} yaw {yaw} pitch {pitch} {fw}x{fh}")
            sys.exit(1)
    frames_total += 2
    # ---- every 8th world: the full raygen (blokb_render = k_gbuffer + k_path<kColor>) against the checker's, 64x64, 4 spp
    if rounds % 8 == 0:
        prev_cam = blok_b200.Camera(tuple(float(v) + 0.37 for v in pos), yaw + 1.5, pitch - 0.5, fov=cam.fov)
        ubo_r = cam.frame(64, 64, int(rng.integers(0, 1000)), 4, prev=prev_cam)
        ref = O.render(to_scene(world), ubo_r.to_numpy(), threads=os.cpu_count() or 1)
        rr = FRAME.get(("render",)) or FRAME.setdefault(("render",), blok_b200.Renderer(64, 64, 0))
        rr.add_voxels(xyz, mid, mats)
        rr.render(ubo_r)
        wp = rr.readback(blok_b200.IMG_WORLDPOS)
        nr = rr.readback(blok_b200.IMG_NORMALROUGH).astype(np.float32)
        am = rr.readback(blok_b200.IMG_ALBEDOMETAL).astype(np.int32)
        mv = rr.readback(blok_b200.IMG_MOTION).astype(np.float32)
        col = rr.readback(blok_b200.IMG_COLOR)
        rel = np.abs(col[..., :3] - ref["color"][..., :3]) / (np.abs(ref["color"][..., :3]) + 1e-3)
        problems = []
        in_table = int(mid.max()) < len(mats)   # ids beyond the material table are undefined in the reference (its checker reads past the array)
        if wp.tobytes() != ref["worldPos"].tobytes(): problems.append("worldPos bits")
        if not np.array_equal(nr, ref["normalRough"].astype(np.float16).astype(np.float32)): problems.append("normalRough")
        if np.abs(am - np.rint(np.clip(ref["albedoMetal"], 0, 1) * 255).astype(np.int32)).max() > 1: problems.append("albedoMetal")
        with np.errstate(over="ignore", invalid="ignore"):
            mref = ref["motion"][..., :2].astype(np.float16).astype(np.float32)   # the image is rg16f: compare at its precision
            mbad = ~np.isclose(mv, mref, rtol=4e-3, atol=2e-3, equal_nan=True)
        if mbad.any():
            i = tuple(np.argwhere(mbad)[0])
            problems.append(f"motion ({int(mbad.sum())} texels, first {i}: got {mv[i]} want {mref[i]})")
        if np.mean(rel.max(-1) < 2e-3) < 0.99: problems.append(f"radiance ({np.mean(rel.max(-1) < 2e-3):.4f} of pixels within 2e-3)")
        if problems and in_table:
            print(f"MISMATCH seed {seed} render: {problems}; camera {pos} yaw {yaw} pitch {pitch}")
            sys.exit(1)
        renders_total += 1
    rounds += 1
    rays_total += 4 * n
    seed += 1
print(f"fuzz ok: {rounds} random worlds (seeds {seed0}..{seed - 1}), {rays_total} ray traces and {frames_total} primary+shadow frames bit-exact against the checker, {renders_total} full-raygen frames (G-buffer exact, radiance within tolerance) "
      f"({type(O).__name__}), device-built == host-built on every world")
a.close(); b.close()
