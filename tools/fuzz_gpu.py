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
    rounds += 1
    rays_total += 4 * n
    seed += 1
print(f"fuzz ok: {rounds} random worlds (seeds {seed0}..{seed - 1}), {rays_total} ray traces bit-exact against the checker "
      f"({type(O).__name__}), device-built == host-built on every world")
a.close(); b.close()
