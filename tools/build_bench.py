"""Scene upload, host path vs device path, on one workload's voxel set.
  python tools/build_bench.py [workload=c3]
host path   = blokb_worldgen_pack (reference builder restated on the CPU) + blokb_upload_world (compact build on the CPU, H2D)
device path = blokb_upload_voxels (H2D of the voxel list, everything else on the GPU)
Prints timings and checks that both leave byte-identical compact arrays on the device."""
import hashlib
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import blok_b200  # noqa: E402
from blok_b200 import WORLD_BLOCKFIRSTLEAF, WORLD_NODES, WORLD_VOXELMAT, WorldSvoGpu, scenes, workloads  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
wl = workloads.WORKLOADS[name]
t0 = time.perf_counter()
xyz, mid, mats = getattr(scenes, wl["scene"])(wl["N"])
t_gen = time.perf_counter() - t0
out = {"workload": name, "voxels": int(len(mid)), "generate_s": round(t_gen, 3)}

b = blok_b200.Renderer(64, 64, 0)
b.add_voxels(xyz[:1000], mid[:1000], mats)  # warm-up: context, cub temp sizing paths
ts = []
for _ in range(3):
    t0 = time.perf_counter()
    b.add_voxels(xyz, mid, mats)
    ts.append(time.perf_counter() - t0)
sb = b.world_stats()
out["device"] = {"wall_ms": round(min(ts) * 1e3, 2), "buildMs": sb["buildMs"], "uploadMs": sb["uploadMs"], "input_MB": round((xyz.nbytes + mid.nbytes) / 1e6, 1)}

a = blok_b200.Renderer(64, 64, 0)
t0 = time.perf_counter()
world = WorldSvoGpu.from_voxels(xyz, mid, mats)
t_pack = time.perf_counter() - t0
t0 = time.perf_counter()
a.add_world(world)
t_up = time.perf_counter() - t0
sa = a.world_stats()
out["host"] = {"worldgen_pack_ms": round(t_pack * 1e3, 1), "upload_world_ms": round(t_up * 1e3, 1), "buildMs": sa["buildMs"], "uploadMs": sa["uploadMs"],
               "input_MB": round((world.globalNodes.nbytes + world.globalSubChunks.nbytes) / 1e6, 1)}
same = True
for k in (WORLD_NODES, WORLD_BLOCKFIRSTLEAF, WORLD_VOXELMAT):
    ha = hashlib.sha256(a.world_export(k).tobytes()).hexdigest()
    hb = hashlib.sha256(b.world_export(k).tobytes()).hexdigest()
    same = same and ha == hb
out["identical_compact_arrays"] = same
out["stats"] = {k: sb[k] for k in ("nodes", "subChunks", "bricks", "blocks", "voxels", "compactBytes", "levels")}
out["speedup_vs_host_path"] = round((t_pack + t_up) / min(ts), 1)
print(json.dumps(out))
assert same
