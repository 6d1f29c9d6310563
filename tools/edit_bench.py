"""Editing a world: full re-upload of all dense chunks (what the reference's updateWorld amounts to) vs blokb_update_dense of the
dirty chunk only. python tools/edit_bench.py [N=768]"""
import hashlib
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import blok_b200  # noqa: E402
from blok_b200 import WORLD_BLOCKFIRSTLEAF, WORLD_NODES, WORLD_VOXELMAT, scenes  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 768
xyz, mid, mats = scenes.terrain(N)
# voxel list -> dense chunk arrays, vectorised (ChunkManager::setVoxel layout: x + 128 y + 16384 z)
c = np.floor_divide(xyz, 128)
l = xyz - c * 128
t = l[:, 0] + 128 * l[:, 1] + 16384 * l[:, 2]
keys, inv = np.unique(c, axis=0, return_inverse=True)
chunks = {}
for k, key in enumerate(keys):
    sel = inv.reshape(-1) == k
    den, ids = np.zeros(128 ** 3, np.float32), np.zeros(128 ** 3, np.uint32)
    den[t[sel]] = 1.0
    ids[t[sel]] = mid[sel]
    chunks[tuple(int(v) for v in key)] = (den, ids)
r = blok_b200.Renderer(64, 64, 0)
r.add_dense(chunks, mats)
ts = []
for _ in range(3):
    t0 = time.perf_counter(); r.add_dense(chunks, mats); ts.append(time.perf_counter() - t0)
full_ms = min(ts) * 1e3
# edit: dig a crater in the middle chunk
key = tuple(int(v) for v in keys[len(keys) // 2])
den = chunks[key][0].reshape(128, 128, 128)
den[32:96, :, 32:96] = 0.0
ts = []
for _ in range(3):
    t0 = time.perf_counter(); r.update_dense({key: chunks[key]}); ts.append(time.perf_counter() - t0)
upd_ms = min(ts) * 1e3
full = blok_b200.Renderer(64, 64, 0)
full.add_dense(chunks, mats)
same = all(hashlib.sha256(r.world_export(k).tobytes()).digest() == hashlib.sha256(full.world_export(k).tobytes()).digest()
           for k in (WORLD_NODES, WORLD_BLOCKFIRSTLEAF, WORLD_VOXELMAT))
print(json.dumps({"scene": f"terrain({N})", "chunks": len(chunks), "voxels": int(r.world_stats()["voxels"]),
                  "full_upload_dense_ms": round(full_ms, 2), "full_input_MB": round(len(chunks) * 16.78, 1),
                  "update_one_dirty_chunk_ms": round(upd_ms, 2), "update_input_MB": 16.8, "identical": bool(same)}))
assert same
