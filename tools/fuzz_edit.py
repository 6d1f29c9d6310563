"""GPU fuzz of the editable-world path: random brush sessions, blokb_update_dense (dirty chunks only) vs blokb_upload_dense
(all chunks) after every edit, plus a scene-dump round trip of the edited world.
  python tools/fuzz_edit.py [seconds=60] [seed=1]
Worlds are small (a few 128^3 chunks around the origin, negative coordinates included); edits are spheres that add (with a
random, sometimes fractional density) or subtract, sometimes in chunks that do not exist yet, sometimes erasing a whole chunk's
content so that it stays behind as an empty chunk. Exits non-zero on the first difference and prints the seed."""
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import blok_b200  # noqa: E402
from blok_b200 import WORLD_BLOCKFIRSTLEAF, WORLD_NODES, WORLD_SUBCHUNKS, WORLD_VOXELMAT, scenes  # noqa: E402

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1
KEYS = (WORLD_NODES, WORLD_BLOCKFIRSTLEAF, WORLD_VOXELMAT, WORLD_SUBCHUNKS)
mats = scenes.default_materials(16)


def sphere(chunks, c, radius, density, material):
    """Writes density (and, when adding, the material id) into every cell within `radius` of c; returns the dirty chunk keys."""
    lo, hi = np.floor(c - radius).astype(int), np.floor(c + radius).astype(int) + 1
    g = np.stack(np.meshgrid(np.arange(lo[0], hi[0]), np.arange(lo[1], hi[1]), np.arange(lo[2], hi[2]), indexing="ij"), -1).reshape(-1, 3)
    g = g[np.sqrt((((g + 0.5) - c) ** 2).sum(1)) <= radius]
    dirty = set()
    if not len(g):
        return dirty
    cc = np.floor_divide(g, 128)
    loc = g - cc * 128
    t = loc[:, 0] + 128 * loc[:, 1] + 16384 * loc[:, 2]
    for key in {tuple(int(v) for v in k) for k in np.unique(cc, axis=0)}:
        sel = (cc == np.array(key)).all(1)
        if key not in chunks:
            chunks[key] = (np.zeros(128 ** 3, np.float32), np.zeros(128 ** 3, np.uint32))
        chunks[key][0][t[sel]] = density
        if density > 0:
            chunks[key][1][t[sel]] = material
        dirty.add(key)
    return dirty


def export(r):
    return [r.world_export(k).tobytes() for k in KEYS]


r = blok_b200.Renderer(64, 64, 0)
full = blok_b200.Renderer(64, 64, 0)
loaded = blok_b200.Renderer(64, 64, 0)
tmp = tempfile.mkdtemp()
t_end = time.time() + budget
seed, sessions, edits = seed0, 0, 0
while time.time() < t_end:
    rng = np.random.default_rng(seed)
    chunks = {}
    for _ in range(int(rng.integers(1, 5))):
        sphere(chunks, rng.uniform(-100, 200, 3), rng.uniform(2, 25), 1.0, int(rng.integers(1, 17)))
    r.add_dense(chunks, mats)
    for e in range(int(rng.integers(2, 7))):
        kind = rng.random()
        if kind < 0.45:
            dirty = sphere(chunks, rng.uniform(-100, 200, 3), rng.uniform(1, 20), float(rng.choice([1.0, 0.5, 0.01])), int(rng.integers(1, 17)))
        elif kind < 0.85:
            dirty = sphere(chunks, rng.uniform(-100, 200, 3), rng.uniform(1, 30), float(rng.choice([0.0, -1.0])), 0)
        else:   # wipe one existing chunk completely
            key = list(chunks)[int(rng.integers(0, len(chunks)))]
            chunks[key][0][:] = 0.0
            dirty = {key}
        if not dirty:
            continue
        r.update_dense({k: chunks[k] for k in dirty})
        full.add_dense(chunks, mats)
        if export(r) != export(full) or r.world_stats()["nodes"] != full.world_stats()["nodes"]:
            print(f"MISMATCH seed {seed} edit {e}: update_dense differs from a full upload_dense (dirty {sorted(dirty)})")
            sys.exit(1)
        edits += 1
    path = os.path.join(tmp, "w.blokb")
    r.save_world(path)
    loaded.load_world(path)
    if export(loaded) != export(r):
        print(f"MISMATCH seed {seed}: scene dump round trip")
        sys.exit(1)
    sessions += 1
    seed += 1
print(f"edit fuzz ok: {sessions} sessions (seeds {seed0}..{seed - 1}), {edits} edits: dirty-chunk update == full upload after every edit, "
      f"scene dump round trip byte-identical")
for x in (r, full, loaded):
    x.close()
