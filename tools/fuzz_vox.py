"""CPU fuzz of the .vox importer (blokb_vox_*; no GPU needed).
  python tools/fuzz_vox.py [seconds=60] [seed=1]
(1) well-formed random files (1-3 models, random sizes / voxels / duplicates, palette or none, random MATL dictionaries, chunks a
    loader must skip, fractional and negative offsets): voxel list + material table must give byte-identical packed arrays and
    materials to the reference's own loader (oracle/_ref: loadAndImportVox + rebuild + pack), when that library is present;
(2) damaged copies of the same files (truncations, flipped bytes, absurd counts): the importer must return or raise BlokbError —
    never crash, hang or over-allocate. The reference loader is NOT run on those (it trusts its input).
Checker-side infrastructure, like tools/fuzz_gpu.py."""
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from blok_b200 import BlokbError, WorldSvoGpu, load_vox  # noqa: E402
from oracle import pyoracle  # noqa: E402
from tests.voxfile import write_vox  # noqa: E402

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1
R = pyoracle.RefLib() if os.path.exists(pyoracle.REF_SO) else None
tmp = tempfile.mkdtemp()
TYPES = ["_diffuse", "_metal", "_glass", "_emit", "_blend", "_media", "_plastic", "weird", ""]
PROPS = ["_rough", "_metal", "_ior", "_emit", "_flux", "_alpha", "_sp", "_ldr", "_weight", "_trans"]


def same(a, b):
    return a.dtype == b.dtype and a.shape == b.shape and a.tobytes() == b.tobytes()


def canonical(nodes, subs):
    """Chunk-order-independent form of the packed world: the reference packs chunks in std::unordered_map order (unspecified),
    the restated builder in (z, y, x) order. Per chunk: its block of the node array and its sub-chunk records minus nodeOffset."""
    if len(subs) == 0:
        return {}
    offs = np.unique(subs["nodeOffset"])
    ends = dict(zip(offs.tolist(), offs.tolist()[1:] + [len(nodes)]))
    out = {}
    for o in offs.tolist():
        ss = subs[subs["nodeOffset"] == o]
        key = tuple(np.floor(np.asarray(ss["worldMin"][0], np.float64) / 128.0).astype(int).tolist())
        rec = sorted((int(r["rootNodeIndex"]), int(r["nodeCount"]), int(r["startDepth"]), np.asarray(r["worldMin"]).tobytes(),
                      np.asarray(r["subChunkSize"]).tobytes(), np.asarray(r["worldMax"]).tobytes()) for r in ss)
        out[key] = (nodes[o:ends[o]].tobytes(), rec)
    return out


def random_file(rng):
    models = []
    for _ in range(int(rng.integers(1, 4))):
        size = tuple(int(v) for v in rng.integers(1, 41, 3))
        n = int(rng.integers(0, 1500))
        v = np.stack([rng.integers(0, size[0], n), rng.integers(0, size[1], n), rng.integers(0, size[2], n), rng.integers(1, 256, n)], 1).astype(np.uint8)
        if n > 10 and rng.random() < 0.5:
            v[n // 2:n // 2 + 5, :3] = v[:5, :3]      # the same cell written twice with another colour
        models.append((size, v))
    palette = [int(x) for x in rng.integers(0, 2 ** 32, 255, dtype=np.uint64)] if rng.random() < 0.7 else None
    matl = {}
    for _ in range(int(rng.integers(0, 12))):
        props = {"_type": str(rng.choice(TYPES))} if rng.random() < 0.9 else {}
        for k in rng.choice(PROPS, int(rng.integers(0, 5)), replace=False):
            val = rng.choice([f"{rng.uniform(-1, 3):.4f}", f"{rng.uniform(0, 1):.2f}", "0", "1", "nan", "abc", "", "1e3", "-0.5", " 0.25"])
            props[str(k)] = str(val)
        matl[int(rng.integers(0, 300))] = props
    return write_vox(models, palette=palette, matl=matl, extra_chunks=bool(rng.random() < 0.6)), len(models)


t_end = time.time() + budget
seed, files, damaged, rejected = seed0, 0, 0, 0
while time.time() < t_end:
    rng = np.random.default_rng(seed)
    data, nmodels = random_file(rng)
    model = int(rng.integers(0, nmodels))
    off = tuple(float(v) for v in (rng.uniform(-300, 300, 3) if rng.random() < 0.7 else rng.integers(-200, 200, 3)))
    try:
        xyz, ids, mats, info = load_vox(data, model, off)
    except BlokbError as ex:
        # only legitimate on a well-formed file when empty models were merged away and the index now points past the end
        # (vox_loader.cpp:216): the reference must refuse the same file
        refused = True
        if R is not None:
            path = os.path.join(tmp, "f.vox")
            with open(path, "wb") as f:
                f.write(data)
            try:
                R.scene_from_vox(path, off, model)
                refused = False
            except RuntimeError:
                pass
        if "requested, but the file has" not in str(ex) or not refused:
            print(f"MISMATCH seed {seed}: importer refuses a file the reference loads: {ex}")
            sys.exit(1)
        files += 1
        seed += 1
        continue
    if R is not None:
        path = os.path.join(tmp, "f.vox")
        with open(path, "wb") as f:
            f.write(data)
        try:
            sc = R.scene_from_vox(path, off, model)
        except RuntimeError:
            if len(ids) == 0:   # loadAndImportVox returns false when nothing was imported (vox_loader.cpp:457): same outcome
                files += 1
                seed += 1
                continue
            print(f"MISMATCH seed {seed}: the reference loader rejects a file the importer accepts ({len(ids)} voxels, model {model})")
            sys.exit(1)
        ok = same(mats, sc.mats)
        if ok and len(ids):
            w = WorldSvoGpu.from_voxels(xyz, ids, mats)
            ok = len(w.globalNodes) == len(sc.nodes) and canonical(w.globalNodes, w.globalSubChunks) == canonical(sc.nodes, sc.subs)
        elif ok:
            ok = len(sc.subs) == 0
        if not ok:
            print(f"MISMATCH seed {seed}: importer differs from the reference loader (model {model}, offset {off})")
            sys.exit(1)
    files += 1
    for _ in range(6):   # damaged copies: ours only
        b = bytearray(data)
        kind = rng.random()
        if kind < 0.4:
            b = b[:int(rng.integers(0, len(b)))]
        elif kind < 0.8:
            for _ in range(int(rng.integers(1, 6))):
                b[int(rng.integers(0, len(b)))] = int(rng.integers(0, 256))
        else:
            p = int(rng.integers(8, max(9, len(b) - 4)))
            b[p:p + 4] = int(rng.choice([0x7FFFFFFF, 0xFFFFFFFF, 0x40000000, 0x80000000])).to_bytes(4, "little")
        try:
            load_vox(bytes(b), model, off)
        except BlokbError:
            rejected += 1
        damaged += 1
    seed += 1
print(f"vox fuzz ok: {files} random files (seeds {seed0}..{seed - 1}) " + ("byte-identical to the reference loader" if R is not None else "(reference library absent: parse only)")
      + f"; {damaged} damaged copies handled ({rejected} rejected, the rest parsed), no crash")
