#!/usr/bin/env python3
"""Generates tests/golden/vox/: a small two-model .vox file written by tests/voxfile.py (our own data, not a reference asset)
and what the REFERENCE's loader + builder (oracle/_ref: loadAndImportVox, rebuildDirtyChunks, packChunksToGpuSvo,
MaterialLibrary::packForGpu) make of it. Pins blokb_vox_* / blokb_upload_vox where /root/reference is not mounted.
Re-run:  python tests/golden/make_golden_vox.py   (needs /root/reference)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from tests.voxfile import write_vox  # noqa: E402


def synth_models():
    rng = np.random.default_rng(42)
    g = np.stack(np.meshgrid(np.arange(40), np.arange(36), np.arange(30), indexing="ij"), -1).reshape(-1, 3)
    keep = ((g[:, 0] - 20) ** 2 + (g[:, 1] - 18) ** 2 + (g[:, 2] - 15) ** 2 < 14 ** 2) & (rng.random(len(g)) < 0.6)
    a = g[keep]
    a = np.concatenate([a, rng.integers(1, 256, (len(a), 1))], 1).astype(np.uint8)
    b = rng.integers(0, 24, (900, 3))
    b = np.concatenate([b, rng.integers(1, 256, (len(b), 1))], 1).astype(np.uint8)  # repeated cells: the last write wins
    return [((40, 36, 30), a), ((24, 24, 24), b)]


def synth_file():
    rng = np.random.default_rng(43)
    palette = [int(v) for v in rng.integers(0, 2 ** 32, 255, dtype=np.uint64)]
    matl = {3: {"_type": "_metal", "_rough": 0.25, "_metal": 0.9, "_sp": 0.7}, 7: {"_type": "_emit", "_emit": 3.5, "_flux": 2},
            9: {"_type": "_emit"}, 11: {"_type": "_glass", "_ior": 1.33, "_alpha": 0.4, "_rough": 0.05}, 200: {"_type": "_diffuse", "_rough": "oops"},
            255: {"_type": "_emit", "_flux": 4}}
    return write_vox(synth_models(), palette, matl)


def main():
    pyoracle.build(ref=True)
    R = pyoracle.RefLib()
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "vox")
    os.makedirs(out, exist_ok=True)
    path = os.path.join(out, "synth.vox")
    with open(path, "wb") as f:
        f.write(synth_file())
    g = {}
    for tag, model, off in (("m0", 0, (0, 0, 0)), ("m1", 1, (60, 10, 70))):
        sc = R.scene_from_vox(path, off, model)
        g[tag + "_nodes"], g[tag + "_subs"], g[tag + "_mats"] = sc.nodes, sc.subs, sc.mats
        g[tag + "_offset"] = np.array(off, np.float32)
        print(tag, len(sc.nodes), len(sc.subs))
    np.savez_compressed(os.path.join(out, "synth_ref.npz"), **g)


if __name__ == "__main__":
    main()
