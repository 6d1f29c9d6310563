#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the REFERENCE ITSELF, run in this container (oracle/_ref = the reference's own
CPU scene-build sources + its RT shader sources compiled in place; see oracle/Makefile). The reference has no tests
or golden vectors of its own, so these outputs are what pins the oracle restatement and, through it,
the CUDA path. Re-run:  python tests/golden/make_golden.py   (needs /root/reference; writes only tests/golden/).

Per case: the packed world arrays the reference would upload (derived data, not reference source), the FrameUBO the
reference would fill for the camera, per-pixel hit records {prim, node, material, face, t} of intersect.rint for the
sample-0 primary rays at 64x64 (full) and 512x512 (FNV-1a-64 + counts), and raygen.rgen's five images at 64x64, 4 spp.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402

ASSETS = "/root/reference/assets/models/"
CASES = [  # name, asset, world offset, camera pos, yaw, pitch   (cameras: SURVEY.md Appendix B.4 where listed)
    ("chr_knight", "chr_knight.vox", (0, 0, 0), (-20, 10, 10), 0.0, 0.0),
    ("castle", "castle.vox", (0, 0, 0), (-30, 20, -25), 40.0, -15.0),
    ("teapot", "teapot.vox", (0, 0, 0), (-60, 40, 30), 0.0, -10.0),
    ("teapot_neg", "teapot.vox", (-30, 0, 100), (-90, 45, 125), 5.0, -12.0),  # 4 chunks, negative coordinates (App. B.2)
]


def main():
    pyoracle.build(ref=True)
    R = pyoracle.RefLib()
    O = pyoracle.OracleLib()
    out_dir = os.path.dirname(os.path.abspath(__file__))
    for name, asset, off, pos, yaw, pitch in CASES:
        sc = R.scene_from_vox(ASSETS + asset, off)
        g = {"nodes": sc.nodes, "subs": sc.subs, "mats": sc.mats}
        # primary hits
        for tag, w in (("s", 64), ("l", 512)):
            ubo = R.make_frame(pos, yaw, pitch, w, w, frame_count=0, sample_count=1)
            rays = R.primary_rays(ubo)
            hit, t, cnt = R.trace(sc, rays)
            rec = np.concatenate([hit, t.view(np.uint32)[:, None]], 1).astype(np.uint32)  # prim,node,mat,face,bits(t)
            g[f"ubo_{tag}"] = ubo
            if tag == "s":
                g["hits_s"] = rec
            else:
                g["fnv_l"] = np.array([int(O.fnv(rec), 16)], np.uint64)
                g["nhit_l"] = np.array([(t >= 0).sum()], np.int64)
                g["meant_l"] = np.array([t[t >= 0].astype(np.float64).mean() if (t >= 0).any() else 0.0])
                g["counters_l"] = cnt  # node fetches, AABB passes, cap hits over the 512x512 frame
        # shadow rays (terminate on first hit) from the 64x64 primary hits towards the sun
        ubo = R.make_frame(pos, yaw, pitch, 64, 64, frame_count=3, sample_count=4)
        imgs = R.render(sc, ubo)
        g["ubo_r"] = ubo
        for k in ("color", "worldPos", "normalRough", "albedoMetal", "motion"):
            g["img_" + k] = imgs[k]
        g["counters_r"] = imgs["counters"]
        path = os.path.join(out_dir, name + ".npz")
        np.savez_compressed(path, **g)
        print(name, "nodes", len(sc.nodes), "subs", len(sc.subs), "hits64", int((g["hits_s"][:, 0] != 0xFFFFFFFF).sum()),
              "hits512", int(g["nhit_l"][0]), f"fnv {int(g['fnv_l'][0]):016x}", os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
