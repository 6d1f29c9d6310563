#!/usr/bin/env python3
"""Generates tests/golden/post/*.npz from the REFERENCE ITSELF: the reference's raygen (G-buffer) and its six compute
shaders (temporal_reproject / variance / atrous / taa / tonemap / sharpen .comp), compiled in place into oracle/_ref and
run in this container over a short camera path. The reference has no tests of its own for these passes;
these outputs pin the CUDA denoiser / post-process chain (tests/test_gpu_post.py) and the CPU emulation
(tests/test_post_cpu.py). Re-run:  python tests/golden/make_golden_post.py   (needs /root/reference).

Per case: for every frame the inputs (FrameUBO, color, worldPos, normalRough, motion as the reference stores them) and
every image the chain produces that frame (history colour / moments / length, variance, denoised, TAA history,
tonemapped, final)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402
from tests.test_post_cpu import KEYS, gbuffer_sequence  # noqa: E402

CASES = {  # name -> (frames, w, h, spp, taa jitter, PostSettings overrides)
    "terrain_default": (4, 64, 40, 2, True, {}),
    "terrain_softclip": (3, 48, 32, 1, False, {"atrousIterations": 2, "tonemapOperator": 0, "saturationBoost": 0.8, "exposure": 2.5}),
}


def main():
    pyoracle.build(ref=True)
    R = pyoracle.RefLib()
    out_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "post")
    os.makedirs(out_dir, exist_ok=True)
    for name, (frames, w, h, spp, jitter, st) in CASES.items():
        seq, _, _ = gbuffer_sequence(R, frames, w, h, spp, jitter)
        chain = pyoracle.PostChain(R.lib, "ref", w, h, threads=8, settings=pyoracle.PostSettings(**st))
        g = {"settings_keys": np.array(list(st.keys())), "settings_vals": np.array([float(v) for v in st.values()], np.float64),
             "frames": np.array([frames])}
        for k, (ubo, col, wp, nr, mv) in enumerate(seq):
            o = chain.frame(ubo, col, wp, nr, mv)
            g[f"f{k}_ubo"], g[f"f{k}_color"], g[f"f{k}_worldPos"], g[f"f{k}_normalRough"], g[f"f{k}_motion"] = ubo, col, wp, nr, mv
            for key in KEYS:
                g[f"f{k}_{key}"] = np.squeeze(o[key]).copy()
        path = os.path.join(out_dir, name + ".npz")
        np.savez_compressed(path, **g)
        print(name, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
