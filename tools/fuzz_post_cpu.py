"""CPU fuzz of the denoiser + post-process pixel functions (post.cuh compiled for the host, tests/host_emul) against the
reference's own compute shaders (oracle/_ref), bit for bit, on SYNTHETIC G-buffers the raygen would never produce:
  python tools/fuzz_post_cpu.py [seconds=60] [seed=1]
random odd image sizes, heavy-tailed / zero / huge / negative colours and now and then NaN or inf, random sky masks, normals that
are not unit length, motion vectors that leave the screen, random camera sequences (jitter on / off), random settings. Three frames
per case so the temporal paths (history, moments, TAA feedback) are live. Needs oracle/_ref (build container only)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from blok_b200 import PostSettings  # noqa: E402
from oracle import pyoracle  # noqa: E402
from tests.host_emul import emu  # noqa: E402
from tests.host_emul.emu import EmuPostChain  # noqa: E402

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1
R = pyoracle.RefLib()
E = emu.load()
KEYS = ("historyColor", "moments", "historyLength", "variance", "denoised", "taaHistory", "tonemapped", "final")


def gbuffer(rng, w, h, wild):
    col = np.abs(rng.standard_cauchy((h, w, 4))).astype(np.float32) * rng.choice([0.01, 0.3, 1.0, 20.0])
    col[..., 3] = 1.0
    if rng.random() < 0.3:
        col[rng.random((h, w)) < 0.2] = 0.0
    wp = np.empty((h, w, 4), np.float32)
    wp[..., :3] = rng.uniform(-200, 200, (h, w, 3))
    depth = rng.uniform(0.01, 400.0, (h, w)).astype(np.float32)
    sky = rng.random((h, w)) < rng.choice([0.0, 0.1, 0.5])
    if rng.random() < 0.5:      # smooth depth so the a-trous / temporal depth tests pass somewhere
        yy, xx = np.mgrid[0:h, 0:w]
        depth = (20.0 + 0.3 * xx + 0.1 * yy + rng.uniform(0, 0.05, (h, w))).astype(np.float32)
        wp[..., 0], wp[..., 1], wp[..., 2] = xx * 0.5, 3.0 + rng.uniform(0, 0.02, (h, w)), yy * 0.5
    wp[..., 3] = np.where(sky, 10000.0, depth)
    nrm = rng.normal(size=(h, w, 3))
    if rng.random() < 0.6:
        nrm = np.tile(np.eye(3)[int(rng.integers(0, 3))] * rng.choice([-1.0, 1.0]), (h, w, 1)) + (rng.normal(size=(h, w, 3)) * 0.01 if rng.random() < 0.5 else 0.0)
    else:
        nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True) * rng.choice([1.0, 0.5, 2.0])
    nr = np.concatenate([nrm, rng.uniform(0, 1, (h, w, 1))], -1).astype(np.float32)
    mv = (rng.normal(size=(h, w, 2)) * rng.choice([0.0, 0.002, 0.05, 1.5])).astype(np.float32)
    if wild:
        k = max(1, (h * w) // 97)
        ys, xs = rng.integers(0, h, k), rng.integers(0, w, k)
        col[ys, xs, int(rng.integers(0, 3))] = rng.choice([np.nan, np.inf, -1.0, 1e30, 65504.0 * 4])
        if rng.random() < 0.5:
            mv[ys[: k // 2], xs[: k // 2]] = rng.choice([np.nan, np.inf, 1e4])
    return col, wp, pyoracle.float_to_half_bits(R.lib, nr), pyoracle.float_to_half_bits(R.lib, mv)


t_end = time.time() + budget
seed, cases, frames = seed0, 0, 0
while time.time() < t_end:
    rng = np.random.default_rng(seed)
    w, h = int(rng.integers(3, 90)), int(rng.integers(3, 70))
    settings = dict(atrousIterations=int(rng.integers(1, 6)), enableTAA=int(rng.random() < 0.8), enableTonemapping=int(rng.random() < 0.8),
                    tonemapOperator=int(rng.integers(0, 3)), exposure=float(rng.choice([0.25, 1.0, 1.7, 8.0])),
                    saturationBoost=float(rng.choice([0.0, 0.8, 1.0, 1.4])), enableSharpening=int(rng.random() < 0.7),
                    sharpenStrength=float(rng.choice([0.0, 0.3, 1.0, 2.5])), feedbackMin=float(rng.choice([0.0, 0.85, 0.88])),
                    feedbackMax=float(rng.choice([0.9, 0.97, 1.0])), whitePoint=float(rng.choice([1.0, 4.0, 11.2])))
    if not settings["enableTAA"] and not settings["enableTonemapping"]:
        settings["enableTonemapping"] = 1   # neither: PostProcess::getOutputImage returns an image nothing wrote (no passthrough); the product refuses it
    try:
        ref = pyoracle.PostChain(R.lib, "ref", w, h, threads=4, settings=pyoracle.PostSettings(**settings))
        ours = EmuPostChain(E, w, h, PostSettings(**settings))
    except TypeError:
        settings = {k: v for k, v in settings.items() if k in ("atrousIterations", "tonemapOperator", "exposure", "saturationBoost", "enableSharpening")}
        ref = pyoracle.PostChain(R.lib, "ref", w, h, threads=4, settings=pyoracle.PostSettings(**settings))
        ours = EmuPostChain(E, w, h, PostSettings(**settings))
    wild = rng.random() < 0.35
    pos = rng.uniform(-50, 50, 3)
    cams = [((float(pos[0] + 0.4 * k), float(pos[1]), float(pos[2] - 0.3 * k)), float(rng.uniform(0, 360) if k == 0 else 0.0) + 0.5 * k, -20.0, 60.0) for k in range(3)]
    cams = [(c[0], cams[0][1] + 0.5 * k, c[2], c[3]) for k, c in enumerate(cams)]
    jitter = rng.random() < 0.6
    for k in range(3):
        ubo = R.make_frame_seq(cams[k], cams[k - 1] if k else None, w, h, k, 8, k if jitter else -1)
        col, wp, nr, mv = gbuffer(rng, w, h, wild)
        a, b = ref.frame(ubo, col, wp, nr, mv), ours.frame(ubo, col, wp, nr, mv)
        for key in KEYS:
            if (key == "tonemapped" and not settings.get("enableTonemapping", 1)) or (key == "taaHistory" and not settings.get("enableTAA", 1)):
                continue   # a disabled stage leaves its image unwritten on both sides: nothing to compare
            x, y = np.ascontiguousarray(np.squeeze(a[key])), np.ascontiguousarray(np.squeeze(b[key]))
            if x.shape == y.shape and x.dtype == y.dtype and x.dtype == np.float32:
                # a NaN is a NaN: which operand's sign / payload survives an x86 addss depends on operand order the compilers chose
                both = np.isnan(x) & np.isnan(y)
                x, y = np.where(both, np.float32(0), x), np.where(both, np.float32(0), y)
            if x.shape != y.shape or x.dtype != y.dtype or x.tobytes() != y.tobytes():
                diff = int((x.view(np.uint8) != y.view(np.uint8)).sum()) if x.shape == y.shape and x.dtype == y.dtype else -1
                print(f"MISMATCH seed {seed} frame {k} image {key}: {diff} differing bytes; size {w}x{h} wild={wild} settings {settings}")
                sys.exit(1)
        frames += 1
    cases += 1
    seed += 1
print(f"post fuzz ok: {cases} cases / {frames} frames (seeds {seed0}..{seed - 1}): all eight images bit-identical to the reference's compute shaders")
