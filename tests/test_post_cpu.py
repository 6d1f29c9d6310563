"""Denoiser + post-process, CPU side: the product's per-pixel functions (post.cuh, compiled
for the host by tests/host_emul) and its pass wiring against the reference's own compute shaders (oracle/_ref), and the
golden fixtures made from them. No GPU here; the kernels themselves are checked in test_gpu_post.py."""
import os

import numpy as np
import pytest

import blok_b200
from blok_b200 import Camera, PostSettings, WorldSvoGpu, scenes
from tests.conftest import to_scene

W, H, FRAMES = 96, 64, 5


def camera_path(n):
    """A slow dolly + pan over the 96^3 terrain, then a jump cut (forces disocclusion / history reset) on the last frame."""
    cams = [((48.0 + 0.35 * k, 62.0 - 0.1 * k, -24.0 + 0.5 * k), 90.0 + 0.4 * k, -27.0, 60.0) for k in range(n)]
    cams[-1] = ((20.0, 70.0, -10.0), 70.0, -35.0, 60.0)
    return cams


def gbuffer_sequence(reflib, frames=FRAMES, w=W, h=H, spp=2, jitter=True):
    """The reference's raygen over a camera path: per frame (ubo, color, worldPos, normalRough16, motion16)."""
    from oracle import pyoracle
    xyz, m, mats = scenes.terrain(96)
    world = WorldSvoGpu.from_voxels(xyz, m, mats)
    sc = to_scene(world)
    cams = camera_path(frames)
    out = []
    for k, cam in enumerate(cams):
        ubo = reflib.make_frame_seq(cam, cams[k - 1] if k else None, w, h, k, spp, k if jitter else -1)
        g = reflib.render(sc, ubo, threads=8)
        nr = pyoracle.float_to_half_bits(reflib.lib, g["normalRough"])
        mv = pyoracle.float_to_half_bits(reflib.lib, g["motion"][..., :2])
        out.append((ubo, g["color"], g["worldPos"], nr, mv))
    return out, world, cams


KEYS = ("historyColor", "moments", "historyLength", "variance", "denoised", "taaHistory", "tonemapped", "final")


def test_camera_frame_seq_matches_reference(reflib):
    cams = camera_path(4)
    for k, cam in enumerate(cams):
        for jit in (-1, k, k + 17):
            ref = reflib.make_frame_seq(cam, cams[k - 1] if k else None, W, H, k, 8, jit)
            ours = Camera(*cam).frame(W, H, k, 8, prev=Camera(*cams[k - 1]) if k else None, jitter_index=jit).to_numpy()
            assert np.array_equal(ref, ours), (k, jit, np.nonzero(ref != ours)[0][:8])


@pytest.mark.parametrize("settings", [{}, {"atrousIterations": 3, "tonemapOperator": 0, "saturationBoost": 0.8},
                                      {"atrousIterations": 1, "enableSharpening": 0, "exposure": 1.7}])
def test_pixel_functions_match_reference_shaders(reflib, emul, settings):
    """Five frames (camera motion, TAA jitter, a jump cut): every image the chain produces is bit-identical to the
    reference's compute shaders run on the CPU (same glibc expf on both sides, so even variance is exact here)."""
    from oracle import pyoracle
    from tests.host_emul.emu import EmuPostChain
    seq, _, _ = gbuffer_sequence(reflib)
    ref = pyoracle.PostChain(reflib.lib, "ref", W, H, threads=8, settings=pyoracle.PostSettings(**settings))
    ours = EmuPostChain(emul, W, H, PostSettings(**settings))
    for k, (ubo, col, wp, nr, mv) in enumerate(seq):
        a = ref.frame(ubo, col, wp, nr, mv)
        b = ours.frame(ubo, col, wp, nr, mv)
        for key in KEYS:
            x, y = np.squeeze(a[key]), np.squeeze(b[key])
            assert x.shape == y.shape and x.dtype == y.dtype, (k, key, x.shape, y.shape, x.dtype, y.dtype)
            same = x.view(np.uint8) == y.view(np.uint8) if x.dtype.kind != "f" else x.view(np.uint32) == y.view(np.uint32)
            assert same.all(), (k, key, int((~same).sum()), x[~same][:4], y[~same][:4])
    # the sequence exercised the interesting paths
    hl = np.squeeze(a["historyLength"]).view(np.float16).astype(np.float32)
    assert hl.max() >= 1.0


GOLD_POST = sorted(__import__("glob").glob(os.path.join(os.path.dirname(__file__), "golden", "post", "*.npz")))


@pytest.mark.parametrize("path", GOLD_POST, ids=lambda p: os.path.basename(p)[:-4])
def test_pixel_functions_match_golden_fixtures(emul, path):
    """The committed fixtures (reference shaders, run by tests/golden/make_golden_post.py) — usable without /root/reference."""
    from tests.host_emul.emu import EmuPostChain
    z = np.load(path)
    flt = ("exposure", "saturationBoost", "feedbackMin", "feedbackMax", "sharpenStrength", "whitePoint")
    st = PostSettings(**{k: (float(v) if k in flt else int(v)) for k, v in zip(z["settings_keys"], z["settings_vals"])})
    h, w = z["f0_color"].shape[:2]
    ours = EmuPostChain(emul, w, h, st)
    for k in range(int(z["frames"][0])):
        b = ours.frame(z[f"f{k}_ubo"], z[f"f{k}_color"], z[f"f{k}_worldPos"], z[f"f{k}_normalRough"], z[f"f{k}_motion"])
        for key in KEYS:
            x, y = z[f"f{k}_{key}"], np.squeeze(b[key])
            assert x.shape == y.shape and x.dtype == y.dtype, (k, key)
            assert np.array_equal(np.ascontiguousarray(x).view(np.uint8), np.ascontiguousarray(y).view(np.uint8)), (k, key)


def test_texel_centre_fetches_are_exact():
    """The fast path of the pinned sampler: (x + 0.5) / W * W - 0.5 quantises to an integer texel for every x at the sizes
    the bench uses (so a-trous / tonemap / sharpen taps read single texels, as the shader comments intend)."""
    for W in (64, 173, 1920, 2160, 3840, 7680):
        x = np.arange(W, dtype=np.float32)
        for off in (-1.0, 0.0, 1.0):
            u = (x + np.float32(0.5)) / np.float32(W) + np.float32(off) * (np.float32(1.0) / np.float32(W))
            xb = u * np.float32(W) - np.float32(0.5)
            q = ((xb - np.floor(xb)) * np.float32(256.0) + np.float32(0.5)).astype(np.int32)
            texel = np.floor(xb).astype(np.int64) + (q == 256)
            assert np.all((q == 0) | (q == 256)) and np.array_equal(texel, np.arange(W) + int(off)), (W, off)
