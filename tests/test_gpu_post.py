"""GPU: the CUDA denoiser + post-process chain (blokb_denoise / blokb_postprocess / blokb_end_frame through the C-ABI)
against the reference's compute shaders — committed golden fixtures (tests/golden/post, made from oracle/_ref) and, where
oracle/_ref travelled to the box, the reference shaders run live on larger frames.

Tolerances. Everything except exp() rounds identically on both sides, so:
  * temporal accumulation (history colour, moments, history length): bit-exact;
  * variance: its spatial half goes through exp() (CUDA expf vs glibc, <= 2 ulp) and then through the cancellation
    m2/W - mean^2 of luminances ~0.3: absolute 2e-6 + relative 1e-5;
  * denoised / TAA history: absolute 2e-5 + relative 1e-4 on >= 99.9 % of values (a 1-ulp change in a weight can cross one
    of the a-trous filter's hard thresholds on a rare pixel), and the image mean within 1e-5 relative;
  * tonemapped / final RGBA8: equal on >= 99.5 % of bytes, never more than 4 steps apart (sharpen amplifies a 1-step
    difference in a neighbour by up to 1.5 x 2).
"""
import glob
import os

import numpy as np
import pytest

import blok_b200
from blok_b200 import (IMG_COLOR, IMG_DENOISED, IMG_FINAL, IMG_HISTORY_COLOR, IMG_HISTORY_LENGTH, IMG_MOMENTS, IMG_MOTION, IMG_NORMALROUGH,
                       IMG_TAA_HISTORY, IMG_TONEMAPPED, IMG_VARIANCE, IMG_WORLDPOS, BlokbError, Camera, PostSettings)

pytestmark = pytest.mark.gpu
GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "post", "*.npz")))
IMAGES = {"historyColor": IMG_HISTORY_COLOR, "moments": IMG_MOMENTS, "historyLength": IMG_HISTORY_LENGTH, "variance": IMG_VARIANCE,
          "denoised": IMG_DENOISED, "taaHistory": IMG_TAA_HISTORY, "tonemapped": IMG_TONEMAPPED, "final": IMG_FINAL}
EXACT = ("historyColor", "moments", "historyLength")


def compare(key, got, want, where=""):
    got, want = np.squeeze(got), np.squeeze(want)
    if want.dtype == np.uint16:
        want = want.view(np.float16)
    assert got.shape == want.shape, (key, got.shape, want.shape)
    if key in EXACT:
        a, b = np.ascontiguousarray(got), np.ascontiguousarray(want)
        same = a.view(np.uint8) == b.view(np.uint8)
        assert same.all(), (where, key, int((~same).sum()))
    elif key == "variance":
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=2e-6, err_msg=f"{where} {key}")
    elif want.dtype == np.uint8:
        d = np.abs(got.astype(np.int32) - want.astype(np.int32))
        assert d.max() <= 4 and (d == 0).mean() >= 0.995, (where, key, int(d.max()), float((d == 0).mean()))
    else:
        ok = np.abs(got - want) <= 2e-5 + 1e-4 * np.abs(want)
        assert ok.mean() >= 0.999, (where, key, float(ok.mean()), float(np.abs(got - want).max()))
        assert abs(float(got.mean()) - float(want.mean())) <= 1e-5 * abs(float(want.mean())) + 1e-7, (where, key)


def run_frame(r, ubo, color, worldPos, normalRough16, motion16, settings):
    r.set_frame(ubo)
    r.write_image(IMG_COLOR, color)
    r.write_image(IMG_WORLDPOS, worldPos)
    r.write_image(IMG_NORMALROUGH, normalRough16)
    r.write_image(IMG_MOTION, motion16)
    r.denoise(settings)
    r.postprocess(settings)
    out = {k: r.readback(i) for k, i in IMAGES.items()}
    r.end_frame()
    return out


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[:-4])
def test_chain_matches_reference_goldens(renderer_factory, path):
    z = np.load(path)
    st = PostSettings(**{k: (float(v) if k in ("exposure", "saturationBoost", "feedbackMin", "feedbackMax", "sharpenStrength", "whitePoint") else int(v))
                         for k, v in zip(z["settings_keys"], z["settings_vals"])})
    h, w = z["f0_color"].shape[:2]
    r = renderer_factory(w, h)
    launches0 = r.launch_count()
    for k in range(int(z["frames"][0])):
        out = run_frame(r, z[f"f{k}_ubo"], z[f"f{k}_color"], z[f"f{k}_worldPos"], z[f"f{k}_normalRough"], z[f"f{k}_motion"], st)
        for key in IMAGES:
            compare(key, out[key], z[f"f{k}_{key}"], f"frame {k}")
    per_frame = 1 + st.atrousIterations + 1 + (1 if st.enableSharpening else 0)
    assert r.launch_count() - launches0 == per_frame * int(z["frames"][0])


@pytest.mark.parametrize("shape,settings", [((256, 144), {}), ((173, 67), {"atrousIterations": 5}), ((64, 64), {"enableSharpening": 0, "atrousIterations": 1}),
                                            ((8, 3), {}), ((1, 1), {})])
def test_chain_matches_reference_shaders_live(renderer_factory, reflib, shape, settings):
    """The reference's own shaders (oracle/_ref) on frames whose sides are not multiples of the block shape, down to 1x1."""
    from oracle import pyoracle
    from tests.test_post_cpu import gbuffer_sequence
    w, h = shape
    seq, _, _ = gbuffer_sequence(reflib, 4, w, h, 1, True)
    ref = pyoracle.PostChain(reflib.lib, "ref", w, h, threads=8, settings=pyoracle.PostSettings(**settings))
    st = PostSettings(**settings)
    r = renderer_factory(w, h)
    for k, (ubo, col, wp, nr, mv) in enumerate(seq):
        want = ref.frame(ubo, col, wp, nr, mv)
        got = run_frame(r, ubo, col, wp, nr, mv, st)
        for key in IMAGES:
            compare(key, got[key], want[key], f"{shape} frame {k}")


def test_rendered_frames_through_the_whole_pipeline(renderer_factory, reflib):
    """blokb_render -> denoise -> postprocess on the GPU against raygen + the six compute shaders of the reference on the CPU.
    Radiance itself is only within 2e-3 (sin/cos/pow), so the chain is compared at that level: final RGBA8 within 2 steps on 99 %."""
    from oracle import pyoracle
    from tests.test_post_cpu import gbuffer_sequence
    w, h = 128, 80
    seq, world, cams = gbuffer_sequence(reflib, 4, w, h, 2, True)
    ref = pyoracle.PostChain(reflib.lib, "ref", w, h, threads=8)
    r = renderer_factory(w, h)
    r.add_world(world)
    st = PostSettings()
    for k, (ubo, col, wp, nr, mv) in enumerate(seq):
        want = ref.frame(ubo, col, wp, nr, mv)
        f = Camera(*cams[k]).frame(w, h, k, 2, prev=Camera(*cams[k - 1]) if k else None, jitter_index=k)
        assert np.array_equal(f.to_numpy(), ubo)
        r.render(f)
        r.denoise(st)
        r.postprocess(st)
        fin, den = r.readback(IMG_FINAL), r.readback(IMG_DENOISED)
        r.end_frame()
        d = np.abs(fin.astype(np.int32) - want["final"].astype(np.int32))
        assert (d <= 2).mean() >= 0.99, (k, float((d <= 2).mean()))
        rel = np.abs(den[..., :3] - want["denoised"][..., :3]) / (np.abs(want["denoised"][..., :3]) + 1e-2)
        assert (rel < 1e-2).mean() >= 0.99, (k, float((rel < 1e-2).mean()))


def test_call_order_and_arguments(renderer_factory):
    r = renderer_factory(32, 16)
    f = Camera((0, 0, 0), 0.0, 0.0).frame(32, 16, 0, 1)
    st = PostSettings()
    with pytest.raises(BlokbError) as e:
        r.denoise(st)  # no frame constants yet
    assert e.value.code == -4
    r.set_frame(f)
    with pytest.raises(BlokbError) as e:
        r.postprocess(st)  # nothing denoised this frame
    assert e.value.code == -4
    for bad in (0, 6):
        with pytest.raises(BlokbError) as e:
            r.denoise(PostSettings(atrousIterations=bad))
        assert e.value.code == -1
    with pytest.raises(BlokbError) as e:
        r.readback(IMG_FINAL)
    assert e.value.code == -4
    r.denoise(st)
    with pytest.raises(BlokbError) as e:
        r.postprocess(PostSettings(enableTAA=0, enableTonemapping=0))
    assert e.value.code == -3
    r.postprocess(st)
    assert r.readback(IMG_FINAL).shape == (16, 32, 4)
    r.end_frame()
    with pytest.raises(BlokbError):
        r.readback(IMG_DENOISED)  # belongs to the finished frame
    r.reset_history()


def test_history_reset_and_determinism(renderer_factory):
    """Same inputs twice (with a history reset in between) give identical bytes; without the reset the history differs."""
    z = np.load(GOLD[0])
    h, w = z["f0_color"].shape[:2]
    r = renderer_factory(w, h)
    st = PostSettings()

    def run():
        outs = []
        for k in range(2):
            outs.append(run_frame(r, z[f"f{k}_ubo"], z[f"f{k}_color"], z[f"f{k}_worldPos"], z[f"f{k}_normalRough"], z[f"f{k}_motion"], st))
        return outs
    a = run()
    r.reset_history()
    b = run()
    for x, y in zip(a, b):
        for key in IMAGES:
            assert np.array_equal(x[key].view(np.uint8), y[key].view(np.uint8)), key


def test_tonemap_off_outputs_float(renderer_factory):
    z = np.load(GOLD[0])
    h, w = z["f0_color"].shape[:2]
    r = renderer_factory(w, h)
    st = PostSettings(enableTonemapping=0)
    out = run_frame(r, z["f0_ubo"], z["f0_color"], z["f0_worldPos"], z["f0_normalRough"], z["f0_motion"], st)
    assert out["final"].dtype == np.float32
    # frame 0: feedback = 0, so taaOutput = current + 0.1 * (current - YCoCgToRGB(mean)) and the history is the denoised frame
    np.testing.assert_array_equal(out["taaHistory"][..., :3], out["denoised"][..., :3])
