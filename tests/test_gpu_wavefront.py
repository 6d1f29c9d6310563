"""Wavefront path tracing (k_path_front -> radix sort -> k_bounce -> k_path_gather) against the fused kernel (k_path): both run
raygen.rgen's expressions in raygen's order, so the accumulation buffer and the colour image must be BIT-identical whatever the
batch size, the sort-key shape or the strip partition. Through the C-ABI (blokb_set_path_options / blokb_trace_path / blokb_render)."""
import os

import numpy as np
import pytest

import blok_b200
from blok_b200 import IMG_ACCUM, IMG_COLOR, PATH_FUSED, PATH_WAVEFRONT, WorldSvoGpu, scenes

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _gold_world(name="castle"):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return z, WorldSvoGpu(z["nodes"], z["subs"], z["mats"])


def _accum(r, a, b):
    r.reset_accum()
    r.trace_path(a, b)
    return r.readback(IMG_ACCUM).copy()


@pytest.mark.parametrize("name", ["castle", "teapot_neg"])
def test_wavefront_accum_and_colour_equal_fused_bitwise(renderer_factory, name):
    z, w = _gold_world(name)
    r = renderer_factory()
    r.add_world(w)
    ubo = blok_b200.FrameUBO.from_bytes(z["ubo_r"])
    ubo.sampleCount = 6
    r.set_frame(ubo)
    r.set_path_options(mode=PATH_FUSED)
    want_acc = _accum(r, 0, 6)
    r.render()
    want_col = r.readback(IMG_COLOR).copy()
    assert np.abs(want_acc[..., :3]).sum() > 0 and np.all(want_acc[..., 3] == 6.0)
    for batch in (1, 4, 6, 16):
        o = r.set_path_options(mode=PATH_WAVEFRONT, batch_samples=batch)
        assert o["mode"] == PATH_WAVEFRONT and o["batchSamples"] == batch
        assert _accum(r, 0, 6).tobytes() == want_acc.tobytes(), f"accum differs at batch {batch}"
        r.render()
        assert r.readback(IMG_COLOR).tobytes() == want_col.tobytes(), f"colour differs at batch {batch}"
    # the sort only permutes independent rays: any key shape gives the same image
    for dc, sh, mb in ((1, 6, 1), (4, 4, 6), (16, 5, 5), (8, 6, 5)):
        r.set_path_options(mode=PATH_WAVEFRONT, batch_samples=4, dir_cells=dc, origin_cell_shift=sh, morton_bits_per_axis=mb)
        assert _accum(r, 0, 6).tobytes() == want_acc.tobytes(), f"accum differs with key ({dc}, {sh}, {mb})"
    # persistent bounce warps whose idle lanes are refilled from the sorted list: which thread walks a ray does not change its result
    for rb in (1, 7, 16, 32):
        o = r.set_path_options(mode=PATH_WAVEFRONT, batch_samples=4, refill_below=rb)
        assert o["refillBelow"] == rb
        assert _accum(r, 0, 6).tobytes() == want_acc.tobytes(), f"accum differs with lanes refilled below {rb}"
    # ... also when a live ray's key is all ones (few direction cells, origin in the top Morton cell): dead slots must still sort last,
    # or the refilling warps, which stop at the first dead key, leave live rays untraced
    for dc, sh, mb in ((1, 6, 1), (2, 5, 2), (4, 4, 3), (2, 6, 1), (4, 6, 1)):
        for rb in (1, 32):
            r.set_path_options(mode=PATH_WAVEFRONT, batch_samples=4, dir_cells=dc, origin_cell_shift=sh, morton_bits_per_axis=mb, refill_below=rb)
            assert _accum(r, 0, 6).tobytes() == want_acc.tobytes(), f"accum differs with key ({dc}, {sh}, {mb}), lanes refilled below {rb}"
    r.set_path_options(mode=PATH_WAVEFRONT, batch_samples=4, dir_cells=8, origin_cell_shift=6, morton_bits_per_axis=2, refill_below=32)
    r.stats_enable(True)
    r.reset_accum(); r.trace_path(0, 6); r.sync()
    refill_counts = r.stats_read()
    r.set_path_options(refill_below=0)
    r.stats_enable(True)   # zeroes the counters
    r.reset_accum(); r.trace_path(0, 6); r.sync()
    assert r.stats_read() == refill_counts   # same rays, same DDA steps
    r.stats_enable(False)
    with pytest.raises(blok_b200.BlokbError):
        r.set_path_options(refill_below=33)
    # a sample range that does not start at 0, accumulated on top of an earlier call
    r.set_path_options(mode=PATH_FUSED)
    r.reset_accum(); r.trace_path(0, 2); r.trace_path(2, 5)
    want2 = r.readback(IMG_ACCUM).copy()
    r.set_path_options(mode=PATH_WAVEFRONT, batch_samples=2)
    r.reset_accum(); r.trace_path(0, 2); r.trace_path(2, 5)
    assert r.readback(IMG_ACCUM).tobytes() == want2.tobytes()


def test_wavefront_on_terrain_odd_frame_size_and_strips(renderer_factory):
    """Frame sizes that are not multiples of the 32x4 block, sky + ground, emissive and metal materials, strips of a tiled frame."""
    xyz, m, mats = scenes.terrain(96)
    w = WorldSvoGpu.from_voxels(xyz, m, mats)
    r = renderer_factory(150, 93)
    r.add_world(w)
    ubo = blok_b200.Camera((48, 60, -20), 90.0, -25.0).frame(150, 93, 3, 5)
    r.set_frame(ubo)
    r.set_path_options(mode=PATH_FUSED)
    want = _accum(r, 0, 5)
    r.stats_enable(True)
    r.reset_accum(); r.trace_path(0, 5); r.sync()
    fused_counts = r.stats_read()
    r.stats_enable(False)
    r.set_path_options(mode=PATH_WAVEFRONT, batch_samples=3)
    assert _accum(r, 0, 5).tobytes() == want.tobytes()
    r.stats_enable(True)
    r.reset_accum(); r.trace_path(0, 5); r.sync()
    assert r.stats_read() == fused_counts  # same rays, same DDA steps
    r.stats_enable(False)
    r.reset_accum()
    for g in range(3):
        r.set_strips(g, 3)
        r.trace_path(0, 5)
    r.set_strips(0, 1)
    assert r.readback(IMG_ACCUM).tobytes() == want.tobytes()


def test_path_options_validation(renderer_factory):
    r = renderer_factory()
    with pytest.raises(blok_b200.BlokbError):
        r.set_path_options(mode=7)
    with pytest.raises(blok_b200.BlokbError):
        r.set_path_options(dir_cells=17)
    o = r.set_path_options(mode=PATH_WAVEFRONT, batch_samples=0, dir_cells=0, origin_cell_shift=0, morton_bits_per_axis=0)
    assert (o["batchSamples"], o["dirCells"], o["originCellShift"], o["mortonBitsPerAxis"]) == (8, 8, 6, 2)
