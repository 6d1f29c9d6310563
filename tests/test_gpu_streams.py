"""Stream ordering across the C-ABI (include/blokb200.h "Streams"): dispatches take a caller stream, while readback / write_image /
stats_read / uploads run on the context's own stream — which must wait for what was enqueued on the caller's streams, so that
"dispatch on my stream, then blokb_readback" needs no synchronisation by the caller."""
import os

import numpy as np
import pytest

import blok_b200
from blok_b200 import IMG_ACCUM, IMG_COLOR, IMG_HIT, WorldSvoGpu, scenes

pytestmark = pytest.mark.gpu


def test_readback_is_ordered_after_dispatches_on_caller_streams(renderer_factory):
    import torch
    xyz, m, mats = scenes.terrain(160)
    w = WorldSvoGpu.from_voxels(xyz, m, mats)
    W, H = 1024, 576
    r = renderer_factory(W, H)
    r.add_world(w)
    cam = blok_b200.Camera((80, 110, -40), 90.0, -25.0)
    r.set_frame(cam.frame(W, H, 0, 4))
    r.trace_primary(); r.reset_accum(); r.trace_path(0, 4); r.resolve(); r.sync()
    want_hit, want_acc, want_col = r.readback(IMG_HIT).copy(), r.readback(IMG_ACCUM).copy(), r.readback(IMG_COLOR).copy()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    spin = torch.empty(64 << 20, dtype=torch.float32, device="cuda")
    for _ in range(3):
        zero16 = np.zeros((H, W, 4), np.float32)
        r.write_image(IMG_HIT, np.zeros((H, W), blok_b200.HIT_DT)); r.write_image(IMG_ACCUM, zero16); r.write_image(IMG_COLOR, zero16)
        with torch.cuda.stream(s1):
            for _ in range(20):
                spin.add_(1.0)   # keeps s1 busy so that the dispatches below are still queued when readback is called
        r.trace_primary(stream=s1.cuda_stream)
        got_hit = r.readback(IMG_HIT)     # no synchronisation by the caller
        assert got_hit.tobytes() == want_hit.tobytes()
        with torch.cuda.stream(s2):
            for _ in range(20):
                spin.add_(1.0)
        r.reset_accum(stream=s2.cuda_stream)
        r.trace_path(0, 4, stream=s2.cuda_stream)
        r.resolve(stream=s2.cuda_stream)
        assert r.readback(IMG_COLOR).tobytes() == want_col.tobytes()
        assert r.readback(IMG_ACCUM).tobytes() == want_acc.tobytes()
    # blokb_sync covers caller streams too
    with torch.cuda.stream(s1):
        for _ in range(20):
            spin.add_(1.0)
    r.trace_primary(stream=s1.cuda_stream)
    r.sync()
    assert s1.query()


def test_wavefront_dispatches_on_two_streams_do_not_race_on_the_shared_scratch(renderer_factory):
    """blokb_render and blokb_trace_path of one context use the same slot-array scratch (wavefront mode, the default). Issued on two
    different caller streams with nothing ordering them, the library serialises them on the scratch: both images equal
    the ones produced one after the other."""
    import torch
    xyz, m, mats = scenes.terrain(160)
    w = WorldSvoGpu.from_voxels(xyz, m, mats)
    W, H = 1024, 576
    r = renderer_factory(W, H)
    r.add_world(w)
    cam = blok_b200.Camera((80, 110, -40), 90.0, -25.0)
    r.set_frame(cam.frame(W, H, 0, 8))
    r.render(); r.sync()
    want_col = r.readback(IMG_COLOR).copy()
    r.reset_accum(); r.trace_path(0, 8); r.sync()
    want_acc = r.readback(IMG_ACCUM).copy()
    a, b = torch.cuda.Stream(), torch.cuda.Stream()
    zero16 = np.zeros((H, W, 4), np.float32)
    for _ in range(3):
        r.write_image(IMG_COLOR, zero16); r.write_image(IMG_ACCUM, zero16)
        torch.cuda.synchronize()
        r.render(stream=a.cuda_stream)                 # writes colour through the scratch ...
        r.trace_path(0, 8, stream=b.cuda_stream)       # ... while this one, on another stream, wants the same scratch for accum
        r.render(stream=a.cuda_stream)
        assert r.readback(IMG_COLOR).tobytes() == want_col.tobytes()
        assert r.readback(IMG_ACCUM).tobytes() == want_acc.tobytes()
