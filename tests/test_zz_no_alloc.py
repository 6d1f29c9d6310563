"""blokb_trace_rays_reordered_static: the reordered ray-list trace for hosts that may not allocate device memory.

Who needs it: an engine that fixes its device-memory budget at start-up and gives the rest of the GPU to its own pool (a cudaMalloc
coming out of this library later would fail, or eat into that budget), and a frame loop that cannot afford the device-wide stall of
the cudaDeviceSynchronize + cudaFree + cudaMalloc with which blokb_trace_rays_reordered grows its scratch. The static variant keeps its
scratch in a fixed array inside the library's device image (one per GPU, shared by every call) and is limited to
BLOKB_STATIC_MAX_RAYS rays per call.

What is checked, through the C-ABI like every other GPU test: its records equal those of blokb_trace_rays_reordered (and of the
unordered blokb_trace_rays) byte for byte — the tolerance the hit-record tests use everywhere — on consecutive calls that reuse the
shared array with lists of different length and content, so counters left over from the previous call would show; and lists beyond
the limit are refused with a message that names it.

This file sorts last on purpose (pytest -x stops at the first failure; a failure here must not hide the other tests) and must pass
both alone in a fresh process — where its first call is also the library's first look-up of the array's address — and after the others.
"""
import glob
import os

import numpy as np
import pytest

from blok_b200 import WorldSvoGpu
from tests.conftest import make_rays

pytestmark = pytest.mark.gpu

STATIC_MAX_RAYS = 1 << 21   # BLOKB_STATIC_MAX_RAYS in include/blokb200.h
GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def _world(path):
    z = np.load(path)
    return WorldSvoGpu(z["nodes"], z["subs"], z["mats"])


def _ck(r, rc):
    assert rc == 0, (rc, r.L.blokb_last_error())


def test_static_scratch_trace_matches_the_allocating_entry_point_call_after_call(renderer_factory):
    import torch
    r = renderer_factory()
    r.add_world(_world(GOLD[0]))
    s = r.world_stats()
    lo = np.array(s["gridOrigin"], np.float32) - 20
    hi = np.array(s["gridOrigin"], np.float32) + np.array(s["gridDim"], np.float32) + 20
    rng = np.random.default_rng(2026)
    fn = r.L.blokb_trace_rays_reordered_static
    hits = 0
    # back-to-back calls on the context's stream, long list first: the second and third calls meet the first one's counters and order
    for n in (100_003, 257, 1, 50_000):
        rays = make_rays(rng, max(n, 64), lo, hi)[:n]
        d = torch.from_numpy(rays).cuda()
        want = torch.full((n, 4), 7, dtype=torch.int32, device="cuda")
        plain = torch.full((n, 4), 8, dtype=torch.int32, device="cuda")
        got = torch.full((n, 4), 9, dtype=torch.int32, device="cuda")
        r.trace_rays_device(d.data_ptr(), n, want.data_ptr(), reorder=True)   # the existing entry point (allocates its scratch)
        r.trace_rays_device(d.data_ptr(), n, plain.data_ptr())
        _ck(r, fn(r._ctx, d.data_ptr(), n, 0, got.data_ptr(), None))
        r.sync()
        assert torch.equal(got, want), f"n={n}: {int((got != want).any(1).sum())} records differ"
        assert torch.equal(got, plain)
        hits += int((got[:, 0] != -1).sum())
    assert hits > 1000
    # the same list twice in a row gives the same bytes (nothing survives from call to call that the next one trusts)
    again = torch.full((n, 4), 5, dtype=torch.int32, device="cuda")
    _ck(r, fn(r._ctx, d.data_ptr(), n, 0, again.data_ptr(), None))
    r.sync()
    assert torch.equal(again, want)
    # a second context on the same GPU uses the same array; on a caller stream the caller orders and waits for the work itself
    r2 = renderer_factory()
    r2.add_world(_world(GOLD[0]))
    q = torch.cuda.Stream()
    q.wait_stream(torch.cuda.current_stream())
    other = torch.full((n, 4), 3, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    _ck(r2, fn(r2._ctx, d.data_ptr(), n, 0, other.data_ptr(), q.cuda_stream))
    q.synchronize()
    assert torch.equal(other, want)
    _ck(r, fn(r._ctx, 0, 0, 0, 0, None))   # empty list: no-op


def test_static_scratch_trace_refuses_lists_beyond_its_limit(renderer_factory):
    import torch
    r = renderer_factory()
    r.add_world(_world(GOLD[0]))
    buf = torch.zeros(64, 8, dtype=torch.float32, device="cuda")
    out = torch.full((64, 4), 9, dtype=torch.int32, device="cuda")
    fn = r.L.blokb_trace_rays_reordered_static
    rc = fn(r._ctx, buf.data_ptr(), STATIC_MAX_RAYS + 1, 0, out.data_ptr(), None)   # refused before anything is read or launched
    assert rc == -1 and b"BLOKB_STATIC_MAX_RAYS" in r.L.blokb_last_error()
    r.sync()
    assert bool((out == 9).all())
    assert fn(r._ctx, 0, 5, 0, out.data_ptr(), None) == -1   # null ray pointer
    # exactly the limit is accepted and fills the array to its last word
    s = r.world_stats()
    lo = np.array(s["gridOrigin"], np.float32) - 20
    hi = np.array(s["gridOrigin"], np.float32) + np.array(s["gridDim"], np.float32) + 20
    base = make_rays(np.random.default_rng(7), 1 << 16, lo, hi)
    d = torch.from_numpy(base).cuda().repeat(STATIC_MAX_RAYS >> 16, 1)
    n = d.shape[0]
    assert n == STATIC_MAX_RAYS
    want = torch.full((n, 4), 7, dtype=torch.int32, device="cuda")
    got = torch.full((n, 4), 9, dtype=torch.int32, device="cuda")
    r.trace_rays_device(d.data_ptr(), n, want.data_ptr())
    _ck(r, fn(r._ctx, d.data_ptr(), n, 0, got.data_ptr(), None))
    r.sync()
    assert torch.equal(got, want)
