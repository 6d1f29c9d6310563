"""GPU: parity of the CUDA path (through the C-ABI) with the oracle and with the reference's golden vectors."""
import ctypes as C
import glob
import os

import numpy as np
import pytest

import blok_b200
from blok_b200 import (IMG_ACCUM, IMG_ALBEDOMETAL, IMG_COLOR, IMG_HIT, IMG_MOTION, IMG_NORMALROUGH, IMG_SHADOW, IMG_WORLDPOS,
                       Camera, WorldSvoGpu, scenes)
from tests.conftest import make_rays, to_scene

pytestmark = pytest.mark.gpu
GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def _gold(path):
    z = np.load(path)
    return z, WorldSvoGpu(z["nodes"], z["subs"], z["mats"])


def _check_hits(out, h, t):
    """out: HIT_DT from the GPU; h (n,4) {prim,node,mat,face}, t (n,) from the oracle. Bit-exact."""
    assert np.array_equal(out["node"], h[:, 1])
    assert np.array_equal(out["material"], h[:, 2])
    assert np.array_equal(out["t"].view(np.uint32), t.view(np.uint32))
    hit = h[:, 1] != 0xFFFFFFFF
    assert np.array_equal(out["facePrim"][hit] & 7, h[hit, 3]) and np.array_equal(out["facePrim"][hit] >> 3, h[hit, 0])
    assert np.all(out["facePrim"][~hit] == 0xFFFFFFFF)


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[:-4])
def test_primary_hits_match_reference_goldens(renderer_factory, oracle, path):
    z, w = _gold(path)
    r = renderer_factory()
    r.add_world(w)
    r.trace_primary(z["ubo_s"])
    out = r.readback(IMG_HIT).reshape(-1)
    g = z["hits_s"]
    _check_hits(out, g[:, :4], g[:, 4].view(np.float32))
    r.trace_primary(z["ubo_l"])
    out = r.readback(IMG_HIT).reshape(-1)
    miss = out["node"] == 0xFFFFFFFF
    rec = np.stack([np.where(miss, 0xFFFFFFFF, out["facePrim"] >> 3), out["node"], out["material"],
                    np.where(miss, 0xFFFFFFFF, out["facePrim"] & 7), out["t"].view(np.uint32)], 1).astype(np.uint32)
    assert int(oracle.fnv(rec), 16) == int(z["fnv_l"][0])
    assert int((~miss).sum()) == int(z["nhit_l"][0])


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[:-4])
def test_render_matches_reference_goldens(renderer_factory, path):
    """Full raygen (4 spp, 2 bounces, sun shadow) against the reference's images. IDs/geometry exact; radiance differs
    only through CUDA's sinf/cosf/powf vs glibc's (<= 2 ulp per call): tolerance 2e-3 relative on 99.5% of pixels
    (a 1-ulp change can flip a Russian-roulette / lobe-choice branch on a rare pixel)."""
    z, w = _gold(path)
    r = renderer_factory()
    r.add_world(w)
    r.render(z["ubo_r"])
    wp = r.readback(IMG_WORLDPOS)
    assert np.array_equal(wp[..., 3].view(np.uint32), z["img_worldPos"][..., 3].view(np.uint32))  # first-hit t: bit-exact
    np.testing.assert_allclose(wp[..., :3], z["img_worldPos"][..., :3], rtol=0, atol=0)           # same unfused o + d*t
    nr = r.readback(IMG_NORMALROUGH).astype(np.float32)
    np.testing.assert_array_equal(nr, z["img_normalRough"].astype(np.float16).astype(np.float32))
    am = r.readback(IMG_ALBEDOMETAL).astype(np.int32)
    want = np.rint(np.clip(z["img_albedoMetal"], 0, 1) * 255).astype(np.int32)
    assert np.abs(am - want).max() <= 1  # unorm8 rounding of sky colour (powf) may differ by one step
    mv = r.readback(IMG_MOTION).astype(np.float32)
    np.testing.assert_allclose(mv, z["img_motion"][..., :2], atol=2e-3)
    col = r.readback(IMG_COLOR)
    ref = z["img_color"]
    assert np.all(col[..., 3] == 1.0)
    rel = np.abs(col[..., :3] - ref[..., :3]) / (np.abs(ref[..., :3]) + 1e-3)
    assert np.mean(rel.max(-1) < 2e-3) > 0.995
    assert abs(col[..., :3].mean() - ref[..., :3].mean()) < 2e-3 * ref[..., :3].mean()


def test_random_and_degenerate_rays_synthetic(renderer_factory, oracle):
    rng = np.random.default_rng(5)
    xyz, m, mats = scenes.terrain(256, origin=(-64, 0, -32))
    w = WorldSvoGpu.from_voxels(xyz, m, mats)
    r = renderer_factory()
    r.add_world(w)
    rays = make_rays(rng, 200000, -90, 260)
    out = r.trace_rays(rays)
    h, t, cnt = oracle.trace(to_scene(w), rays, threads=8)
    _check_hits(out, h, t)
    assert cnt[2] == 0
    # shadow-ray rules (terminate on first hit): occlusion must agree
    out2 = r.trace_rays(rays, any_hit=True)
    assert np.array_equal(out2["node"] != 0xFFFFFFFF, t >= 0)


def test_primary_shadow_config2_style(renderer_factory, oracle):
    """Primary + sun shadow rays on a dense blob 'mesh' (config 2 at test scale): shadow flags vs oracle shadow rays."""
    xyz, m, mats = scenes.blob_mesh(128)
    w = WorldSvoGpu.from_voxels(xyz, m, mats)
    r = renderer_factory(160, 96)
    r.add_world(w)
    f = Camera((-0.4 * 128, 0.9 * 128, -0.4 * 128), 45.0, -25.0).frame(160, 96, 0, 1)
    r.trace_primary(f, shadow=True)
    hits = r.readback(IMG_HIT).reshape(-1)
    sh = r.readback(IMG_SHADOW).reshape(-1)
    sc = to_scene(w)
    rays = oracle.primary_rays(f.to_numpy())
    h, t, _ = oracle.trace(sc, rays, threads=8)
    _check_hits(hits, h, t)
    hit = t >= 0
    assert hit.sum() > 1000
    # rebuild the shadow rays exactly like raygen.rgen:280-298 (fp32, unfused)
    normals = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], np.float32)
    N = normals[np.where(hit, h[:, 3], 0)]
    d = rays[:, 3:6]
    flip = ((N[:, 0] * d[:, 0] + N[:, 1] * d[:, 1]) + N[:, 2] * d[:, 2]) > 0
    N = np.where(flip[:, None], -N, N)
    s = np.array([0.5, 0.8, 0.3], np.float32)
    inv = np.float32(1.0) / np.sqrt((s[0] * s[0] + s[1] * s[1]) + s[2] * s[2], dtype=np.float32)
    sun = (s * inv).astype(np.float32)
    ndl = (N[:, 0] * sun[0] + N[:, 1] * sun[1]) + N[:, 2] * sun[2]
    cast = hit & (ndl > 0)
    o = (rays[:, 0:3] + d * t[:, None]).astype(np.float32) + (N * np.float32(0.001)).astype(np.float32)
    srays = np.concatenate([o, np.tile(sun, (len(o), 1)), np.full((len(o), 1), 0.001, np.float32), np.full((len(o), 1), 1000, np.float32)], 1).astype(np.float32)
    _, st, _ = oracle.trace(sc, srays[cast], any_hit=True, threads=8)
    want = np.zeros(len(o), np.uint8)
    want[cast] = np.where(st >= 0, 2, 1)
    assert np.array_equal(sh, want)
    assert (want == 2).sum() > 50 and (want == 1).sum() > 50


def test_path_accumulate_sharded_equals_render(renderer_factory):
    """trace_path over any partition of the sample range + resolve == render's colour (fp32 sum order only), and the
    accumulator carries (sum.rgb, spp) like the reference's CUDA tracer."""
    z, w = _gold(GOLD[0])
    r = renderer_factory()
    r.add_world(w)
    ubo = blok_b200.FrameUBO.from_bytes(z["ubo_r"])
    ubo.sampleCount = 8
    r.render(ubo)
    want = r.readback(IMG_COLOR)
    r.reset_accum()
    for a, b in ((0, 3), (3, 4), (4, 8)):
        r.trace_path(a, b)
    acc = r.readback(IMG_ACCUM)
    assert np.all(acc[..., 3] == 8.0)
    r.resolve()
    got = r.readback(IMG_COLOR)
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)
    r.reset_accum()
    assert np.all(r.readback(IMG_ACCUM) == 0)


def test_empty_world_and_errors(renderer_factory):
    r = renderer_factory()
    with pytest.raises(blok_b200.BlokbError) as e:
        r.trace_rays(np.zeros((1, 8), np.float32))
    assert e.value.code == -4  # no world yet
    mats = scenes.default_materials(4)
    r.add_world(WorldSvoGpu.from_voxels(np.zeros((0, 3), np.int32), np.zeros(0, np.uint32), mats))
    out = r.trace_rays(np.array([[0, 0, 0, 0, 0, 1, 0.001, 100]], np.float32))
    assert out["node"][0] == 0xFFFFFFFF and out["t"][0] == -1.0
    r.render(Camera().frame(32, 16, 0, 2))
    col = r.readback(IMG_COLOR)
    assert col.shape == (16, 32, 4) and np.all(np.isfinite(col)) and col[..., :3].min() > 0.3  # sky everywhere
    assert r.trace_rays(np.zeros((0, 8), np.float32)).shape == (0,)
    bad = WorldSvoGpu.from_voxels([[1, 1, 1]], [1], mats)
    bad.globalSubChunks["startDepth"][0] = 2
    with pytest.raises(blok_b200.BlokbError) as e:
        r.add_world(bad)
    assert e.value.code == -3


def test_full_size_properties_terrain(renderer_factory, oracle):
    """Config-3-shaped scene at 1024^3 / 1920x1080: too big for a full oracle pass, so (a) a strided pixel subset goes
    through the oracle bit-exactly, (b) every reported hit satisfies the invariants: the reported leaf is filled in the
    reference node array with the reported material, and t is inside (tmin, tmax)."""
    xyz, m, mats = scenes.terrain(1024)
    w = WorldSvoGpu.from_voxels(xyz, m, mats)
    r = renderer_factory(1920, 1080)
    r.add_world(w)
    f = Camera((512, 0.6 * 1024, -0.2 * 1024), 90.0, -25.0).frame(1920, 1080, 0, 1)
    r.trace_primary(f)
    out = r.readback(IMG_HIT).reshape(-1)
    hit = out["node"] != 0xFFFFFFFF
    assert hit.mean() > 0.3
    leaf = w.globalNodes[out["node"][hit]]
    assert np.all(leaf["occupancy"] > 0) and np.all(leaf["childMask"] == 0)
    assert np.array_equal(leaf["materialId"], out["material"][hit])
    assert np.all(out["t"][hit] > 0.001) and np.all(out["t"][hit] < 10000)
    rays = oracle.primary_rays(f.to_numpy())
    idx = np.arange(0, len(rays), 97)
    h, t, cnt = oracle.trace(to_scene(w), rays[idx], threads=8)
    _check_hits(out[idx], h, t)
    assert cnt[2] == 0


def test_frame_pipeline_matches_blocking_path(renderer_factory):
    """Frames in flight: records delivered by the pipeline == trace_primary + readback, frame by frame."""
    z, w = _gold(GOLD[2])
    r = renderer_factory(160, 96)
    r.add_world(w)
    cams = [Camera((-60 + 3 * i, 40, 30 - 2 * i), 0.0 + 2 * i, -10.0) for i in range(5)]
    pipe = blok_b200.FramePipeline(r, 160, 96, depth=2)
    tickets = []
    got = []
    for c in cams:
        tickets.append(pipe.submit(c.frame(160, 96, 0, 1)))
        if len(tickets) > 1:
            got.append(pipe.result(tickets[-2]).copy())
    got.append(pipe.result(tickets[-1]).copy())
    pipe.drain()
    with pytest.raises(blok_b200.BlokbError):
        pipe.result(0)  # no longer in flight
    pipe.close()
    for c, g in zip(cams, got):
        r.trace_primary(c.frame(160, 96, 0, 1))
        want = r.readback(IMG_HIT)
        assert np.array_equal(g.view(np.uint32), want.view(np.uint32))
    assert sum(int((g["node"] != 0xFFFFFFFF).sum()) for g in got) > 1000


def test_packed_frame_pipeline_loses_nothing(renderer_factory):
    """blokb_pipeline_*_packed: the 8-byte records of k_primary_packed, unpacked on the host against the uploaded arrays,
    are byte for byte the 16-byte records of trace_primary; frame sizes that are not a multiple of the 32x8 block."""
    z, w = _gold(GOLD[2])
    r = renderer_factory(150, 91)
    r.add_world(w)
    cams = [Camera((-60 + 3 * i, 40, 30 - 2 * i), 0.0 + 2 * i, -10.0) for i in range(4)]
    pipe = blok_b200.FramePipeline(r, 150, 91, depth=2, packed=True)
    assert pipe.nbytes == 150 * 91 * 8
    tickets = [pipe.submit(c.frame(150, 91, 0, 1)) for c in cams[:2]]
    got = [pipe.result(tickets[0]).copy()]
    for c in cams[2:]:
        tickets.append(pipe.submit(c.frame(150, 91, 0, 1)))
        got.append(pipe.result(tickets[-2]).copy())
    got.append(pipe.result(tickets[-1]).copy())
    pipe.drain()
    pipe.close()
    plain = blok_b200.FramePipeline(r, 150, 91, depth=1)
    f0 = cams[0].frame(150, 91, 0, 1)
    rc = plain.L.blokb_pipeline_submit_packed(plain._p, C.byref(f0), plain._host[0], None)  # record formats do not mix
    assert rc == -1 and b"packed" in plain.L.blokb_last_error()
    plain.close()
    hits = 0
    for c, g in zip(cams, got):
        r.trace_primary(c.frame(150, 91, 0, 1))
        want = r.readback(IMG_HIT)
        assert np.array_equal(w.unpack_hits(g).view(np.uint32), want.view(np.uint32))
        hits += int((want["node"] != 0xFFFFFFFF).sum())
    assert hits > 1000


def test_reordered_ray_list_gives_the_same_records(renderer_factory):
    """blokb_trace_rays_reordered (sort by direction cell + origin Morton code, trace, scatter back) == blokb_trace_rays,
    byte for byte, on incoherent rays: random directions from random origins inside and outside the grid, degenerate
    families included, list sizes that are not a multiple of anything."""
    import torch
    z, w = _gold(GOLD[0])
    r = renderer_factory()
    r.add_world(w)
    s = r.world_stats()
    lo = np.array(s["gridOrigin"], np.float32) - 20
    hi = np.array(s["gridOrigin"], np.float32) + np.array(s["gridDim"], np.float32) + 20
    rng = np.random.default_rng(11)
    for n in (1, 257, 100_003):
        rays = make_rays(rng, max(n, 64), lo, hi)[:n]
        d = torch.from_numpy(rays).cuda()
        a = torch.full((n, 4), 7, dtype=torch.int32, device="cuda")
        b = torch.full((n, 4), 9, dtype=torch.int32, device="cuda")
        r.trace_rays_device(d.data_ptr(), n, a.data_ptr())
        r.trace_rays_device(d.data_ptr(), n, b.data_ptr(), reorder=True)
        r.sync()
        assert torch.equal(a, b)
        assert int((a[:, 0] != -1).sum()) > 0 or n == 1
    r.trace_rays_device(0, 0, 0, reorder=True)  # empty list: no-op


def test_material_ids_beyond_the_table_read_as_zero_records(renderer_factory):
    """Found by tools/fuzz_gpu.py: a voxel whose material id has no entry in the uploaded table made the shading kernels read
    past the table. The reference's behaviour is undefined there (storage-buffer overrun); the product reads an all-zero record
    (robustBufferAccess semantics): the frame renders, geometry images are unaffected, albedo is black."""
    xyz, mid, mats = scenes.terrain(96)
    r = renderer_factory(64, 64)
    cam = Camera((48.0, 70.0, -30.0), 90.0, -25.0)
    f = cam.frame(64, 64, 0, 4)
    r.add_voxels(xyz, mid, mats)
    r.render(f)
    wp = r.readback(IMG_WORLDPOS).copy()
    r.add_voxels(xyz, (mid + 70000).astype(np.uint32), mats)   # clamps to 65535 in hit.rchit:62 — far beyond the 17 entries
    r.render(f)
    r.sync()
    assert r.readback(IMG_WORLDPOS).tobytes() == wp.tobytes()
    am = r.readback(IMG_ALBEDOMETAL)
    hit = wp[..., 3] < 9999.0
    assert hit.any() and np.all(am[hit][:, :3] == 0)
    assert np.isfinite(r.readback(IMG_COLOR)).all()
    r.reset_accum(); r.trace_path(0, 2); r.resolve(); r.sync()
