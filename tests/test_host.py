"""CPU: host logic — C-ABI surface, scene upload layout, world builder, camera, sharding — and the DDA's logic through
tests/host_emul (a CPU compile of the very device functions the kernels run). No GPU compute here."""
import ctypes as C
import glob
import os
import re

import numpy as np
import pytest

import blok_b200
from blok_b200 import WorldSvoGpu, scenes, sharding
from blok_b200.api import EXPORTS, LIB_PATH, FrameUBO
from oracle.pyoracle import Scene
from tests.conftest import ROOT, make_rays, to_scene

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def _gold(path):
    z = np.load(path)
    return z, WorldSvoGpu(z["nodes"], z["subs"], z["mats"])


def test_cabi_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "blokb200.h")).read()
    declared = set(re.findall(r"\b(blokb_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(EXPORTS), declared ^ set(EXPORTS)
    lib = C.CDLL(LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), name


def test_cabi_header_is_plain_c():
    """The drop-in boundary is a C ABI: the header must compile as C11 on its own (no C++ types, no torch, no CUDA headers)."""
    import subprocess
    hdr = os.path.join(ROOT, "include", "blokb200.h")
    r = subprocess.run(["gcc", "-x", "c", "-std=c11", "-fsyntax-only", "-Wall", "-Werror", hdr], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    code = re.sub(r"/\*.*?\*/", "", open(hdr).read(), flags=re.S)   # comments may mention torch / std:: as context
    assert "torch" not in code and "cuda_runtime" not in code and "std::" not in code and "#include <cuda" not in code


def test_no_cpu_fallback_without_device():
    """On a box without a GPU the product must fail loudly, not compute on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(blok_b200.BlokbError) as e:
        blok_b200.Renderer(64, 64)
    assert e.value.code == -2


def test_product_does_not_reference_oracle():
    for path in glob.glob(os.path.join(ROOT, "blok_b200", "**", "*"), recursive=True):
        if os.path.isfile(path) and path.endswith((".py", ".cpp", ".cu", ".cuh", ".h")):
            src = open(path).read()
            assert "pyoracle" not in src and "libblokoracle" not in src and "libblokref" not in src and "host_emul/" not in src.replace("tests/host_emul", ""), path


def test_worldgen_matches_reference_builder(reflib):
    xyz, m, mats = scenes.blob_mesh(64)
    w = WorldSvoGpu.from_voxels(xyz, m, mats)
    s = reflib.scene_from_voxels(xyz, m, mats)
    assert np.array_equal(w.globalNodes, s.nodes) and np.array_equal(w.globalSubChunks, s.subs)  # single chunk: bit-identical
    xyz, m, mats = scenes.terrain(192, origin=(-100, -20, -50))  # 4+ chunks incl. negative coordinates
    w = WorldSvoGpu.from_voxels(xyz, m, mats)
    s = reflib.scene_from_voxels(xyz, m, mats)

    def chunks(nodes, subs):
        d = {}
        for off in np.unique(subs["nodeOffset"]):
            ss = subs[subs["nodeOffset"] == off].copy()
            key = tuple(np.floor(ss["worldMin"].min(0) / 128).astype(int))
            cnt = ss["nodeCount"][0]
            ss["nodeOffset"] = 0
            d[key] = (nodes[off:off + cnt].tobytes(), ss.tobytes())
        return d
    a, b = chunks(w.globalNodes, w.globalSubChunks), chunks(s.nodes, s.subs)
    assert sorted(a) == sorted(b) and all(a[k] == b[k] for k in a)  # identical up to the unordered_map chunk order


def test_worldgen_last_write_wins_and_empty():
    mats = scenes.default_materials(4)
    w = WorldSvoGpu.from_voxels([[1, 2, 3], [1, 2, 3]], [1, 2], mats)
    leaf = w.globalNodes[w.globalNodes["occupancy"] > 0]
    assert len(leaf) == 1 and leaf["materialId"][0] == 2
    w = WorldSvoGpu.from_voxels(np.zeros((0, 3), np.int32), np.zeros(0, np.uint32), mats)
    assert len(w.globalNodes) == 0 and len(w.globalSubChunks) == 0


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[:-4])
def test_dda_logic_matches_reference_goldens(emul, oracle, path):
    """The compact-layout DDA (host compile of traverse.cuh) on reference-built worlds: hit records bit-exact."""
    z, w = _gold(path)
    e = emul.world(w)
    assert e.ok(), e.err
    rays = oracle.primary_rays(z["ubo_s"])
    g = z["hits_s"]
    hit = g[:, 0] != 0xFFFFFFFF
    for locate in (True, False):  # both start-cell strategies of Walker::init
        out, _ = e.trace(rays, locate=locate)
        assert np.array_equal(out["node"], g[:, 1]) and np.array_equal(out["material"], g[:, 2])
        assert np.array_equal(out["t"].view(np.uint32), g[:, 4])
        assert np.array_equal(out["facePrim"][hit] & 7, g[hit, 3]) and np.array_equal(out["facePrim"][hit] >> 3, g[hit, 0])
    # 512x512 via the checksum of the reference's records
    rays = oracle.primary_rays(z["ubo_l"])
    out, _ = e.trace(rays)
    miss = out["node"] == 0xFFFFFFFF
    rec = np.stack([np.where(miss, 0xFFFFFFFF, out["facePrim"] >> 3), out["node"], out["material"],
                    np.where(miss, 0xFFFFFFFF, out["facePrim"] & 7), out["t"].view(np.uint32)], 1).astype(np.uint32)
    assert int(oracle.fnv(rec), 16) == int(z["fnv_l"][0])


def test_dda_logic_random_and_degenerate_rays(emul, oracle):
    rng = np.random.default_rng(3)
    xyz, m, mats = scenes.terrain(192, origin=(-64, 0, -32))
    w = WorldSvoGpu.from_voxels(xyz, m, mats)
    e = emul.world(w)
    rays = make_rays(rng, 60000, -90, 200)
    h, t, _ = oracle.trace(to_scene(w), rays, threads=8)
    for locate in (True, False):
        out, _ = e.trace(rays, locate=locate)
        assert np.array_equal(out["node"], h[:, 1]) and np.array_equal(out["t"].view(np.uint32), t.view(np.uint32))
        assert np.array_equal(out["material"], h[:, 2])


def _pack(hits):
    """HIT_DT -> PACKED_DT the way k_primary_packed stores it (kernels.cu)."""
    from blok_b200.api import PACKED_DT
    p = np.empty(hits.shape, PACKED_DT)
    miss = hits["node"] == 0xFFFFFFFF
    p["nodeFace"] = np.where(miss, 0xFFFFFFFF, hits["node"] | (hits["facePrim"] & 7) << 29)
    p["t"] = hits["t"]
    return p


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[:-4])
def test_packed_records_unpack_to_the_full_records(emul, oracle, path):
    """blokb_node_prims + blokb_unpack_hits (host side of the packed readback): the 8-byte record loses nothing —
    material and gl_PrimitiveID come back from the uploaded arrays, equal to the reference's own records."""
    z, w = _gold(path)
    e = emul.world(w)
    out, _ = e.trace(oracle.primary_rays(z["ubo_s"]))
    assert (out["node"] != 0xFFFFFFFF).sum() > 100 and (out["node"] == 0xFFFFFFFF).sum() > 0
    back = w.unpack_hits(_pack(out))
    assert np.array_equal(back.view(np.uint32), out.view(np.uint32))
    g = z["hits_s"]  # the reference's records: prim, node, material, face, t bits
    hit = g[:, 0] != 0xFFFFFFFF
    assert np.array_equal(back["facePrim"][hit] >> 3, g[hit, 0]) and np.array_equal(back["material"], g[:, 2])
    no_prim = w.unpack_hits(_pack(out), with_prim=False)
    assert np.array_equal(no_prim["facePrim"][hit], g[hit, 3]) and np.all(no_prim["facePrim"][~hit] == 0xFFFFFFFF)
    # every node below a sub-chunk root maps to exactly that sub-chunk; roots map to themselves
    prims = w.node_prims()
    subs = w.globalSubChunks
    roots = (subs["nodeOffset"].astype(np.int64) + subs["rootNodeIndex"]).astype(np.int64)
    assert np.array_equal(prims[roots], np.arange(len(subs), dtype=np.uint32))
    # a record naming a node outside the array is refused (and stored as a miss)
    from blok_b200.api import PACKED_DT
    bad = np.array([(len(w.globalNodes), 1.0)], PACKED_DT)
    with pytest.raises(blok_b200.BlokbError):
        w.unpack_hits(bad)
    assert len(w.unpack_hits(np.zeros(0, PACKED_DT))) == 0
    # batches large enough to be split over the host's cores: same records, and a bad one anywhere is still reported
    big = np.tile(_pack(out), 400_000 // len(out) + 1)
    assert len(big) > 4 * (1 << 16)
    want = np.tile(out, 400_000 // len(out) + 1)
    assert np.array_equal(w.unpack_hits(big).view(np.uint32), want.view(np.uint32))
    big[-5] = (len(w.globalNodes), 1.0)
    with pytest.raises(blok_b200.BlokbError):
        w.unpack_hits(big)


def test_walk_profile_counts_the_same_walk(emul, oracle):
    """tools/walk_profile.py's per-level counters restate traverse()'s loop around the Walker: same hits, same step total."""
    z, w = _gold(GOLD[0])
    e = emul.world(w)
    rays = oracle.primary_rays(z["ubo_s"])
    for locate in (False, True):
        out, steps = e.trace(rays, threads=1, locate=locate)
        prof, hits = e.walk_profile(rays, locate=locate)
        assert hits == int((out["node"] != 0xFFFFFFFF).sum()) and int(prof[2].sum()) == steps
        assert np.all(prof[1] <= prof[0]) and prof[0, 0] >= hits  # every hit descended into a brick; tests are a subset of descents
        n32 = len(rays) // 32 * 32  # the lock-step (warp) restatement executes the same lane-level decisions
        p32, _ = e.walk_profile(rays[:n32], locate=locate)
        wp = e.warp_profile(rays[:n32], locate=locate)
        assert int(wp[1]) == int(p32[0].sum()) and int(wp[3]) == int(p32[1].sum()) and int(wp[5]) == int(p32[2].sum())
        assert int(wp[9]) == int(p32[3].sum()) and int(wp[10]) == n32 // 32
        assert all(int(wp[2 * k]) <= int(wp[2 * k + 1]) <= 32 * int(wp[2 * k]) for k in range(5))


def test_secondary_rays_started_from_the_previous_walks_stack(emul, oracle):
    """WalkResume (traverse.cuh): a shadow / bounce ray traced right after the camera ray begins in the deepest cell of that walk's
    node stack that holds its start voxel. Same records as a start at the root, bit for bit — for the rays the kernels cast (just off
    the voxel that was hit) and for arbitrary second rays (far away, outside the grid, after a miss)."""
    import bench
    rng = np.random.default_rng(5)
    xyz, m, mats = scenes.terrain(160)   # lit terrain: thousands of sun shadow rays, multi-chunk
    terrain = (WorldSvoGpu.from_voxels(xyz, m, mats), blok_b200.Camera((80, 110, -40), 90.0, -25.0).frame(160, 96, 0, 1).to_numpy())
    for path in (GOLD[1], GOLD[3], None):
        if path is None:
            w, ubo_s = terrain
            z = {"ubo_s": ubo_s, "subs": w.globalSubChunks}
        else:
            z, w = _gold(path)
        e = emul.world(w)
        prim = oracle.primary_rays(z["ubo_s"])
        h, t, _ = oracle.trace(to_scene(w), prim, threads=8)
        srays, cast = bench.shadow_rays(prim, h, t)
        # bounce-like rays off the same hits: random directions in the hemisphere of the face normal, origin 0.002 off the face
        normals = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], np.float32)
        hit = t >= 0
        N = normals[np.where(hit, h[:, 3], 0)]
        d = rng.normal(size=(len(prim), 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        d = np.where(((d * N).sum(1) < 0)[:, None], -d, d).astype(np.float32)
        o = (prim[:, 0:3] + prim[:, 3:6] * t[:, None] + N * np.float32(0.002)).astype(np.float32)
        brays = np.concatenate([o, d, np.full((len(o), 1), 0.001, np.float32), np.full((len(o), 1), 10000, np.float32)], 1).astype(np.float32)
        assert hit.sum() > 200 and (path is not None or cast.sum() > 2000)
        for first, second in ((prim[cast], srays), (prim[hit], brays[hit])):
            got, resumed = e.trace_chain(first, second)
            want, _ = e.trace(second, locate=True)
            assert got.tobytes() == want.tobytes()
            assert resumed == len(first)
        # arbitrary second rays (incl. after misses, from outside the grid, degenerate directions)
        lo, hi = z["subs"]["worldMin"].min(0) - 20, z["subs"]["worldMax"].max(0) + 20
        second = make_rays(rng, len(prim), lo, hi)
        got, resumed = e.trace_chain(prim, second)
        want, _ = e.trace(second, locate=True)
        assert got.tobytes() == want.tobytes()
        assert 0 < resumed < len(prim)
        # and chains of secondary rays: the second walk's own end state feeds the next start
        got, _ = e.trace_chain(second, second[::-1].copy(), locate_first=True)
        want, _ = e.trace(second[::-1].copy(), locate=True)
        assert got.tobytes() == want.tobytes()


def test_dda_path_samples_match_oracle(emul, oracle):
    """path_sample (shade.cuh) compiled for the host uses the same libm as the oracle => radiance sums agree to fp32
    rounding of the sum; also: any split of the sample range reproduces the whole (sharding invariant)."""
    z, w = _gold(GOLD[1])
    e = emul.world(w)
    ubo = z["ubo_r"]
    full = e.path(ubo, 0, 4)
    ref = oracle.render(to_scene(w), ubo, threads=8)["color"]
    got = full[..., :3] / full[..., 3:4]
    np.testing.assert_allclose(got, ref[..., :3], rtol=1e-5, atol=1e-6)
    a, b = e.path(ubo, 0, 1), e.path(ubo, 1, 4)
    np.testing.assert_allclose(a + b, full, rtol=1e-6, atol=1e-6)


def test_upload_rejects_layouts_the_reference_cannot_produce(emul):
    z, w = _gold(GOLD[0])
    subs = w.globalSubChunks.copy()
    subs["subChunkSize"][0] = 32.0
    e = emul.world(WorldSvoGpu(w.globalNodes, subs, w.materials))
    assert not e.ok() and e.rc == -3
    nodes = w.globalNodes.copy()
    root = subs_root = int(w.globalSubChunks["nodeOffset"][0] + w.globalSubChunks["rootNodeIndex"][0])
    nodes["childMask"][root] = 0
    nodes["occupancy"][root] = 1.0  # a filled 16^3 "leaf": not something SvoTree::insertVoxel emits
    e = emul.world(WorldSvoGpu(nodes, w.globalSubChunks, w.materials))
    assert not e.ok() and e.rc == -3


def test_compact_layout_counts(emul):
    z, w = _gold(GOLD[2])
    st = emul.world(w).stats()
    filled = int((w.globalNodes["occupancy"] > 0).sum())
    assert st["voxels"] == filled and st["levels"] >= 3 and st["compactBytes"] < w.globalNodes.nbytes


def test_camera_frame_matches_reference_ubo():
    z, _ = _gold(os.path.join(os.path.dirname(__file__), "golden", "castle.npz"))
    ref = FrameUBO.from_bytes(z["ubo_r"])
    cam = blok_b200.Camera((-30, 20, -25), 40.0, -15.0)
    got = cam.frame(64, 64, frame_count=3, sample_count=4)
    for name in ("view", "proj", "invView", "invProj", "prevViewProj"):
        np.testing.assert_allclose(np.array(getattr(got, name)), np.array(getattr(ref, name)), rtol=2e-5, atol=2e-5, err_msg=name)
    assert (got.frameCount, got.sampleCount, got.screenWidth, got.screenHeight) == (3, 4, 64, 64)
    assert list(got.camPos) == list(ref.camPos)
    for name in ("temporalAlpha", "momentAlpha", "phiColor", "phiNormal", "stepSize", "minHistoryLength"):
        assert getattr(got, name) == getattr(ref, name)


def test_sharding_partitions():
    for world in (1, 2, 3, 4, 8):
        for spp in (1, 8, 64, 7):
            r = [sharding.sample_range(g, world, spp) for g in range(world)]
            assert r[0][0] == 0 and r[-1][1] == spp and all(r[i][1] == r[i + 1][0] for i in range(world - 1))
        for h in (8, 720, 2160, 61):
            b = [sharding.row_band(g, world, h) for g in range(world)]
            assert b[0][0] == 0 and b[-1][1] == h and all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            assert all(x[0] % 8 == 0 for x in b)
    with pytest.raises(ValueError):
        sharding.sample_range(2, 2, 8)
