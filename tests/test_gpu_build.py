"""GPU: the device scene builder (blokb_upload_voxels) against the reference's own builder.

The reference builds a chunk by replaying SvoTree::insertVoxel over the dense grid and packs chunks with
packChunksToGpuSvo; its node numbering is part of the contract because hit records carry node indices. The device
builder never materialises those arrays, so parity is checked on everything that depends on them:
  * the compact world (nodes, blockFirstLeaf = reference node index of every leaf block, voxelMat) is byte-identical to
    the one blokb_upload_world builds from the reference-format arrays;
  * the SubChunkGpu records (nodeOffset, rootNodeIndex, nodeCount, bounds) are byte-identical to the reference's;
  * hit records traced through either world are identical.
Reference-format arrays come from oracle/_ref (the reference's ChunkManager + packChunksToGpuSvo, single-chunk worlds:
its multi-chunk order is that of a std::unordered_map) and from blokb_worldgen_pack (the host restatement, chunks in
(z, y, x) order, itself checked against the reference in tests/test_host.py)."""
import numpy as np
import pytest

import blok_b200
from blok_b200 import IMG_HIT, WORLD_BLOCKFIRSTLEAF, WORLD_NODES, WORLD_SUBCHUNKS, WORLD_VOXELMAT, BlokbError, Camera, WorldSvoGpu, scenes
from tests.conftest import make_rays

pytestmark = pytest.mark.gpu


def _cloud(rng, n, lo, hi):
    xyz = rng.integers(lo, hi, (n, 3)).astype(np.int32)
    return xyz, rng.integers(1, 9, n).astype(np.uint32)


def _cases():
    rng = np.random.default_rng(7)
    mats = scenes.default_materials(16)
    out = {}
    out["terrain96"] = scenes.terrain(96)
    out["blob64"] = scenes.blob_mesh(64)
    out["terrain320_multi_chunk"] = scenes.terrain(320)
    xyz, m = _cloud(rng, 5000, -300, 300)
    out["sparse_cloud_negative"] = (xyz, m, mats)
    xyz, m = _cloud(rng, 40000, 0, 40)
    out["dense_cloud_duplicates"] = (xyz, m, mats)  # 64 000 cells, 40 000 draws: many repeated voxels, last write wins
    out["single_voxel"] = (np.array([[5, 6, 7]], np.int32), np.array([3], np.uint32), mats)
    out["single_voxel_negative"] = (np.array([[-1, -1, -1]], np.int32), np.array([2], np.uint32), mats)
    g = np.stack(np.meshgrid(np.arange(120, 136), np.arange(120, 136), np.arange(-8, 8), indexing="ij"), -1).reshape(-1, 3).astype(np.int32)
    out["cube_across_8_chunks"] = (g, (1 + (g.sum(1) % 7)).astype(np.uint32), mats)
    xyz, m = _cloud(rng, 3000, 0, 128)
    out["wide_material_ids"] = (xyz, (m + 70000).astype(np.uint32), mats)
    full = np.stack(np.meshgrid(np.arange(16), np.arange(16), np.arange(16), indexing="ij"), -1).reshape(-1, 3).astype(np.int32)
    out["full_subchunk"] = (full + 32, np.ones(len(full), np.uint32), mats)
    return out


CASES = _cases()


def _export(r):
    return {k: r.world_export(k) for k in (WORLD_NODES, WORLD_BLOCKFIRSTLEAF, WORLD_VOXELMAT)}


@pytest.mark.parametrize("name", list(CASES))
def test_device_build_equals_reference_numbering(renderer_factory, name):
    xyz, mid, mats = CASES[name]
    world = WorldSvoGpu.from_voxels(xyz, mid, mats)  # reference-format arrays (host restatement of the reference builder)
    a = renderer_factory(64, 64)
    a.add_world(world)
    b = renderer_factory(64, 64)
    b.add_voxels(xyz, mid, mats)
    ea, eb = _export(a), _export(b)
    for k in ea:
        assert ea[k].shape == eb[k].shape, (name, k, ea[k].shape, eb[k].shape)
        assert np.array_equal(ea[k].view(np.uint8), eb[k].view(np.uint8)), (name, k)
    subs = b.world_export(WORLD_SUBCHUNKS)
    assert np.array_equal(subs.view(np.uint8), world.globalSubChunks.view(np.uint8)), name
    sa, sb = a.world_stats(), b.world_stats()
    for key in ("nodes", "subChunks", "bricks", "blocks", "voxels", "compactBytes", "gridOrigin", "gridDim", "levels"):
        assert sa[key] == sb[key], (name, key, sa[key], sb[key])
    assert sb["nodes"] == len(world.globalNodes)
    # and the traversal sees the same world
    lo = xyz.min(0).astype(np.float32) - 20
    hi = xyz.max(0).astype(np.float32) + 20
    rays = make_rays(np.random.default_rng(1), 4000, lo, hi)
    ha, hb = a.trace_rays(rays), b.trace_rays(rays)
    assert np.array_equal(ha.view(np.uint8), hb.view(np.uint8)), name


@pytest.mark.parametrize("name", ["terrain96", "blob64", "dense_cloud_duplicates", "single_voxel", "full_subchunk"])
def test_device_build_equals_the_reference_builder(renderer_factory, reflib, name):
    """Single-chunk worlds through the reference's own ChunkManager + packChunksToGpuSvo (oracle/_ref)."""
    xyz, mid, mats = CASES[name]
    assert xyz.min() >= 0 and xyz.max() < 128
    # the reference keeps the last write too (setVoxelMaterial overwrites the dense cell)
    sc = reflib.scene_from_voxels(xyz, mid, mats)
    a = renderer_factory(64, 64)
    a.add_world(WorldSvoGpu(sc.nodes, sc.subs, sc.mats))
    b = renderer_factory(64, 64)
    b.add_voxels(xyz, mid, mats)
    ea, eb = _export(a), _export(b)
    for k in ea:
        assert np.array_equal(ea[k].view(np.uint8), eb[k].view(np.uint8)), (name, k)
    assert np.array_equal(b.world_export(WORLD_SUBCHUNKS).view(np.uint8), sc.subs.view(np.uint8))
    assert b.world_stats()["nodes"] == len(sc.nodes)


def test_empty_and_bad_input(renderer_factory):
    r = renderer_factory(32, 32)
    mats = scenes.default_materials(4)
    r.add_voxels(np.zeros((0, 3), np.int32), np.zeros(0, np.uint32), mats)
    st = r.world_stats()
    assert st["voxels"] == 0 and st["subChunks"] == 0
    r.trace_primary(Camera((0, 0, -10), 90.0, 0.0).frame(32, 32, 0, 1))
    assert np.all(r.readback(IMG_HIT)["node"] == 0xFFFFFFFF)
    with pytest.raises(BlokbError) as e:
        r.add_voxels(np.array([[1 << 21, 0, 0]], np.int32), np.array([1], np.uint32), mats)
    assert e.value.code == -3
    # a failed build leaves the previous world in place
    assert r.world_stats()["voxels"] == 0
    r.add_voxels(np.array([[3, 3, 3]], np.int32), np.array([1], np.uint32), mats)  # update_world: replaces it
    assert r.world_stats()["voxels"] == 1


def test_rendered_frame_identical(renderer_factory):
    xyz, mid, mats = CASES["terrain96"]
    a = renderer_factory(96, 64)
    a.add_world(WorldSvoGpu.from_voxels(xyz, mid, mats))
    b = renderer_factory(96, 64)
    b.add_voxels(xyz, mid, mats)
    f = Camera((48, 60, -20), 90.0, -25.0).frame(96, 64, 2, 2)
    for r in (a, b):
        r.render(f)
    for img in (blok_b200.IMG_COLOR, blok_b200.IMG_WORLDPOS, blok_b200.IMG_NORMALROUGH, blok_b200.IMG_ALBEDOMETAL):
        assert np.array_equal(a.readback(img).view(np.uint8), b.readback(img).view(np.uint8))


def test_vox_file_straight_to_the_device(renderer_factory):
    """blokb_upload_vox on the committed synthetic file == blokb_upload_world of what the reference's loader + builder made of it
    (tests/golden/vox/synth_ref.npz, from oracle/_ref), for both models, incl. a world offset."""
    import os
    here = os.path.dirname(os.path.abspath(__file__))
    z = np.load(os.path.join(here, "golden", "vox", "synth_ref.npz"))
    for tag, model in (("m0", 0), ("m1", 1)):
        a = renderer_factory(64, 64)
        a.add_world(WorldSvoGpu(z[tag + "_nodes"], z[tag + "_subs"], z[tag + "_mats"]))
        b = renderer_factory(64, 64)
        b.add_vox(os.path.join(here, "golden", "vox", "synth.vox"), model, tuple(z[tag + "_offset"]))
        ea, eb = _export(a), _export(b)
        for k in ea:
            assert np.array_equal(ea[k].view(np.uint8), eb[k].view(np.uint8)), (tag, k)
        assert np.array_equal(b.world_export(WORLD_SUBCHUNKS).view(np.uint8), z[tag + "_subs"].view(np.uint8))
        f = Camera((-30, 40, -30), 45.0, -25.0).frame(64, 64, 1, 2)
        for r in (a, b):
            r.render(f)
        for img in (blok_b200.IMG_COLOR, blok_b200.IMG_ALBEDOMETAL):  # materials (MATL: metal / glass / emissive) reach the shading
            assert np.array_equal(a.readback(img).view(np.uint8), b.readback(img).view(np.uint8)), (tag, img)
    with pytest.raises(BlokbError):
        b.add_vox(os.path.join(here, "golden", "vox", "missing.vox"))


def _dense_from_voxels(xyz, mid, density=None):
    """Voxel writes -> {chunk: (density, materialIds)} exactly as ChunkManager::setVoxel fills its dense arrays."""
    chunks = {}
    den = np.ones(len(mid), np.float32) if density is None else density
    c = np.floor_divide(xyz, 128)
    l = xyz - c * 128
    for i in range(len(mid)):
        key = (int(c[i, 0]), int(c[i, 1]), int(c[i, 2]))
        if key not in chunks:
            chunks[key] = (np.zeros(128 ** 3, np.float32), np.zeros(128 ** 3, np.uint32))
        t = int(l[i, 0]) + 128 * int(l[i, 1]) + 16384 * int(l[i, 2])
        chunks[key][0][t] = den[i]
        chunks[key][1][t] = mid[i]
    return chunks


@pytest.mark.parametrize("name", ["terrain96", "sparse_cloud_negative", "cube_across_8_chunks"])
def test_dense_chunks_equal_voxel_list(renderer_factory, name):
    """blokb_upload_dense (the reference's Chunk arrays) == blokb_upload_voxels == the host path."""
    xyz, mid, mats = CASES[name]
    if name == "sparse_cloud_negative":
        xyz, mid = xyz[:600], mid[:600]  # keeps the number of 16 MiB dense chunks small
    a = renderer_factory(64, 64)
    a.add_world(WorldSvoGpu.from_voxels(xyz, mid, mats))
    b = renderer_factory(64, 64)
    b.add_dense(_dense_from_voxels(xyz, mid), mats)
    ea, eb = _export(a), _export(b)
    for k in ea:
        assert np.array_equal(ea[k].view(np.uint8), eb[k].view(np.uint8)), (name, k)
    assert a.world_stats()["nodes"] == b.world_stats()["nodes"]


def test_dense_density_and_empty_chunk_follow_the_reference(renderer_factory, reflib):
    """Fractional densities fill (> 0), zero / negative ones do not, and a chunk whose only voxel was erased is still packed by
    the reference as a lone root node — every later node index moves up by one. Checked against the reference's own
    ChunkManager; its chunk order is a hash map's, so the erased voxel is placed where both orders put that chunk first."""
    xyz, mid, mats = CASES["blob64"]
    rng = np.random.default_rng(3)
    den = rng.choice(np.array([1.0, 0.25, 0.0, -1.0], np.float32), len(mid), p=[0.6, 0.2, 0.1, 0.1])
    found = False
    for ghost in ([-5, 3, 3], [3, -5, 3], [3, 3, -5], [-5, -5, -5], [-200, 3, 3]):
        # written last: libstdc++'s unordered_map iterates the most recently created chunk first
        wx = np.concatenate([xyz, np.array([ghost], np.int32)])
        wm = np.concatenate([mid, np.array([9], np.uint32)])
        wd = np.concatenate([den, np.array([0.0], np.float32)])
        sc = reflib.scene_from_writes(wx, wm, wd, mats)
        if len(sc.subs) and sc.subs["nodeOffset"][0] == 1:  # the reference packed the empty chunk first, like our (z, y, x) order
            found = True
            break
    assert found, "no ghost position where the reference's unordered_map puts the empty chunk first"
    a = renderer_factory(64, 64)
    a.add_world(WorldSvoGpu(sc.nodes, sc.subs, sc.mats))
    b = renderer_factory(64, 64)
    b.add_dense(_dense_from_voxels(wx, wm, wd), mats)
    ea, eb = _export(a), _export(b)
    for k in ea:
        assert np.array_equal(ea[k].view(np.uint8), eb[k].view(np.uint8)), k
    assert np.array_equal(b.world_export(WORLD_SUBCHUNKS).view(np.uint8), sc.subs.view(np.uint8))
    assert b.world_stats()["nodes"] == len(sc.nodes)
    with pytest.raises(BlokbError):
        z = (np.zeros(128 ** 3, np.float32), np.zeros(128 ** 3, np.uint32))
        b.L.blokb_upload_dense  # same chunk twice is rejected
        import ctypes as C
        from blok_b200.api import DenseChunkC, _ck
        arr = (DenseChunkC * 2)(DenseChunkC(0, 0, 0, z[0].ctypes.data, z[1].ctypes.data), DenseChunkC(0, 0, 0, z[0].ctypes.data, z[1].ctypes.data))
        _ck(b.L.blokb_upload_dense(b._ctx, C.cast(arr, C.c_void_p), 2, None, 0))


# ---- scene dump: blokb_world_save / blokb_world_load -----------------------------------------
def _trace_all(r, cam, w=96, h=64):
    r.trace_primary(cam.frame(w, h, 0, 1), shadow=True)
    return r.readback(IMG_HIT).tobytes(), r.readback(blok_b200.IMG_SHADOW).tobytes()


@pytest.mark.parametrize("name", ["terrain96", "sparse_cloud_negative", "wide_material_ids", "single_voxel_negative"])
@pytest.mark.parametrize("built_on", ["device", "host"])
def test_scene_dump_round_trip(tmp_path, name, built_on):
    xyz, mid, mats = CASES[name]
    a = blok_b200.Renderer(96, 64, 0)
    if built_on == "device":
        a.add_voxels(xyz, mid, mats)
    else:
        a.add_world(WorldSvoGpu.from_voxels(xyz, mid, mats))
    lo, hi = xyz.min(0).astype(np.float64), xyz.max(0).astype(np.float64)
    cam = Camera(tuple((lo + hi) / 2 + np.array([0.0, (hi - lo).max() * 0.8 + 20, -(hi - lo).max() * 0.9 - 20])), 90.0, -35.0)
    want = _trace_all(a, cam)
    path = tmp_path / "world.blokb"
    a.save_world(path)
    b = blok_b200.Renderer(96, 64, 0)
    b.load_world(path)
    ea, eb = _export(a), _export(b)
    for k in ea:
        assert ea[k].tobytes() == eb[k].tobytes()
    sa, sb = a.world_stats(), b.world_stats()
    for k in ("nodes", "subChunks", "materials", "bricks", "blocks", "voxels", "compactBytes", "levels"):
        assert sa[k] == sb[k], k
    if built_on == "device":
        assert a.world_export(WORLD_SUBCHUNKS).tobytes() == b.world_export(WORLD_SUBCHUNKS).tobytes()
    assert _trace_all(b, cam) == want
    # a dump of the loaded world is the same file
    path2 = tmp_path / "again.blokb"
    b.save_world(path2)
    r1, r2 = path.read_bytes(), path2.read_bytes()   # 200-byte header; bytes 172..184 are the stats' two timing floats
    assert r1[:8] == b"BLOKBWD1" and r1[:172] == r2[:172] and r1[184:] == r2[184:]
    a.close(); b.close()


def test_scene_dump_rejects_damaged_files(tmp_path):
    xyz, mid, mats = CASES["terrain96"]
    a = blok_b200.Renderer(64, 64, 0)
    a.add_voxels(xyz, mid, mats)
    path = tmp_path / "w.blokb"
    a.save_world(path)
    raw = path.read_bytes()
    b = blok_b200.Renderer(64, 64, 0)
    with pytest.raises(BlokbError, match="no world"):
        b.save_world(tmp_path / "none.blokb")
    cases = {"truncated payload": raw[:-100], "truncated header": raw[:40], "not a blok-b200": b"XXXXXXXX" + raw[8:],
             "checksum": raw[:-9] + bytes([raw[-9] ^ 1]) + raw[-8:], "trailing": raw + b"\0", "version": raw[:8] + b"\x07" + raw[9:]}
    for why, data in cases.items():
        p = tmp_path / "bad.blokb"
        p.write_bytes(data)
        with pytest.raises(BlokbError, match=why.split()[0]):
            b.load_world(p)
    with pytest.raises(BlokbError, match="cannot open"):
        b.load_world(tmp_path / "missing.blokb")
    # files that pass the checksum but would send the traversal out of bounds: every index it follows is range-checked on load
    import struct

    def fnv(payload):
        h = 1469598103934665603
        for byte in payload:
            h = ((h ^ byte) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
        return h

    def crafted(edit_header=None, edit_payload=None):
        head, payload = bytearray(raw[:200]), bytearray(raw[200:])
        if edit_payload:
            edit_payload(payload)
            head[192:200] = struct.pack("<Q", fnv(payload))
        if edit_header:
            edit_header(head)
        return bytes(head) + bytes(payload)

    n_nodes = struct.unpack_from("<Q", raw, 48)[0]
    assert fnv(raw[200:]) == struct.unpack_from("<Q", raw, 192)[0]
    forged = {
        "points outside": crafted(edit_payload=lambda pl: struct.pack_into("<I", pl, 8, 0x7FFFFF00)),                     # root.base
        "past voxelMat": crafted(edit_payload=lambda pl: struct.pack_into("<I", pl, 16 * (n_nodes - 1) + 8, 0xFFFFFF00)),    # last brick.base
        "past blockFirstLeaf": crafted(edit_payload=lambda pl: struct.pack_into("<I", pl, 16 * (n_nodes - 1) + 12, 0xFFFFFF00)),
        "children|points outside|past|longer|run past": crafted(edit_payload=lambda pl: struct.pack_into("<Q", pl, 0, 0xFFFFFFFFFFFFFFFF)),  # root.mask
        "multiple of 16": crafted(edit_header=lambda h: struct.pack_into("<i", h, 16, 8)),                                  # origin.x
        "exceeds 4": crafted(edit_header=lambda h: struct.pack_into("<i", h, 28, 1 << 20)),                                 # dim.x
    }
    for why, data in forged.items():
        p = tmp_path / "forged.blokb"
        p.write_bytes(data)
        with pytest.raises(BlokbError, match=why):
            b.load_world(p)
    # a failed load leaves the context usable
    b.load_world(path)
    assert b.world_stats()["voxels"] == a.world_stats()["voxels"]
    a.close(); b.close()


# ---- editing: dirty chunks only (blokb_update_dense) ------------------------------------------------------------------
def _brush(chunks, centre, radius, mode, value):
    """applyBrush (blok/src/brush.cpp:13-63) on the dict-of-dense-arrays stand-in for ChunkManager (voxelSize 1): every chunk the
    bounding box touches is created, cells within `radius` of the centre get max(d, value) / min(d, value). Returns dirty keys."""
    c = np.asarray(centre, np.float32)
    lo = np.floor(c - np.float32(radius)).astype(int)
    hi = np.floor(c + np.float32(radius)).astype(int) + 1
    dirty = set()
    for gz in range(lo[2], hi[2]):
        for gy in range(lo[1], hi[1]):
            gx = np.arange(lo[0], hi[0])
            cz, cy = gz // 128, gy // 128
            for cx in np.unique(gx // 128):
                key = (int(cx), int(cy), int(cz))
                if key not in chunks:
                    chunks[key] = (np.zeros(128 ** 3, np.float32), np.zeros(128 ** 3, np.uint32))
                xs = gx[gx // 128 == cx]
                ws = np.stack([xs + 0.5, np.full(len(xs), gy + 0.5), np.full(len(xs), gz + 0.5)], 1).astype(np.float32)
                inside = np.sqrt(((ws - c) ** 2).sum(1, dtype=np.float32)) <= np.float32(radius)
                if not inside.any():
                    continue
                t = (xs[inside] - cx * 128) + 128 * (gy - cy * 128) + 16384 * (gz - cz * 128)
                d = chunks[key][0]
                d[t] = np.maximum(d[t], value) if mode == "add" else np.minimum(d[t], value)
                dirty.add(key)
    return dirty


def test_update_dense_equals_full_upload(renderer_factory):
    """A session of brush edits: after each one only the dirty chunks go to blokb_update_dense; the world must be byte-identical
    (compact arrays, SubChunkGpu records, stats, traced frame) to blokb_upload_dense of ALL chunks into a fresh context."""
    xyz, mid, mats = scenes.terrain(200)
    chunks = _dense_from_voxels(xyz, mid)
    r = renderer_factory(96, 64)
    r.add_dense(chunks, mats)
    cam = Camera((100.0, 160.0, -60.0), 90.0, -35.0)
    edits = [((100.0, 70.0, 100.0), 9.0, "sub", 0.0),      # dig a crater (inside chunk (0,0,0))
             ((126.0, 80.0, 126.0), 6.0, "add", 1.0),      # a ball of material-0 voxels across the 8-chunk corner
             ((300.0, 40.0, 40.0), 3.0, "add", 0.5),       # a brand-new chunk far away (fractional density fills)
             ((300.0, 40.0, 40.0), 4.0, "sub", 0.0),       # ... erased again: that chunk stays, as a lone root node
             ((-20.0, 60.0, 30.0), 5.0, "add", 1.0)]       # negative coordinates
    for k, (centre, radius, mode, value) in enumerate(edits):
        dirty = _brush(chunks, centre, radius, mode, value)
        assert dirty
        r.update_dense({key: chunks[key] for key in dirty})
        full = blok_b200.Renderer(96, 64, 0)
        full.add_dense(chunks, mats)
        ea, eb = _export(r), _export(full)
        for which in ea:
            assert ea[which].tobytes() == eb[which].tobytes(), (k, which)
        assert r.world_export(WORLD_SUBCHUNKS).tobytes() == full.world_export(WORLD_SUBCHUNKS).tobytes(), k
        sa, sb = r.world_stats(), full.world_stats()
        for key in ("nodes", "subChunks", "bricks", "blocks", "voxels", "compactBytes", "levels"):
            assert sa[key] == sb[key], (k, key)
        assert _trace_all(r, cam) == _trace_all(full, cam), k
        full.close()
    # the erased far chunk still counts one node (checked against the reference's ChunkManager in
    # test_dense_density_and_empty_chunk_follow_the_reference): total = sum over chunks of 1 + 8 * interior
    assert any(not (d > 0).any() for d, _ in chunks.values())


def test_update_dense_needs_an_editable_world(renderer_factory):
    xyz, mid, mats = CASES["terrain96"]
    r = renderer_factory(64, 64)
    z = {(0, 0, 0): (np.zeros(128 ** 3, np.float32), np.zeros(128 ** 3, np.uint32))}
    with pytest.raises(BlokbError):
        r.update_dense(z)                    # no world at all
    r.add_voxels(xyz, mid, mats)
    with pytest.raises(BlokbError, match="blokb_upload_dense"):
        r.update_dense(z)                    # world not uploaded as dense chunks
    r.add_dense(_dense_from_voxels(xyz, mid), mats)
    r.update_dense({})                       # nothing dirty: no-op
    want = _export(r)
    r.update_dense(_dense_from_voxels(xyz, mid))   # the same content again: unchanged world
    got = _export(r)
    for k in want:
        assert want[k].tobytes() == got[k].tobytes()
    r.add_voxels(xyz, mid, mats)             # replaced by a non-editable world again
    with pytest.raises(BlokbError, match="blokb_upload_dense"):
        r.update_dense(z)


def test_vox_model_without_voxels_is_a_failed_load(renderer_factory, tmp_path):
    """loadAndImportVox returns `count > 0` (vox_loader.cpp:457): a model that imports nothing is a failed load in the reference's
    app; blokb_upload_vox reports it the same way and leaves the previous world in place. (Found by tools/fuzz_vox.py.)"""
    from tests.voxfile import write_vox
    xyz, mid, mats = CASES["terrain96"]
    r = renderer_factory(64, 64)
    r.add_voxels(xyz, mid, mats)
    before = _export(r)
    p = tmp_path / "empty.vox"
    p.write_bytes(write_vox([((8, 8, 8), np.zeros((0, 4), np.uint8))]))
    with pytest.raises(BlokbError, match="imports no voxels"):
        r.add_vox(str(p))
    after = _export(r)
    for k in before:
        assert before[k].tobytes() == after[k].tobytes()
