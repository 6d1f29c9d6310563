""".vox import, CPU side: blokb_vox_* against the reference's loadVoxFile / importVoxMaterials /
importVoxToChunks + its builder (oracle/_ref) — live on the reference's seven assets where /root/reference is mounted, and on
the committed fixture tests/golden/vox (our own synthetic file + what the reference made of it) everywhere."""
import glob
import os

import numpy as np
import pytest

import blok_b200
from blok_b200 import BlokbError, WorldSvoGpu, load_vox
from tests.voxfile import write_vox

HERE = os.path.dirname(os.path.abspath(__file__))
SYNTH = os.path.join(HERE, "golden", "vox", "synth.vox")
ASSETS = sorted(glob.glob("/root/reference/assets/models/*.vox"))


def _same(a, b):
    return a.shape == b.shape and np.array_equal(np.ascontiguousarray(a).view(np.uint8), np.ascontiguousarray(b).view(np.uint8))


@pytest.mark.parametrize("tag,model", [("m0", 0), ("m1", 1)])
def test_synthetic_file_matches_reference_fixture(tag, model):
    z = np.load(os.path.join(HERE, "golden", "vox", "synth_ref.npz"))
    xyz, ids, mats, info = load_vox(SYNTH, model, tuple(z[tag + "_offset"]))
    assert info["models"] == 2
    assert _same(mats, z[tag + "_mats"])  # 256 x MaterialGpu incl. the MATL cases (metal, glass, emissive defaults, unparsable float)
    w = WorldSvoGpu.from_voxels(xyz, ids, mats)
    assert _same(w.globalNodes, z[tag + "_nodes"]) and _same(w.globalSubChunks, z[tag + "_subs"])
    with open(SYNTH, "rb") as f:  # the in-memory entry point parses the same bytes
        x2, i2, m2, _ = load_vox(f.read(), model, tuple(z[tag + "_offset"]))
    assert _same(xyz, x2) and _same(ids, i2) and _same(mats, m2)


@pytest.mark.skipif(not ASSETS, reason="/root/reference not mounted")
@pytest.mark.parametrize("path", ASSETS, ids=lambda p: os.path.basename(p)[:-4])
def test_reference_assets_match_reference_loader(reflib, path, capfd):
    for model in ((0, 1) if path.endswith("cars.vox") else (0,)):
        sc = reflib.scene_from_vox(path, (0, 0, 0), model)
        xyz, ids, mats, _ = load_vox(path, model)
        assert _same(mats, sc.mats)
        w = WorldSvoGpu.from_voxels(xyz, ids, mats)
        assert _same(w.globalNodes, sc.nodes) and _same(w.globalSubChunks, sc.subs), (path, model)


def test_default_palette_and_fractional_offset(reflib, tmp_path, capfd):
    """No RGBA chunk: MagicaVoxel's default palette (generated here, a table in the reference); offset with a fraction: floor()."""
    rng = np.random.default_rng(5)
    v = np.concatenate([rng.integers(0, 30, (500, 3)), np.arange(500)[:, None] % 255 + 1], 1).astype(np.uint8)
    p = tmp_path / "nopal.vox"
    p.write_bytes(write_vox([((30, 30, 30), v)], palette=None, matl=None, extra_chunks=False))
    off = (3.5, 0.25, 40.75)
    sc = reflib.scene_from_vox(str(p), off, 0)
    xyz, ids, mats, _ = load_vox(str(p), 0, off)
    assert _same(mats, sc.mats)
    w = WorldSvoGpu.from_voxels(xyz, ids, mats)
    assert _same(w.globalNodes, sc.nodes) and _same(w.globalSubChunks, sc.subs)


def test_errors(tmp_path):
    with pytest.raises(BlokbError):
        load_vox(str(tmp_path / "missing.vox"))
    with pytest.raises(BlokbError) as e:
        load_vox(b"NOPE" + b"\0" * 32)
    assert "magic" in str(e.value)
    v = np.array([[1, 2, 3, 4]], np.uint8)
    with pytest.raises(BlokbError) as e:
        load_vox(write_vox([((4, 4, 4), v)], version=100))
    assert e.value.code == -3
    with pytest.raises(BlokbError) as e:
        load_vox(write_vox([((4, 4, 4), v)]), model_index=3)
    assert "model 3 requested" in str(e.value)
    with pytest.raises(BlokbError):
        load_vox(b"VOX " + (150).to_bytes(4, "little") + b"MAIN" + b"\0" * 8)  # no models
    xyz, ids, mats, info = load_vox(write_vox([((4, 4, 4), v)], extra_chunks=False)[:-2])  # truncated voxel record: dropped, like a short read
    assert len(ids) == 0 and info["models"] == 1
    xyz, ids, _, _ = load_vox(write_vox([((4, 4, 4), v)]), offset=(-10, 0, 0))
    assert xyz.tolist() == [[-9, 3, 2]] and ids.tolist() == [4]  # VOX (x, y, z) -> world (x, z, y)
