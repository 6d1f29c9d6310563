"""CPU: the oracle restatement against the golden vectors produced by the reference itself (tests/golden/make_golden.py)
and, where oracle/_ref is available, against the reference live."""
import glob
import os

import numpy as np
import pytest

from oracle.pyoracle import Scene

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


def _load(path):
    z = np.load(path)
    return z, Scene(z["nodes"], z["subs"], z["mats"])


def test_golden_present():
    assert len(GOLD) >= 4


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[:-4])
def test_restatement_hits_match_reference_goldens(oracle, path):
    z, sc = _load(path)
    rays = oracle.primary_rays(z["ubo_s"])
    hit, t, _ = oracle.trace(sc, rays, threads=4)
    rec = np.concatenate([hit, t.view(np.uint32)[:, None]], 1).astype(np.uint32)
    assert np.array_equal(rec, z["hits_s"])  # prim, node, material, face, bits(t): bit-exact
    rays = oracle.primary_rays(z["ubo_l"])
    hit, t, cnt = oracle.trace(sc, rays, threads=8)
    rec = np.concatenate([hit, t.view(np.uint32)[:, None]], 1).astype(np.uint32)
    assert int(oracle.fnv(rec), 16) == int(z["fnv_l"][0])
    assert int((t >= 0).sum()) == int(z["nhit_l"][0])
    assert int(cnt[2]) == 0  # MAX_ITER cap (intersect.rint:43,123) never reached


@pytest.mark.parametrize("path", GOLD, ids=lambda p: os.path.basename(p)[:-4])
def test_restatement_render_matches_reference_goldens(oracle, path):
    z, sc = _load(path)
    imgs = oracle.render(sc, z["ubo_r"], threads=8)
    for k in ("worldPos", "normalRough", "albedoMetal", "motion", "color"):
        # same compiler flags and libm as the generator => identical; 1e-6 leaves room for a different libm build
        np.testing.assert_allclose(imgs[k], z["img_" + k], rtol=1e-6, atol=1e-6, err_msg=k)


def test_restatement_pinned_to_live_reference(oracle, reflib):
    """Random + degenerate rays and a path-traced frame: restatement == reference sources, bit for bit."""
    from tests.conftest import make_rays
    z, sc = _load(GOLD[-1])
    rng = np.random.default_rng(7)
    rays = make_rays(rng, 40000, -60, 160)
    for any_hit in (False, True):
        h1, t1, _ = reflib.trace(sc, rays, any_hit=any_hit)
        h2, t2, _ = oracle.trace(sc, rays, any_hit=any_hit, threads=8)
        if any_hit:  # any-hit result depends on candidate order (terminate on first hit): compare occlusion only
            assert np.array_equal(t1 >= 0, t2 >= 0)
        else:
            assert np.array_equal(h1, h2) and np.array_equal(t1.view(np.uint32), t2.view(np.uint32))
    ubo = reflib.make_frame((-90, 45, 125), 5.0, -12.0, 48, 48, frame_count=11, sample_count=3)
    a, b = reflib.render(sc, ubo), oracle.render(sc, ubo, threads=8)
    for k in ("color", "worldPos", "normalRough", "albedoMetal", "motion"):
        assert np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32)), k


def test_edge_cases(oracle):
    """Empty world, ray starting inside a filled voxel (t = tmin), tmax shorter than the hit, zero-length range."""
    from blok_b200 import WorldSvoGpu, scenes
    mats = scenes.default_materials(16)
    from oracle.pyoracle import NODE_DT, SUB_DT
    empty = Scene(np.zeros(0, NODE_DT), np.zeros(0, SUB_DT), mats)
    rays = np.array([[0, 0, 0, 1, 0, 0, 0.001, 100]], np.float32)
    _, t, _ = oracle.trace(empty, rays)
    assert t[0] == -1.0
    w = WorldSvoGpu.from_voxels([[5, 5, 5], [9, 5, 5]], [3, 4], mats)
    sc = Scene(w.globalNodes, w.globalSubChunks, w.materials)
    rays = np.array([[5.5, 5.5, 5.5, 1, 0, 0, 0.001, 100],   # starts inside voxel (5,5,5): t = tmin
                     [0.5, 5.5, 5.5, 1, 0, 0, 0.001, 100],   # hits (5,5,5) at t = 4.5 on its -X face
                     [0.5, 5.5, 5.5, 1, 0, 0, 0.001, 4.0],   # tmax before the voxel: miss
                     [7.0, 5.5, 5.5, 1, 0, 0, 0.001, 100],   # between the two: hits (9,5,5)
                     [20.0, 5.5, 5.5, -1, 0, 0, 0.001, 100]], np.float32)
    hit, t, _ = oracle.trace(sc, rays)
    assert t[0] == np.float32(0.001) and hit[0, 2] == 3
    assert t[1] == 4.5 and hit[1, 2] == 3 and hit[1, 3] == 1
    assert t[2] == -1.0
    assert t[3] == 2.0 and hit[3, 2] == 4
    assert t[4] == 10.0 and hit[4, 2] == 4 and hit[4, 3] == 0
