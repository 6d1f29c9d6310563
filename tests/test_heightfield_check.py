"""Pins the heightfield property checker (tests/heightfield.py — the full-size parity check of configs[4]) against the
oracle at test scale: it accepts the oracle's hit records and rejects corrupted ones."""
import numpy as np

from blok_b200 import WorldSvoGpu, scenes, workloads
from tests import heightfield
from tests.conftest import to_scene


def _case(oracle, N=96, w=160, h=90):
    xyz, mid, mats = scenes.terrain(N)
    world = WorldSvoGpu.from_voxels(xyz, mid, mats)
    cam = workloads.WORKLOADS["t1"]["cam"](N)
    ubo = cam.frame(w, h, 0, 1).to_numpy()
    rays = oracle.primary_rays(ubo)
    hits, t, _ = oracle.trace(to_scene(world), rays, threads=4)
    return scenes.terrain_heights(N), rays, hits, t


def test_checker_accepts_the_oracles_hits(oracle):
    hgt, rays, hits, t = _case(oracle)
    res = heightfield.check_hits(hgt, 3, rays, hits[:, 1], hits[:, 2], t, hits[:, 3])
    assert res["hits"] > 1000 and res["hits"] < res["rays"]      # the frame holds both ground and sky
    assert heightfield.passed(res), res


def test_checker_accepts_random_rays_inside_the_grid(oracle):
    N = 96
    xyz, mid, mats = scenes.terrain(N)
    world = WorldSvoGpu.from_voxels(xyz, mid, mats)
    rng = np.random.default_rng(5)
    o = rng.uniform(0, N, (4000, 3)).astype(np.float32)
    o[:, 1] = rng.uniform(N / 4 - 4, N / 2, 4000)
    d = rng.normal(size=(4000, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.concatenate([o, d, np.full((4000, 1), 0.001, np.float32), np.full((4000, 1), 1000, np.float32)], 1).astype(np.float32)
    hgt = scenes.terrain_heights(N)
    # rays that start inside a filled voxel report that voxel at tmin: outside the checker's "face of a voxel" model
    vo = np.floor(o).astype(np.int64)
    keep = ~heightfield.filled(hgt, 3, vo[:, 0], vo[:, 1], vo[:, 2])
    rays = rays[keep]
    hits, t, _ = oracle.trace(to_scene(world), rays, threads=4)
    res = heightfield.check_hits(hgt, 3, rays, hits[:, 1], hits[:, 2], t, hits[:, 3])
    assert heightfield.passed(res), res


def test_checker_rejects_corrupted_records(oracle):
    hgt, rays, hits, t = _case(oracle)
    node, mat, face = hits[:, 1].copy(), hits[:, 2].copy(), hits[:, 3].copy()
    hit = node != 0xFFFFFFFF
    k = np.nonzero(hit)[0]
    # (a) a hit reported 3 voxels too far: the true voxel is now "filled before the hit" (or the far point is in the air)
    t2 = t.copy(); t2[k[:50]] += 3.0
    r = heightfield.check_hits(hgt, 3, rays, node, mat, t2, face)
    assert r["filled_before_hit"] + r["bad_hit_voxel"] >= 50
    # (b) hits turned into misses
    n2 = node.copy(); n2[k[:40]] = 0xFFFFFFFF
    r = heightfield.check_hits(hgt, 3, rays, n2, mat, t, face)
    assert r["filled_on_missed_ray"] == 40
    # (c) a hit reported too early (in the air)
    t3 = t.copy(); t3[k[:30]] *= 0.9
    r = heightfield.check_hits(hgt, 3, rays, node, mat, t3, face)
    assert r["bad_hit_voxel"] >= 30
    # (d) wrong material
    m2 = mat.copy(); m2[k[:20]] += 1
    r = heightfield.check_hits(hgt, 3, rays, node, m2, t, face)
    assert r["bad_material"] == 20
