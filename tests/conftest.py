import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

HAVE_REF_SRC = os.path.exists("/root/reference/blok/src/svo.cpp")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU restatement (oracle/restate.cpp), built on demand."""
    from oracle import pyoracle
    if not os.path.exists(pyoracle.ORACLE_SO):
        pyoracle.build(ref=False)
    return pyoracle.OracleLib()


@pytest.fixture(scope="session")
def reflib():
    """The reference's own sources compiled in place (only where /root/reference is mounted or _ref was prebuilt)."""
    from oracle import pyoracle
    if not os.path.exists(pyoracle.REF_SO):
        if not HAVE_REF_SRC:
            pytest.skip("oracle/_ref not built and /root/reference not mounted")
        pyoracle.build(ref=True)
    return pyoracle.RefLib()


@pytest.fixture(scope="session")
def emul():
    """CPU build of the product's device functions (tests/host_emul) — logic check only."""
    from tests.host_emul import emu
    return emu.load()


@pytest.fixture(scope="session")
def renderer_factory():
    import blok_b200
    made = []

    def make(w=64, h=64, device=0):
        r = blok_b200.Renderer(w, h, device)
        made.append(r)
        return r
    yield make
    for r in made:
        r.close()


def to_scene(world):
    """blok_b200.WorldSvoGpu -> oracle Scene (same arrays)."""
    from oracle.pyoracle import Scene
    return Scene(world.globalNodes, world.globalSubChunks, world.materials)


def make_rays(rng, n, lo, hi, tmax=1000.0):
    """Random rays plus the degenerate families: axis-aligned, near-zero components, lattice-plane origins."""
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    k = max(n // 100, 8)
    d[:k] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, k)] * rng.choice([-1, 1], (k, 1)).astype(np.float32)
    d[k:2 * k, rng.integers(0, 3)] *= np.float32(1e-7)
    d[2 * k:3 * k, rng.integers(0, 3)] = 0.0
    o[3 * k:5 * k] = np.round(o[3 * k:5 * k])
    o[5 * k:6 * k] = np.round(o[5 * k:6 * k]) + 0.5
    return np.concatenate([o, d, np.full((n, 1), 0.001, np.float32), np.full((n, 1), tmax, np.float32)], 1).astype(np.float32)
