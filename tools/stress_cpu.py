"""CPU stress: host emulation of the device walk vs the oracle restatement on millions of random + degenerate rays."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from blok_b200 import WorldSvoGpu, scenes  # noqa: E402
from oracle import pyoracle  # noqa: E402
from tests.conftest import make_rays, to_scene  # noqa: E402
from tests.host_emul import emu  # noqa: E402

O = pyoracle.OracleLib()
E = emu.load()
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 12)
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1000000
bad = 0
cases = [("terrain300", lambda: scenes.terrain(300, origin=(-100, -20, 37)), -150, 400), ("terrain128", lambda: scenes.terrain(128), -10, 140),
         ("blob128", lambda: scenes.blob_mesh(128), -20, 150)]
for name, gen, lo, hi in cases:
    xyz, m, mats = gen()
    w = WorldSvoGpu.from_voxels(xyz, m, mats)
    e = E.world(w)
    rays = make_rays(rng, n, lo, hi)
    # a third of the rays start inside the grid box, half of those exactly on lattice planes
    k = n // 3
    rays[:k, 0:3] = rng.uniform(lo * 0.3 + hi * 0.2, hi * 0.6, (k, 3)).astype(np.float32)
    rays[: k // 2, 0:3] = np.round(rays[: k // 2, 0:3])
    h, t, _ = O.trace(to_scene(w), rays, threads=8)
    hit = t >= 0
    for locate in (True, False):
      out, steps = e.trace(rays, locate=locate)
      mm = [(out["node"] != h[:, 1]).sum(), (out["t"].view(np.uint32) != t.view(np.uint32)).sum(), (out["material"] != h[:, 2]).sum(),
            ((out["facePrim"] >> 3)[hit] != h[hit, 0]).sum(), ((out["facePrim"] & 7)[hit] != h[hit, 3]).sum()]
      bad += sum(int(v) for v in mm)
      print(f"{name} locate={locate}: {n} rays, hits {hit.mean():.3f}, steps/ray {steps / n:.2f}, mismatches node/t/mat/prim/face = {mm}")
sys.exit(1 if bad else 0)
