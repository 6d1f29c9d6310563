"""Where the walk spends its decisions, per level, counted on the CPU build of the product's own Walker (tests/host_emul):
descents, descents that needed plane tests, DDA steps and climbs by child-cell size, for a strided sample of a workload's
primary rays and for the sun shadow rays of their hits. Input for DESIGN.md 12 (what a different leaf size would remove).
  python tools/walk_profile.py [c3] [stride]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from blok_b200 import workloads  # noqa: E402
from oracle import pyoracle  # noqa: E402
from tests.host_emul import emu  # noqa: E402
import bench  # noqa: E402


def table(title, prof, hits, n):
    print(f"\n{title}: {n} rays, {hits} hits")
    print(f"{'child cell':>12} {'descents':>10} {'w/ tests':>10} {'steps':>10} {'climbs to':>10}   per ray: desc  steps")
    used = [k for k in range(prof.shape[1]) if prof[:, k].any()]
    for k in used:
        d, s, st, c = (int(v) for v in prof[:, k])
        print(f"{4 ** k:>10}^3 {d:>10} {s:>10} {st:>10} {c:>10}            {d / n:6.2f} {st / n:6.2f}")
    tot = prof.sum(1)
    print(f"{'total':>12} {int(tot[0]):>10} {int(tot[1]):>10} {int(tot[2]):>10} {int(tot[3]):>10}            {tot[0] / n:6.2f} {tot[2] / n:6.2f}")


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c3"
    stride = int(sys.argv[2]) if len(sys.argv) > 2 else 37
    wl = workloads.WORKLOADS[name]
    world = workloads.build_world(name)
    emu.build()
    e = emu.load().world(world)
    assert e.ok(), e.err
    if not os.path.exists(pyoracle.ORACLE_SO):
        pyoracle.build(ref=False)
    O = pyoracle.OracleLib()
    ubo = workloads.camera_for(name).frame(wl["width"], wl["height"], 0, 1).to_numpy()
    rays = O.primary_rays(ubo)[::stride].copy()
    prof, hits = e.walk_profile(rays, locate=False)
    table(f"{name} primary rays (every {stride}th pixel)", prof, hits, len(rays))
    out, _ = e.trace(rays, locate=False)
    h = np.stack([out["facePrim"] >> 3, out["node"], out["material"], out["facePrim"] & 7], 1)
    sh = bench.shadow_rays(rays, h, out["t"])
    if len(sh):
        prof, hits = e.walk_profile(sh, locate=True)
        table(f"{name} sun shadow rays of those hits", prof, hits, len(sh))


if __name__ == "__main__":
    main()
