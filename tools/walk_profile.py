"""Where the walk spends its decisions, per level, counted on the CPU build of the product's own Walker (tests/host_emul):
descents, descents that needed plane tests, DDA steps and climbs by child-cell size, for a strided sample of a workload's
primary rays and for the sun shadow rays of their hits. Input for DESIGN.md 12 (what a different leaf size would remove).
  python tools/walk_profile.py [c3] [stride]            per-ray counts by level
  python tools/walk_profile.py --warps [c3] [stride]    32 Walkers in lock step per 8x4 tile: lanes per issued decision"""
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
    sh, _ = bench.shadow_rays(rays, h, out["t"])
    if len(sh):
        prof, hits = e.walk_profile(sh, locate=True)
        table(f"{name} sun shadow rays of those hits", prof, hits, len(sh))


def warp_table(title, c):
    names = ("descend", "  select inside descend", "step", "  widen inside step", "  climb inside step")
    print(f"\n{title}: {int(c[10])} warps in lock step")
    print(f"{'decision':>26} {'warp issues':>12} {'lane execs':>12} {'lanes / issue':>14} {'issues / warp':>14}")
    for k, nm in enumerate(names):
        iss, ln = int(c[2 * k]), int(c[2 * k + 1])
        print(f"{nm:>26} {iss:>12} {ln:>12} {ln / max(iss, 1):>14.2f} {iss / max(int(c[10]), 1):>14.2f}")


def warps(name="c3", stride=37):
    """The same walk as a warp executes it: 8x4-pixel tiles (lane = x + 8 y, as thread_pixel maps them), every stride-th tile."""
    wl = workloads.WORKLOADS[name]
    W, H = wl["width"], wl["height"]
    world = workloads.build_world(name)
    emu.build()
    e = emu.load().world(world)
    if not os.path.exists(pyoracle.ORACLE_SO):
        pyoracle.build(ref=False)
    ubo = workloads.camera_for(name).frame(W, H, 0, 1).to_numpy()
    rays = pyoracle.OracleLib().primary_rays(ubo).reshape(H, W, 8)
    assert H % 4 == 0 and W % 8 == 0
    tiles = rays.reshape(H // 4, 4, W // 8, 8, 8).transpose(0, 2, 1, 3, 4).reshape(-1, 32, 8)[::stride]
    warp_table(f"{name} primary rays, every {stride}th 8x4 tile", e.warp_profile(tiles.reshape(-1, 8), locate=False))
    # what-if loop shapes: at most K descend passes before each step pass (K = 1 is the `if occupied descend else step` loop §5 dropped).
    # Cost in issue slots with rough per-decision instruction counts read off the SASS (descend 24, select 42, step 40, widen 14, climb 10).
    wts = np.array([24, 42, 40, 14, 10], np.float64)
    print("\nloop shape (descend passes per step pass) -> weighted issues per warp; lanes-per-issue of descend / step")
    for k in (0, 1, 2, 3, 4):
        c = e.warp_profile(tiles.reshape(-1, 8), locate=False, max_desc_passes=k).astype(np.float64)
        cost = float((c[0:10:2] * wts).sum() / c[10])
        print(f"  K = {'inf (product)' if k == 0 else k:<14} {cost:8.1f}   {c[1] / c[0]:5.2f} / {c[5] / c[4]:5.2f}")


if __name__ == "__main__":
    if "--warps" in sys.argv:
        sys.argv.remove("--warps")
        warps(sys.argv[1] if len(sys.argv) > 1 else "c3", int(sys.argv[2]) if len(sys.argv) > 2 else 37)
    else:
        main()
