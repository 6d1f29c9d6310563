"""One rank's share of a tiled frame on ONE GPU, no exchange: strips (0, G) of the workload's frame — primary pass and a 1-spp path
pass in both schedules. Separates the per-rank kernel time of a tiled run from what the exchange and the fences add.
  python tools/strip_share_time.py [workload=c5] [G=8] [iters=5]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import blok_b200  # noqa: E402
from blok_b200 import PATH_FUSED, PATH_WAVEFRONT, workloads  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c5"
G = int(sys.argv[2]) if len(sys.argv) > 2 else 8
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 5
wl = workloads.WORKLOADS[name]
W, H = wl["width"], wl["height"]
r = blok_b200.Renderer(W, H, 0)
r.add_voxels(*workloads.voxels(name))
r.set_frame(workloads.camera_for(name).frame(W, H, 0, 1))


def timed(fn):
    fn(); r.sync()
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter(); fn(); r.sync(); ts.append((time.perf_counter() - t0) * 1e3)
    return min(ts)


for first, stride in ((0, 1), (0, G)):
    r.set_strips(first, stride)
    tag = "whole frame" if stride == 1 else f"strips (0, {G})"
    print(f"{name} {tag}: primary {timed(lambda: r.trace_primary()):.3f} ms", flush=True)
    for mode, label in ((PATH_FUSED, "fused"), (PATH_WAVEFRONT, "wavefront")):
        r.set_path_options(mode=mode)
        print(f"{name} {tag}: path 1 spp {label} {timed(lambda: (r.reset_accum(), r.trace_path(0, 1))):.3f} ms", flush=True)
r.close()
