"""Profiling driver: runs one workload's kernel a few times (for ncu) and prints per-ray DDA statistics.
  python tools/prof.py c3 [iters] [--stats]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import blok_b200  # noqa: E402
from blok_b200 import workloads  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 4
wl = workloads.WORKLOADS[name]
W, H = wl["width"], wl["height"]
world = workloads.build_world(name)
r = blok_b200.Renderer(W, H, 0)
r.add_world(world)
print("world", r.world_stats())
r.set_frame(workloads.camera_for(name).frame(W, H, 0, min(wl["spp"], 8)))


def step():
    if wl["kind"] == "primary":
        r.trace_primary()
    elif wl["kind"] == "primary_shadow":
        r.trace_primary(shadow=True)
    else:
        r.trace_path(0, min(wl["spp"], 8))


ts = []
for i in range(iters):
    t0 = time.perf_counter()
    step()
    r.sync()
    ts.append((time.perf_counter() - t0) * 1e3)
print(f"mode={os.environ.get('BLOKB_PRIMARY_MODE', '-')} refill={os.environ.get('BLOKB_REFILL', '-')} "
      f"min {min(ts):.3f} ms  median {sorted(ts)[len(ts) // 2]:.3f} ms over {iters} iters")
if "--stats" in sys.argv:
    r.stats_enable(True)
    step()
    r.sync()
    steps, rays = r.stats_read()
    rays = rays or W * H
    print(f"DDA steps {steps} rays {rays} steps/ray {steps / rays:.2f}")
r.close()
