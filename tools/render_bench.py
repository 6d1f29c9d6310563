"""k_render (the reference's full raygen: frame.sampleCount spp -> five G-buffer images) timed alone on a workload's scene.
  python tools/render_bench.py [workload=c3] [spp=8] [iters=5]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import blok_b200  # noqa: E402
from blok_b200 import workloads  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 8
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 5
wl = workloads.WORKLOADS[name]
W, H = wl["width"], wl["height"]
r = blok_b200.Renderer(W, H, 0)
r.add_voxels(*workloads.voxels(name))
r.set_frame(workloads.camera_for(name).frame(W, H, 0, spp))
r.render(); r.sync()
ts = []
for _ in range(iters):
    t0 = time.perf_counter(); r.render(); r.sync(); ts.append((time.perf_counter() - t0) * 1e3)
print(f"lib={os.environ.get('BLOKB_LIB_PATH', 'default')} render {spp} spp {W}x{H}: min {min(ts):.3f} ms median {sorted(ts)[len(ts) // 2]:.3f} ms")
r.close()
