"""trace_path(0, spp) timed alone on a workload's scene, with whatever library BLOKB_LIB_PATH names (old builds included: only calls
that every version of the C-ABI has).  python tools/path_time.py [workload=c4] [spp=16] [iters=3] [fused|wavefront]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import blok_b200  # noqa: E402
from blok_b200 import workloads  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c4"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 3
mode = sys.argv[4] if len(sys.argv) > 4 else None
wl = workloads.WORKLOADS[name]
W, H = wl["width"], wl["height"]
L = blok_b200.load_library()
r = blok_b200.Renderer(W, H, 0)
r.add_voxels(*workloads.voxels(name))
r.set_frame(workloads.camera_for(name).frame(W, H, 0, spp))
if mode:
    try:
        r.set_path_options(mode=blok_b200.PATH_FUSED if mode == "fused" else blok_b200.PATH_WAVEFRONT)
    except blok_b200.BlokbError:
        pass   # a build from before blokb_set_path_options: it only has the fused kernel
r.reset_accum(); r.trace_path(0, spp); r.sync()
ts = []
for _ in range(iters):
    r.reset_accum(); r.sync()
    t0 = time.perf_counter(); r.trace_path(0, spp); r.sync(); ts.append((time.perf_counter() - t0) * 1e3)
print(f"lib={os.path.basename(os.environ.get('BLOKB_LIB_PATH', 'default'))} mode={mode} {name} {spp} spp: min {min(ts):.3f} ms  median {sorted(ts)[len(ts) // 2]:.3f} ms")
r.close()
