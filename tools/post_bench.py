"""Times the reference's full frame on one workload: raygen (sampleCount spp) -> denoise -> postprocess -> end_frame,
per stage with CUDA events on the launching stream, over a slow camera dolly (so the temporal paths are live).
  python tools/post_bench.py [workload=c3] [frames=12] [spp=8]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import blok_b200  # noqa: E402
from blok_b200 import Camera, PostSettings, workloads  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 12
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 8
wl = workloads.WORKLOADS[name]
W, H = wl["width"], wl["height"]
r = blok_b200.Renderer(W, H, 0)
r.add_world(workloads.build_world(name))
base = workloads.camera_for(name)
cams = [Camera((base.position[0] + 0.5 * k, base.position[1], base.position[2] + 0.25 * k), base.yaw + 0.02 * k, base.pitch) for k in range(frames)]
st = PostSettings()
stream = torch.cuda.Stream()
sp = stream.cuda_stream
names = ("render", "denoise", "postprocess")
times = {n: [] for n in names}
for k, cam in enumerate(cams):
    f = cam.frame(W, H, k, spp, prev=cams[k - 1] if k else None, jitter_index=k)
    r.set_frame(f)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    ev[0].record(stream)
    r.render(stream=sp)
    ev[1].record(stream)
    r.denoise(st, stream=sp)
    ev[2].record(stream)
    r.postprocess(st, stream=sp)
    ev[3].record(stream)
    stream.synchronize()
    r.end_frame()
    if k >= 2:
        for i, n in enumerate(names):
            times[n].append(ev[i].elapsed_time(ev[i + 1]))
px = W * H
for n in names:
    t = float(np.median(times[n]))
    print(f"{n:12s} median {t:8.3f} ms   ({px / t / 1e6:8.2f} Gpx/s)")
hl = r.L  # keep lib alive
print("launches per frame:", 1 + 1 + st.atrousIterations + 2)
r.close()
