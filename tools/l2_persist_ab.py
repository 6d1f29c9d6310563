"""A/B of the persisting-L2 window over the node pool (BLOKB_L2_PERSIST_MB) on a path-traced workload: one process per setting (the
set-aside is a device-wide limit fixed at context creation), CUDA-event time of trace_path(0, spp) on a caller stream, accumulator
checked identical between the settings.
  python tools/l2_persist_ab.py [workload=c4] [spp=16] [iters=5]            -> runs itself once per setting and prints a table
"""
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one(name, spp, iters):
    import torch
    import blok_b200
    from blok_b200 import IMG_ACCUM, workloads
    wl = workloads.WORKLOADS[name]
    W, H = wl["width"], wl["height"]
    r = blok_b200.Renderer(W, H, 0)
    r.add_voxels(*workloads.voxels(name))
    r.set_frame(workloads.camera_for(name).frame(W, H, 0, spp))
    q = torch.cuda.Stream()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    out = {}
    for label, kw in (("primary", None), ("path", None)):
        ts = []
        for i in range(iters + 2):
            with torch.cuda.stream(q):
                flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                if label == "path":
                    r.reset_accum(stream=q.cuda_stream)
                a.record(q)
                if label == "path":
                    r.trace_path(0, spp, stream=q.cuda_stream)
                else:
                    r.trace_primary(stream=q.cuda_stream)
                b.record(q)
            q.synchronize()
            if i >= 2:
                ts.append(a.elapsed_time(b))
        out[label + "_ms"] = min(ts)
        out[label + "_ms_median"] = sorted(ts)[len(ts) // 2]
    out["accum_sha"] = hashlib.sha256(r.readback(IMG_ACCUM).tobytes()).hexdigest()[:16]
    r.close()
    print("RESULT " + json.dumps(out), flush=True)


if __name__ == "__main__":
    if os.environ.get("L2AB_CHILD"):
        one(sys.argv[1], int(sys.argv[2]), int(sys.argv[3]))
        sys.exit(0)
    name = sys.argv[1] if len(sys.argv) > 1 else "c4"
    spp = sys.argv[2] if len(sys.argv) > 2 else "16"
    iters = sys.argv[3] if len(sys.argv) > 3 else "5"
    rows = {}
    for mb in ("0", "8", "16", "32", "64"):
        env = dict(os.environ, L2AB_CHILD="1")
        env.pop("BLOKB_L2_PERSIST_MB", None)
        if mb != "0":
            env["BLOKB_L2_PERSIST_MB"] = mb
        p = subprocess.run([sys.executable, os.path.abspath(__file__), name, spp, iters], env=env, capture_output=True, text=True)
        line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")]
        rows[mb] = json.loads(line[0][7:]) if line else {"error": p.stderr[-400:]}
        print(f"persist {mb:>3s} MB: {rows[mb]}", flush=True)
    shas = {v.get("accum_sha") for v in rows.values() if "accum_sha" in v}
    print("accumulators identical across settings:", len(shas) == 1)
