"""Fused (k_path) against wavefront (k_path_front / sort / k_bounce / k_path_gather) path tracing on a workload's scene: ms per
call of trace_path(0, spp) for a list of batch sizes and sort-key shapes, all checked bit-identical to the fused accumulator.
  python tools/wave_bench.py [workload=c4] [spp=16] [iters=3] [--quick] [--json out.json]
Wall-clock around trace_path + sync (the calls are tens of ms); prints one line per variant and, with --json, the table."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import blok_b200  # noqa: E402
from blok_b200 import IMG_ACCUM, PATH_FUSED, PATH_WAVEFRONT, workloads  # noqa: E402

args = [a for a in sys.argv[1:] if not a.startswith("--")]
name = args[0] if len(args) > 0 else "c4"
spp = int(args[1]) if len(args) > 1 else 16
iters = int(args[2]) if len(args) > 2 else 3
quick = "--quick" in sys.argv
wl = workloads.WORKLOADS[name]
W, H = wl["width"], wl["height"]
r = blok_b200.Renderer(W, H, 0)
r.add_voxels(*workloads.voxels(name))
r.set_frame(workloads.camera_for(name).frame(W, H, 0, spp))


def timed(label, **opt):
    o = r.set_path_options(**opt)
    r.reset_accum(); r.trace_path(0, spp); r.sync()
    acc = r.readback(IMG_ACCUM).tobytes()
    ts = []
    for _ in range(iters):
        r.reset_accum(); r.sync()
        t0 = time.perf_counter(); r.trace_path(0, spp); r.sync(); ts.append((time.perf_counter() - t0) * 1e3)
    return {"label": label, "options": o, "ms_min": min(ts), "ms_median": sorted(ts)[len(ts) // 2], "ms_per_spp": min(ts) / spp}, acc


rows = []
base, want = timed("fused k_path", mode=PATH_FUSED)
r.stats_enable(True); r.reset_accum(); r.trace_path(0, spp); r.sync(); steps, rays = r.stats_read(); r.stats_enable(False)
base["rays"] = rays
rows.append(base)
print(f"{name} {W}x{H} {spp} spp: {rays} rays ({rays / (W * H * spp):.3f} per sample)", flush=True)
print(f"{base['label']:44s} {base['ms_min']:9.3f} ms  {rays / base['ms_min'] / 1e6:7.3f} Grays/s", flush=True)
variants = [(b, 8, 6, 2) for b in ((8,) if quick else (2, 4, 8, 16))]
if not quick:
    variants += [(8, dc, sh, mb) for dc, sh, mb in ((4, 6, 5), (16, 6, 5), (8, 5, 5), (8, 4, 6), (8, 5, 6), (16, 5, 6), (8, 7, 4), (1, 6, 5), (8, 6, 1))]
    variants += [(16, dc, sh, mb) for dc, sh, mb in ((8, 5, 6), (16, 5, 6))]
for b, dc, sh, mb in variants:
    if b > spp:
        continue
    row, acc = timed(f"wavefront batch {b} key dir {dc}^3 cell 2^{sh} morton 3x{mb}", mode=PATH_WAVEFRONT, batch_samples=b, dir_cells=dc,
                     origin_cell_shift=sh, morton_bits_per_axis=mb, refill_below=0)
    row["identical_to_fused"] = acc == want
    rows.append(row)
    print(f"{row['label']:44s} {row['ms_min']:9.3f} ms  {rays / row['ms_min'] / 1e6:7.3f} Grays/s  identical={row['identical_to_fused']}", flush=True)
refills = (8, 16, 24) if quick else (4, 8, 12, 16, 20, 24, 28, 32)
for rb in refills:
    row, acc = timed(f"wavefront batch 8, bounce lanes refilled below {rb}", mode=PATH_WAVEFRONT, batch_samples=8, dir_cells=8,
                     origin_cell_shift=6, morton_bits_per_axis=2, refill_below=rb)
    row["identical_to_fused"] = acc == want
    rows.append(row)
    print(f"{row['label']:44s} {row['ms_min']:9.3f} ms  {rays / row['ms_min'] / 1e6:7.3f} Grays/s  identical={row['identical_to_fused']}", flush=True)
r.set_path_options(refill_below=0)
for a in sys.argv:
    if a.startswith("--json"):
        path = sys.argv[sys.argv.index(a) + 1]
        with open(path, "w") as f:
            json.dump({"workload": name, "width": W, "height": H, "spp": spp, "rows": rows}, f, indent=1)
r.close()
