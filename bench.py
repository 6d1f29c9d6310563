#!/usr/bin/env python
"""bench.py — Mrays/s of the voxel ray-traversal hot path on N B200s (contract: see the task statement / DESIGN.md §6).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c2|c4|c5] [--impl ours|reference] [--no-extras] [--no-cpu]

A step = one frame of the workload (default c3: 2048^3 SVO terrain, 3840x2160 primary rays = 8 294 400 rays).
  N = 1     value: rays/s with the scene and frame constants resident in HBM; CUDA events on the launching stream around every
            step, L2 flushed (256 MiB memset) between steps outside the event pairs.
  N > 1     ONE frame tiled over the N GPUs (strong scaling): rank g traces the 8-row strips g, g + N, ... and its kernel stores
            every hit record (8-byte packed form) into all ranks' images through an NVLink multicast address (P2P stores where the
            box has no multicast); the frame fence is the library's own (blokb_frame_fence: multimem.red counter + acquire spin), a
            4-byte all-reduce where multicast is missing. value = the frame's rays / max over ranks of the device time. The same
            exchange with 16-byte records is extras.tiled_full_records; every rank its own frame, no exchange, is extras.replicas.
  e2e       the same frames through the public API with HOST buffers, two frames in flight (blokb_pipeline_*): per step the
            host builds a FrameUBO, submits it, and the hit records arrive in pinned host memory (at N > 1 each rank delivers
            the strips it traced: the host receives the frame once).
  roofline  algorithmic bytes per ray B = 16 V + 48 A + O; V, A counted by the CPU checker on the
            cpu_baseline sample of the same frame; achieved = rays * B / kernel time; peak = MEASURED_PEAKS.json hbm_gbs.
            roofline.issue = the bound the kernels actually sit under (warp instructions of one launch, from the committed ncu
            capture in profiles/traffic.json, per second of live kernel time, against 148 SMs x 4 issue slots x the sampled clock).
  cpu_baseline / --impl reference   the reference's own sources compiled in place (oracle/_ref, kind "reference") or, when that
            library is absent, the CPU restatement (kind "port"), all host threads, on a strided pixel sample.
  extras    c2 (primary + sun shadow) and c4 (64 spp path traced; at N > 1 spp-sharded and reduced through blokb_reduce_accum),
            each with its own roofline and a parity block against the CPU checker at bench size; at N = 8 also c5 (8192^3, 8K).
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CPU_SAMPLE_STRIDE = 7  # every 7th pixel in row-major order (3840 % 7 != 0 -> a diagonal lattice over the frame)
PATH_SAMPLE_STRIDE = CPU_SAMPLE_STRIDE * 64  # path workloads: full raygen at 64 spp on every 448th pixel


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=100)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--workload", default="c3", choices=["c2", "c3", "c4", "c5", "t1"])
    p.add_argument("--no-extras", action="store_true", help="skip the secondary workloads (c2, c4, c5) and warm-L2 numbers")
    p.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (then roofline uses committed V, A)")
    p.add_argument("--write-counts", action="store_true", help="(--impl reference) store V, A in blok_b200/workload_counts.json")
    return p.parse_args()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
    return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md; MEASURED_PEAKS.json absent)"


class ClockSampler:
    """Samples SM clock / throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._th = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80, "sync_boost": 0x10, "applications_clocks": 0x2}
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.005)

    def __enter__(self):
        if self.nv is not None:
            self._th = threading.Thread(target=self._run, daemon=True)
            self._th.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._th is not None:
            self._th.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": "NVML unavailable"}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------ CPU checker legs
def cpu_checker():
    """(library, kind): the reference's own sources compiled in place when present, else the restatement."""
    from oracle import pyoracle
    if os.path.exists(pyoracle.REF_SO):
        return pyoracle.RefLib(), "reference"
    if not os.path.exists(pyoracle.ORACLE_SO):
        pyoracle.build(ref=False)
    return pyoracle.OracleLib(), "port"


def run_cpu(world, wl, cam, threads=None):
    """Times the CPU checker on the strided sample of the workload's frame. Returns its counts (V, A per ray, rays traced, MAX_ITER cap
    hits) and results (hit records / shadow verdicts / radiance) for the parity blocks."""
    from oracle import pyoracle
    lib, kind = cpu_checker()
    sc = pyoracle.Scene(world.globalNodes, world.globalSubChunks, world.materials)
    W, H = wl["width"], wl["height"]
    ubo = cam.frame(W, H, 0, wl["spp"]).to_numpy()
    threads = threads or os.cpu_count() or 1
    if wl["kind"] == "path":
        stride = PATH_SAMPLE_STRIDE
        t0 = time.perf_counter()
        img = lib.render(sc, ubo, 0, W * H, stride, threads=threads, want=("color",))
        dt = time.perf_counter() - t0
        cnt = img["counters"]
        npx = (W * H + stride - 1) // stride
        rays = int(cnt[3])
        return dict(seconds=dt, pixels=npx, rays=rays, value=rays / dt / 1e6, V=int(cnt[0]) / rays, A=int(cnt[1]) / rays, cap_hits=int(cnt[2]),
                    kind=kind, cores=threads, color=img["color"].reshape(-1, 4)[::stride, :3].copy(), stride=stride,
                    rays_per_pixel=rays / npx,
                    sample=f"every {stride}th pixel of the {W}x{H} frame, {wl['spp']} spp full raygen ({npx} pixels, {rays} rays)")
    rays = lib.primary_rays(ubo)[::CPU_SAMPLE_STRIDE]
    t0 = time.perf_counter()
    h, t, cnt = lib.trace(sc, rays, threads=threads)
    n = len(rays)
    fetch, aabb, caps = int(cnt[0]), int(cnt[1]), int(cnt[2])
    shadow = None
    if wl["kind"] == "primary_shadow":
        srays, cast = shadow_rays(rays, h, t)
        _, st, c2 = lib.trace(sc, srays, any_hit=True, threads=threads)
        n += len(srays)
        fetch += int(c2[0]); aabb += int(c2[1]); caps += int(c2[2])
        shadow = np.zeros(len(rays), np.uint8)       # the product's flag: 0 = no shadow ray cast, 1 = lit, 2 = shadowed
        shadow[cast] = np.where(st >= 0, 2, 1)
    dt = time.perf_counter() - t0
    return dict(seconds=dt, rays=n, value=n / dt / 1e6, V=fetch / n, A=aabb / n, cap_hits=caps, kind=kind, cores=threads,
                hit_fraction=float((t >= 0).mean()), hits=(h, t), shadow=shadow, rays_per_pixel=n / len(rays),
                sample=f"every {CPU_SAMPLE_STRIDE}th pixel (row-major) of the {W}x{H} frame: {n} rays")


def shadow_rays(rays, h, t):
    """Sun shadow rays of the lit primary hits, built like raygen.rgen:280-298 (fp32, unfused). Returns (rays, cast mask)."""
    normals = np.array([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]], np.float32)
    hit = t >= 0
    N = normals[np.where(hit, h[:, 3], 0)]
    d = rays[:, 3:6]
    flip = ((N[:, 0] * d[:, 0] + N[:, 1] * d[:, 1]) + N[:, 2] * d[:, 2]) > 0
    N = np.where(flip[:, None], -N, N)
    s = np.array([0.5, 0.8, 0.3], np.float32)
    sun = (s * (np.float32(1.0) / np.sqrt((s[0] * s[0] + s[1] * s[1]) + s[2] * s[2], dtype=np.float32))).astype(np.float32)
    ndl = (N[:, 0] * sun[0] + N[:, 1] * sun[1]) + N[:, 2] * sun[2]
    cast = hit & (ndl > 0)
    o = (rays[:, 0:3] + d * t[:, None]).astype(np.float32) + (N * np.float32(0.001)).astype(np.float32)
    n = int(cast.sum())
    out = np.concatenate([o[cast], np.tile(sun, (n, 1)), np.full((n, 1), 0.001, np.float32), np.full((n, 1), 1000, np.float32)], 1).astype(np.float32)
    return out, cast


def hit_parity(got, res, got_shadow=None):
    """GPU hit records (HIT_DT, the cpu_baseline sample's pixels) against the CPU checker's: node, material, bits of t, face and
    gl_PrimitiveID, all exact; for primary_shadow also the per-pixel shadow verdict."""
    h, t = res["hits"]
    miss = t < 0
    want_fp = np.where(miss, np.uint32(0xFFFFFFFF), (h[:, 3] | (h[:, 0] << np.uint32(3))).astype(np.uint32))
    bad = (got["node"] != h[:, 1]) | (got["material"] != h[:, 2]) | (got["t"].view(np.uint32) != t.view(np.uint32)) | (got["facePrim"] != want_fp)
    out = {"checked_rays": int(len(t)), "mismatches": int(np.count_nonzero(bad)), "cap_hits": int(res["cap_hits"]),
           "what": f"GPU hit records (node, material, bits of t, face | gl_PrimitiveID << 3) vs the CPU checker ({res['kind']}) on the "
                   "cpu_baseline sample; cap_hits = rays on which the checker ran into the reference's MAX_ITER (outside the parity claim)"}
    if got_shadow is not None and res.get("shadow") is not None:
        out["shadow_checked"] = int(np.count_nonzero(res["shadow"]))
        out["shadow_mismatches"] = int(np.count_nonzero(got_shadow != res["shadow"]))
        out["mismatches"] += out["shadow_mismatches"]
        out["what"] += "; shadow flags (0 none cast / 1 lit / 2 shadowed) vs the checker's any-hit trace of the same shadow rays"
    return out


def radiance_parity(got_rgb, res):
    """Resolved colour of the sampled pixels against the CPU checker's full raygen (same seeds): relative error per pixel =
    max over rgb of |gpu - ref| / (|ref| + 1e-3); tolerance 2e-3 on >= 99.5 % of pixels and on the image mean."""
    ref = res["color"]
    rel = (np.abs(got_rgb - ref) / (np.abs(ref) + 1e-3)).max(-1)
    mean_rel = float(np.abs(got_rgb.mean(0) - ref.mean(0)).max() / (np.abs(ref.mean(0)).max() + 1e-12))
    frac = float(np.mean(rel < 2e-3))
    return {"checked_pixels": int(len(ref)), "max_rel_err": float(rel.max()), "p99_5_rel_err": float(np.percentile(rel, 99.5)),
            "fraction_within_2e-3": frac, "mean_image_rel_err": mean_rel, "cap_hits": int(res["cap_hits"]),
            "passed": bool(frac >= 0.995 and mean_rel < 2e-3 and res["cap_hits"] == 0),
            "what": f"GPU colour (accum / spp, firefly clamp) vs the CPU checker's ({res['kind']}) full raygen on {res['sample']}; "
                    "sinf/cosf differ from glibc's by <= 2 ulp, which can flip a lobe choice on a rare pixel"}


COUNTS_PATH = os.path.join(ROOT, "blok_b200", "workload_counts.json")


def committed_counts(name):
    """V, A (and rays per pixel) of the workload's strided sample as counted by the CPU checker in the build container — used when the
    cpu_baseline leg does not run (N > 1 or --no-cpu). Regenerate with `python bench.py --impl reference --workload X --write-counts`."""
    if os.path.exists(COUNTS_PATH):
        with open(COUNTS_PATH) as f:
            return json.load(f).get(name)
    return None


def out_bytes_per_ray(kind, rays_per_pixel=1.0):
    """O of SURVEY.md §8(d): 16 B hit record per primary ray; +1 B shadow flag per pixel; 32 B per pixel per pass
    (RGBA32F accumulate read + write) amortised over the pass's rays for path tracing."""
    if kind == "primary":
        return 16.0
    if kind == "primary_shadow":
        return 17.0 / rays_per_pixel
    return 32.0 / rays_per_pixel


def roofline_for(name, kind, rays, ms, V, A, rays_per_pixel, va_src, clocks, sms, share=1.0):
    """roofline block of one workload: algorithmic bytes / live kernel time against the measured HBM peak, plus the issue-slot bound.
    share = the fraction of the frame's launch this rank executes (1 / G when one frame is tiled or sample-sharded over G GPUs): the
    committed warp-instruction count is that of the whole frame on one GPU."""
    peak, peak_src = measured_peaks()
    B = 16.0 * V + 48.0 * A + out_bytes_per_ray(kind, rays_per_pixel)
    achieved = rays * B / (ms * 1e-3) / 1e9
    roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
            "bytes_per_ray": B, "V": V, "A": A, "counts_source": va_src, "peak_source": peak_src,
            "kernel": {"primary": "k_primary<false>", "primary_shadow": "k_primary<true>",
                       "path": "k_path_front + radix sort + k_bounce + k_path_gather (wavefront)"}[kind],
            "note": "algorithmic bytes are defined in the reference's 16-B-node / 48-B-sub-chunk format (SURVEY.md §8(d)); the kernels "
                    "read a compact L2-resident layout instead, so DRAM traffic (profiles/) is far below this figure"}
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        with open(tr) as f:
            t = json.load(f).get(name)
        if t:  # T = dram__bytes_read.sum + dram__bytes_write.sum of one step's launches (ncu --set full), source beside it
            roof["traffic"] = t.get("total")
            roof["traffic_detail"] = t
            if t.get("warp_instructions") and (clocks or {}).get("sm_mhz") and sms:
                ach = t["warp_instructions"] * share / (ms * 1e-3) / 1e9
                pk = sms * 4 * clocks["sm_mhz"] * 1e6 / 1e9
                roof["issue"] = {"achieved": ach, "peak": pk, "unit": "G warp-instr/s", "frac": ach / pk,
                                 "threads_per_instruction": t.get("threads_per_instruction")}
                if ach / pk > roof["frac"]:
                    # the roof that actually binds: warp-instruction issue (HBM is a few % busy, profiles/). achieved / peak / frac above stay
                    # the SURVEY 8(d) HBM-algorithmic figure the contract asks for
                    roof["bound"] = "issue"
                    roof["bound_note"] = ("binding roof = warp-instruction issue slots (roofline.issue); achieved / peak / frac are the HBM-algorithmic "
                                          "figure of SURVEY.md 8(d), kept for comparison")
    return roof


# ------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    """The reference's own CPU code on the box's host cores. Maps oracle/ and the host-only helper library only — never the
    product's CUDA binary (scene generator and camera come from blok_b200/libblokbhost.so, the same sources built without CUDA)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from blok_b200 import workloads
    name = args.workload
    wl = workloads.WORKLOADS[name]
    world = workloads.build_world(name)
    cam = workloads.camera_for(name, 0)
    for _ in range(max(args.warmup, 0)):
        run_cpu(world, wl, cam)
    res = None
    rays = 0
    dt = 0.0  # sum of the traced regions (ray set-up and result allocation are outside, as on the GPU arm)
    for _ in range(args.steps):
        res = run_cpu(world, wl, cam)
        rays += res["rays"]
        dt += res["seconds"]
    value = rays / dt / 1e6
    line = {"impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{name}: {wl['desc']}", "step": "one bounded CPU sample of the frame: " + res["sample"]},
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": res["cores"], "kind": res["kind"], "sample": res["sample"]},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    line["counts"] = {"V": res["V"], "A": res["A"], "rays_per_pixel": res["rays_per_pixel"], "cap_hits": res["cap_hits"]}
    if "hit_fraction" in res:
        line["counts"]["hit_fraction"] = res["hit_fraction"]
    if getattr(args, "write_counts", False):
        allc = {}
        if os.path.exists(COUNTS_PATH):
            with open(COUNTS_PATH) as f:
                allc = json.load(f)
        allc.setdefault(name, {}).update(dict(line["counts"], kind=res["kind"], sample=res["sample"]))
        with open(COUNTS_PATH, "w") as f:
            json.dump(allc, f, indent=1, sort_keys=True)
    emit(line)


# ------------------------------------------------------------------------------------------ our arm
class Env:
    """Per-process plumbing: device, torch.distributed group, launching stream, L2 flush, reductions over ranks."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world_size = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world_size > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.stream = torch.cuda.Stream(device=self.dev)
        self.sp = self.stream.cuda_stream
        self.flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2
        self.sms = torch.cuda.get_device_properties(self.dev).multi_processor_count

    def barrier(self):
        if self.world_size > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def _all(self, x, op):
        if self.world_size == 1:
            return float(x)
        t = self.torch.tensor([float(x)], device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=op)
        return float(t.item())

    def max(self, x):
        return self._all(x, self.dist.ReduceOp.MAX)

    def min(self, x):
        return self._all(x, self.dist.ReduceOp.MIN)

    def sum(self, x):
        return self._all(x, self.dist.ReduceOp.SUM)

    def flush_l2(self):
        with self.torch.cuda.stream(self.stream):
            self.flush_buf.zero_()

    def timed(self, step, steps, warmup, flush=True):
        """W warm-up steps, then K steps each inside its own CUDA-event pair on the launching stream (L2 flushed between them, outside
        the pairs), bracketed by barrier + synchronize on both sides. Returns (mean ms of this rank, max over ranks, clock summary)."""
        torch = self.torch
        for _ in range(warmup):
            step()
        self.barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        with ClockSampler(self.local) as clk:
            self.barrier()
            for a, b in ev:
                if flush:
                    self.flush_l2()
                a.record(self.stream)
                step()
                b.record(self.stream)
            self.barrier()
        ms = sum(a.elapsed_time(b) for a, b in ev) / steps
        return ms, self.max(ms), clk.summary()

    def timed_back_to_back(self, step, steps):
        """Same steps without flushes inside ONE event pair (the steady state of a render loop)."""
        torch = self.torch
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        a.record(self.stream)
        for _ in range(steps):
            step()
        b.record(self.stream)
        self.barrier()
        return self.max(a.elapsed_time(b) / steps)


def open_workload(env, name):
    """Scene on this rank's GPU (replicated), renderer sized for the workload, frame constants of camera 0."""
    import blok_b200
    from blok_b200 import workloads
    wl = workloads.WORKLOADS[name]
    t0 = time.perf_counter()
    world = workloads.build_world(name)
    gen_s = time.perf_counter() - t0
    r = blok_b200.Renderer(wl["width"], wl["height"], env.local)
    r.add_world(world)
    cam = workloads.camera_for(name, 0)
    ubo = cam.frame(wl["width"], wl["height"], 0, wl["spp"])
    r.set_frame(ubo)
    return dict(name=name, wl=wl, world=world, renderer=r, cam=cam, ubo=ubo, gen_s=gen_s, stats=r.world_stats())


def count_rays(env, ctx, s0=0, s1=None):
    """Rays one step casts on this rank (deterministic): primary rays + the shadow / bounce rays the kernels actually trace."""
    from blok_b200 import IMG_SHADOW
    r, wl = ctx["renderer"], ctx["wl"]
    n = wl["width"] * wl["height"]
    if wl["kind"] == "primary":
        return n
    if wl["kind"] == "primary_shadow":
        r.trace_primary(shadow=True)
        return n + int(np.count_nonzero(r.readback(IMG_SHADOW)))
    r.stats_enable(True)
    r.reset_accum()
    r.trace_path(s0, wl["spp"] if s1 is None else s1)
    r.sync()
    rays = r.stats_read()[1]
    r.stats_enable(False)
    return rays


def measure_single(env, ctx, steps, warmup, cam_rank=0):
    """Device-timed numbers of a primary / primary_shadow workload rendered whole on this rank's GPU (N = 1 headline; replicas extra)."""
    from blok_b200 import workloads
    r, wl = ctx["renderer"], ctx["wl"]
    W, H = wl["width"], wl["height"]
    if cam_rank:
        r.set_frame(workloads.camera_for(ctx["name"], cam_rank).frame(W, H, 0, wl["spp"]))
    shadow = wl["kind"] == "primary_shadow"
    rays = count_rays(env, ctx)

    def step():
        r.trace_primary(shadow=shadow, stream=env.sp)

    l0 = r.launch_count()
    ms, ms_max, clocks = env.timed(step, steps, warmup)
    launches = (r.launch_count() - l0) * steps // (steps + warmup)
    warm = env.timed_back_to_back(step, steps)
    if cam_rank:
        r.set_frame(ctx["ubo"])
    return dict(rays_rank=rays, ms_rank=ms, ms=ms_max, launches=launches, clocks=clocks, ms_warm_l2=warm)


def e2e_pipeline(env, ctx, steps, packed=False):
    """Frames in flight through the public API with host buffers (blokb_pipeline_*, depth 2 like the reference's 2 frames in flight,
    blok/include/renderer.hpp:185): per step the host builds the FrameUBO from the Camera and submits the frame (primary pass on the
    context's stream, hit records copied to pinned host memory on a copy stream), then reads the PREVIOUS frame's records. With strips
    set (a tiled frame) each rank traces and delivers only its own strips: the host receives every frame exactly once."""
    import blok_b200
    r, wl, cam = ctx["renderer"], ctx["wl"], ctx["cam"]
    W, H, spp = wl["width"], wl["height"], wl["spp"]
    rec = 8 if packed else 16
    pipe = blok_b200.FramePipeline(r, W, H, depth=2, packed=packed)
    key = "nodeFace" if packed else "node"
    G, g = env.world_size, env.rank
    probe_row = min(H - 1, (g if G > 1 else (H // 16)) * 8)   # a row this rank delivers
    n = max(3, min(steps, 50))
    for i in range(2):
        pipe.result(pipe.submit(cam.frame(W, H, i, spp)))
    pipe.drain()
    env.barrier()
    t0 = time.perf_counter()
    last, seen = None, 0
    for i in range(n):
        tk = pipe.submit(cam.frame(W, H, i, spp))
        if last is not None:
            seen += int(pipe.result(last)[key][probe_row, W // 2] != 0xFFFFFFFF)
        last = tk
    seen += int(pipe.result(last)[key][probe_row, W // 2] != 0xFFFFFFFF)
    pipe.drain()
    env.barrier()
    dt = env.max(time.perf_counter() - t0)
    pipe.close()
    strips = (H + 7) // 8
    mine = (strips - g + G - 1) // G if G > 1 else strips
    d2h = min(mine * 8, H) * W * rec if G > 1 else W * H * rec
    return {"value": W * H * n / dt / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 560, "d2h_bytes_per_step": int(env.sum(d2h)),
            "d2h_bytes_per_step_per_gpu": int(d2h), "steps": n, "ms_per_step": dt / n * 1e3,
            "what": ("FramePipeline (blokb_pipeline_*, 2 frames in flight): Camera -> FrameUBO on host, submit = primary pass + async copy of the "
                     f"{rec}-B hit records to pinned host memory, result(previous frame) read on the host; scene uploaded once beforehand "
                     "(Renderer::addWorld)" + ("; frame tiled over the GPUs: each rank traces and delivers its own 8-row strips" if G > 1 else ""))}


def measure_tiled(env, ctx, steps, warmup, packed=False):
    """ONE frame tiled over the G GPUs (pixel partition): rank g traces strips g, g + G, ... (blokb_set_strips) and its
    kernel stores each record into every rank's image — one multimem.st through the NVSwitch when the box has NVLink multicast
    (blokb_bind_image), else P2P stores into IPC-mapped peer images (blokb_set_peers); a 4-byte all-reduce on the dispatch stream is
    the frame fence. Parity: the assembled image on every rank is byte-identical to that rank's own undivided dispatch."""
    from blok_b200 import IMG_HIT, IMG_HIT_PACKED, sharding
    torch = env.torch
    r, wl = ctx["renderer"], ctx["wl"]
    W, H = wl["width"], wl["height"]
    G, rank = env.world_size, env.rank
    img = IMG_HIT_PACKED if packed else IMG_HIT
    r.set_strips(0, 1)
    r.set_frame(ctx["ubo"])
    r.trace_primary(packed=packed)
    full = r.readback(img).copy()
    # this rank's share of the frame with NO exchange bound (stores stay local): what the traversal of 1/G of the strips costs by itself —
    # launch + wave tail; trace_only_ms below minus this is what the peer stores add, ms minus trace_only_ms what the fence adds
    r.set_strips(*sharding.strip_set(rank, G))
    _, share_ms, _ = env.timed(lambda: r.trace_primary(packed=packed, stream=env.sp), max(5, steps // 4), 2, flush=False)
    r.set_strips(0, 1)
    keep = sharding.connect_multicast(r, (img,))
    if keep is None:
        sharding.connect_peers(r, (img,))
    r.set_strips(*sharding.strip_set(rank, G))
    token = torch.zeros(1, device=env.dev)
    fence_keep = sharding.connect_fence(r) if keep is not None else None   # the library's own fence over a multicast word, else NCCL

    def step():
        r.trace_primary(packed=packed, stream=env.sp)
        if fence_keep is not None:
            r.frame_fence(stream=env.sp)
        else:
            with torch.cuda.stream(env.stream):
                sharding.frame_barrier(token)

    l0 = r.launch_count()
    ms, ms_max, clocks = env.timed(step, steps, warmup)
    launches = (r.launch_count() - l0) * steps // (steps + warmup)
    warm = env.timed_back_to_back(step, steps)
    # the traversal alone (no fence): what the exchange adds on top
    def trace_only():
        r.trace_primary(packed=packed, stream=env.sp)
    _, trace_ms, _ = env.timed(trace_only, max(5, steps // 4), 2, flush=False)
    step()
    env.barrier()
    got_img = r.readback(img)
    same = env.min(int(got_img.tobytes() == full.tobytes()))
    assembled = got_img.reshape(-1)[::CPU_SAMPLE_STRIDE].copy() if rank == 0 else None   # checked against the CPU checker by the caller
    del got_img
    fence_name = "4-byte NCCL all-reduce"
    if fence_keep is not None:
        r.fence_status()
        sharding.disconnect_fence(r, fence_keep)
        fence_name = "blokb_frame_fence (multimem.red on a multicast counter + acquire spin, one thread)"
    if keep is None:
        sharding.disconnect_peers(r, (img,))
    else:
        sharding.disconnect_multicast(r, keep)
    r.set_strips(0, 1)
    recb = 8 if packed else 16
    return dict(ms=ms_max, ms_rank=ms, launches=launches, clocks=clocks, ms_warm_l2=warm, identical=bool(same), trace_only_ms=trace_ms,
                exchange="NVLink multicast (multimem.st into symmetric memory)" if keep is not None else "P2P stores into IPC-mapped peer images",
                fence=fence_name, assembled_sample=assembled, share_only_ms=share_ms,
                nvlink_ingress_bytes_per_gpu=int(W * H * recb * (G - 1) / G), record_bytes=recb)


def measure_path(env, ctx, steps, warmup, e2e_steps=2):
    """c4: the frame's spp sharded over the ranks (raygen's per-(pixel, frame, sample) seeds => the union is the single-GPU sample set),
    per-GPU accumulation buffers reduced and resolved through the C-ABI (blokb_reduce_accum over the context's own NCCL communicator).
    value: device time of reset + trace + reduce (every rank ends with the whole colour image). e2e: the same frame with host buffers —
    set_frame (560 B in), reduce to slices, every rank reads back ITS slice of the colour image (the host receives the frame once)."""
    from blok_b200 import IMG_COLOR, REDUCE_ALL, REDUCE_SLICE, sharding
    torch = env.torch
    r, wl = ctx["renderer"], ctx["wl"]
    W, H, spp = wl["width"], wl["height"], wl["spp"]
    G, rank = env.world_size, env.rank
    s0, s1 = sharding.sample_range(rank, G, spp)
    if G > 1:
        sharding.connect_comm(r)
    rays = count_rays(env, ctx, s0, s1)

    def step(mode=REDUCE_ALL):
        r.reset_accum(stream=env.sp)
        r.trace_path(s0, s1, stream=env.sp)
        r.reduce_accum(mode, stream=env.sp)

    l0 = r.launch_count()
    ms, ms_max, clocks = env.timed(step, steps, warmup)
    launches = (r.launch_count() - l0) * steps // (steps + warmup)
    first, count = r.reduce_slice() if G > 1 else (0, W * H)
    nbytes = count * 16
    host = torch.empty(max(nbytes, 16), dtype=torch.uint8, pin_memory=True)
    step(REDUCE_SLICE)
    env.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        r.set_frame(ctx["ubo"])
        step(REDUCE_SLICE)
        if G > 1:
            r.readback_slice_into(IMG_COLOR, host.data_ptr(), nbytes)
        else:
            r.readback_into(IMG_COLOR, host.data_ptr(), nbytes)
    env.barrier()
    dt = env.max(time.perf_counter() - t0)
    total = env.sum(rays)
    e2e = {"value": total * e2e_steps / dt / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 560, "d2h_bytes_per_step": int(env.sum(nbytes)),
           "d2h_bytes_per_step_per_gpu": int(nbytes), "steps": e2e_steps, "ms_per_step": dt / e2e_steps * 1e3,
           "what": "blokb_set_frame, reset_accum, trace_path(sample shard), blokb_reduce_accum(SLICE) = reduce-scatter + resolve of the owned "
                   "slice, blokb_readback_slice of the RGBA32F colour slice into pinned host memory"}
    step(REDUCE_ALL)
    env.barrier()
    mc = None
    if G > 1:
        # the same frame with accum in symmetric memory: the sum over the GPUs happens in the NVSwitch (multimem.ld_reduce inside
        # k_reduce_resolve_mc) and the resolved texels are broadcast by multicast stores — timed, and left in the colour image for the parity check
        nccl_color = r.readback(IMG_COLOR).reshape(-1, 4)[::PATH_SAMPLE_STRIDE, :3].copy() if rank == 0 else None
        try:
            from blok_b200 import IMG_ACCUM
            keep = sharding.connect_multicast(r, (IMG_ACCUM, IMG_COLOR), reduce_only=(IMG_ACCUM,))
            if keep is None:
                mc = {"unavailable": "no NVLink multicast on this box"}
            else:
                _, mc_ms, _ = env.timed(step, steps, warmup)
                step(REDUCE_ALL)
                r.sync(); env.barrier()
                mc = {"ms_per_step": mc_ms, "value": total / (mc_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                      "what": "blokb_bind_accum_for_reduce + colour bound to a multicast address: blokb_reduce_accum = two 4-byte all-reduces around ONE "
                              "kernel (k_reduce_resolve_mc: multimem.ld_reduce of the G accum copies, resolve, multimem.st of the colour texel)"}
                if rank == 0:
                    got = r.readback(IMG_COLOR).reshape(-1, 4)[::PATH_SAMPLE_STRIDE, :3].copy()
                    mc["color_sample"] = got
                    mc["max_abs_diff_vs_nccl_path"] = float(np.abs(got - nccl_color).max())
                env.barrier()
                sharding.disconnect_multicast(r, keep)
                step(REDUCE_ALL)      # leave the NCCL-path result in the (context-owned) colour image again
                env.barrier()
        except Exception as ex:
            mc = {"error": repr(ex)}
    return dict(rays_rank=rays, rays_total=total, ms=ms_max, ms_rank=ms, launches=launches, clocks=clocks, e2e=e2e, samples=(s0, s1), mc=mc)


def extra_device_build(env, ctx):
    """Scene upload on the device (blokb_upload_voxels) beside the host path the headline used, and the
    full-size parity property: both leave byte-identical compact arrays (incl. the reference node index of every leaf block)."""
    import hashlib
    import blok_b200
    from blok_b200 import WORLD_BLOCKFIRSTLEAF, WORLD_NODES, WORLD_VOXELMAT, workloads
    r = ctx["renderer"]
    xyz, mid, mats = workloads.voxels(ctx["name"])
    b = blok_b200.Renderer(64, 64, env.local)
    b.add_voxels(xyz[:4096], mid[:4096], mats)
    ts = []
    for _ in range(3):
        t0 = time.perf_counter()
        b.add_voxels(xyz, mid, mats)
        ts.append(time.perf_counter() - t0)
    same = all(hashlib.sha256(r.world_export(k).tobytes()).digest() == hashlib.sha256(b.world_export(k).tobytes()).digest()
               for k in (WORLD_NODES, WORLD_BLOCKFIRSTLEAF, WORLD_VOXELMAT))
    st = b.world_stats()
    b.close()
    return {"what": "blokb_upload_voxels: voxel list H2D (pageable) + sort/dedupe + compact 64-tree + reference node numbering on the GPU",
            "voxels": int(len(mid)), "wall_ms": min(ts) * 1e3, "input_bytes": int(xyz.nbytes + mid.nbytes),
            "identical_to_host_path": bool(same), "nodes_reference_format": int(st["nodes"])}


def extra_full_frame(env, ctx):
    """The reference's whole frame (Renderer::render, blok/src/renderer_draw.cpp): raygen at the reference's 8 spp, Denoiser::denoise,
    PostProcess::process, over a slow dolly so that both temporal passes are live. Per stage, CUDA events on the launching stream."""
    from blok_b200 import Camera, PostSettings, workloads
    torch = env.torch
    r, wl = ctx["renderer"], ctx["wl"]
    W, H = wl["width"], wl["height"]
    base = workloads.camera_for(ctx["name"])
    frames = 8
    cams = [Camera((base.position[0] + 0.5 * k, base.position[1], base.position[2] + 0.25 * k), base.yaw + 0.02 * k, base.pitch) for k in range(frames)]
    st = PostSettings()
    stream = torch.cuda.Stream()
    sp = stream.cuda_stream
    names = ("raygen_8spp", "denoise", "postprocess")
    times = {n: [] for n in names}
    l0 = r.launch_count()
    for k, cam in enumerate(cams):
        r.set_frame(cam.frame(W, H, k, 8, prev=cams[k - 1] if k else None, jitter_index=k))
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
    out = {n + "_ms": float(np.median(times[n])) for n in names}
    out["frame_ms"] = sum(out.values())
    out["launches_per_frame"] = (r.launch_count() - l0) // frames
    out["what"] = f"{W}x{H}, 8 spp raygen (wavefront) + temporal/variance + 4 a-trous + TAA/tonemap + sharpen (reference defaults)"
    r.set_frame(ctx["ubo"])
    return out


def c2_lit_side(env, ctx):
    """c2 once more from the sunny side. From the SURVEY 8(d) pose almost every visible face looks away from the sun, so only ~2 % of the
    hits cast a shadow ray; this camera sits 40 degrees off the sun direction (0.5, 0.8, 0.3), where most hits do: the shadow-ray half of
    k_primary<true> measured and checked (flags vs the checker's any-hit trace) at bench size."""
    from blok_b200 import IMG_HIT, IMG_SHADOW, workloads
    r, wl = ctx["renderer"], ctx["wl"]
    W, H, N = wl["width"], wl["height"], wl["N"]
    cam = workloads._look_at((1.59 * N, 1.105 * N, 0.258 * N), (0.5 * N, 0.5 * N, 0.5 * N), 40.0)
    ubo = cam.frame(W, H, 0, 1)
    r.set_frame(ubo)
    try:
        m = measure_single(env, dict(ctx, cam=cam, ubo=ubo), 10, 3)
        res = run_cpu(ctx["world"], wl, cam)
        r.trace_primary(shadow=True)
        got = r.readback(IMG_HIT).reshape(-1)[::CPU_SAMPLE_STRIDE]
        sh = r.readback(IMG_SHADOW).reshape(-1)[::CPU_SAMPLE_STRIDE]
        par = hit_parity(got, res, sh)
        hits = int(np.count_nonzero(got["node"] != 0xFFFFFFFF))
        return {"camera": "(1.59N, 1.105N, 0.258N) -> centre, fov 40: 40 degrees off the sun direction", "ms_per_step": m["ms"],
                "rays_per_step": m["rays_rank"], "value": m["rays_rank"] / (m["ms"] * 1e-3) / 1e6, "unit": "Mrays/s",
                "shadow_rays_per_step": m["rays_rank"] - W * H, "hit_fraction": res["hit_fraction"],
                "shadow_rays_per_hit_in_sample": (par.get("shadow_checked", 0) / hits) if hits else None, "parity": par}
    finally:
        r.set_frame(ctx["ubo"])


def mc_block(m, res):
    """extras block of measure_path's multicast-reduction leg; its colour sample is compared with the CPU checker when one ran."""
    mc = m.get("mc")
    if not mc:
        return None
    mc = dict(mc)
    got = mc.pop("color_sample", None)
    if got is not None and res is not None:
        mc["parity"] = radiance_parity(got, res)
    return mc


def extra_workload(env, name, steps, warmup, do_cpu, do_check=None):
    """A secondary workload measured like the headline: value, roofline (V, A from the CPU checker on this run's sample at N = 1,
    else the committed counts) and a parity block against the checker at bench size."""
    from blok_b200 import IMG_COLOR, IMG_HIT, IMG_SHADOW
    ctx = open_workload(env, name)
    r, wl = ctx["renderer"], ctx["wl"]
    W, H, kind = wl["width"], wl["height"], wl["kind"]
    G = env.world_size
    try:
        if kind == "path":
            m = measure_path(env, ctx, steps, warmup)
            rays_total, ms = m["rays_total"], m["ms"]
            out = {"workload": wl["desc"], "value": rays_total / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": ms, "rays_per_step": rays_total,
                   "scaling": "strong (spp sharded over the ranks; blokb_reduce_accum: reduce-scatter + resolve + all-gather over NCCL)" if G > 1 else "one GPU",
                   "samples_this_rank": list(m["samples"]), "gpu_launches": m["launches"], "e2e": m["e2e"], "clocks": m["clocks"],
                   "path_options": r.set_path_options()}
        else:
            # replicas: every rank renders the whole frame (at N > 1 a weak-scaling figure; the strong one is the tiled headline)
            m = measure_single(env, ctx, steps, warmup)
            rays_total = env.sum(m["rays_rank"])
            ms = m["ms"]
            out = {"workload": wl["desc"], "value": rays_total / (ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": ms, "rays_per_step": rays_total,
                   "value_warm_l2": rays_total / (m["ms_warm_l2"] * 1e-3) / 1e6, "scaling": "weak (every rank renders the frame)" if G > 1 else "one GPU",
                   "gpu_launches": m["launches"], "clocks": m["clocks"]}
        res = None
        if do_check is None:
            do_check = do_cpu
        if do_cpu or do_check:   # at N > 1 rank 0 still runs the checker — as the parity check of the exchanged / reduced result, not as a baseline
            res = run_cpu(ctx["world"], wl, ctx["cam"])
        if do_cpu:
            V, A, src = res["V"], res["A"], f"counted by the CPU checker ({res['kind']}) on this run's sample: {res['sample']}"
            out["cpu_baseline"] = {"value": res["value"], "unit": "Mrays/s", "cores": res["cores"], "kind": res["kind"], "sample": res["sample"]}
        else:
            cc = committed_counts(name)
            V, A, src = (cc["V"], cc["A"], "blok_b200/workload_counts.json (CPU checker counts of the same sample, committed)") if cc else (None, None, None)
        if V is not None:
            rpp = rays_total / (W * H) / (1 if kind == "path" else G)
            out["roofline"] = roofline_for(name, kind, m["rays_rank"], m["ms_rank"], V, A, rpp, src, m["clocks"], env.sms,
                                           share=(1.0 / G if kind == "path" else 1.0))
        if res is not None:
            r.set_frame(ctx["ubo"])
            if kind == "path":
                got = r.readback(IMG_COLOR).reshape(-1, 4)[::res["stride"], :3]
                out["parity"] = radiance_parity(got, res)
                if G > 1:
                    out["parity"]["what"] = f"the REDUCED image of the {G} ranks' sample shards (blokb_reduce_accum over NCCL): " + out["parity"]["what"]
            else:
                r.trace_primary(shadow=(kind == "primary_shadow"))
                got = r.readback(IMG_HIT).reshape(-1)[::CPU_SAMPLE_STRIDE]
                sh = r.readback(IMG_SHADOW).reshape(-1)[::CPU_SAMPLE_STRIDE] if kind == "primary_shadow" else None
                out["parity"] = hit_parity(got, res, sh)
                out["hit_fraction"] = res["hit_fraction"]
        if kind == "path" and m.get("mc"):
            out["multicast_reduce"] = mc_block(m, res)
        if name == "c2" and do_cpu:
            try:
                out["lit_side"] = c2_lit_side(env, ctx)
            except Exception as ex:
                out["lit_side"] = {"error": repr(ex)}
        return out
    finally:
        r.close()


def extra_c5(env, check=True):
    """BASELINE configs[4] under the bench: 8192^3 terrain (201 M voxels, built by the device builder from the voxel list), one
    7680x4320 frame tiled over the ranks with the multicast exchange — primary rays, and a 1-spp path-traced frame whose incoherent
    bounce rays go through the wavefront reordering. Parity at full size: (a) rank 0 packs the same voxel list into the reference's
    arrays (host) and the CPU checker traces every 251st pixel: hit records bit-exact; (b) the size-independent heightfield properties
    (tests/heightfield.py) on 20 000 random pixels."""
    import blok_b200
    from blok_b200 import IMG_ACCUM, IMG_COLOR, IMG_HIT, IMG_HIT_PACKED, sharding, workloads
    from blok_b200.api import WorldSvoGpu
    torch = env.torch
    G, rank = env.world_size, env.rank
    wl = workloads.WORKLOADS["c5"]

    def colour_sample():
        """One more fenced 1-spp frame; rank 0 keeps every 251st pixel of the resolved colour for the check against the CPU raygen."""
        p_path()
        fence()
        torch.cuda.synchronize()
        env.barrier()
        return r.readback(IMG_COLOR).reshape(-1, 4)[::251, :3].copy() if rank == 0 else None
    W, H, N = wl["width"], wl["height"], wl["N"]
    t0 = time.perf_counter()
    xyz, mid, mats = workloads.voxels("c5")
    t_gen = time.perf_counter() - t0
    r = blok_b200.Renderer(W, H, env.local)
    t0 = time.perf_counter()
    r.add_voxels(xyz, mid, mats)
    r.sync()
    t_build = time.perf_counter() - t0
    stats = r.world_stats()
    cam = workloads.camera_for("c5", 0)
    ubo = cam.frame(W, H, 0, 1)
    r.set_frame(ubo)
    out = {"workload": wl["desc"], "n_gpus": G, "scene_gen_s": t_gen, "device_build_wall_s": t_build,
           "world": {k: stats[k] for k in ("nodes", "subChunks", "voxels", "bricks", "compactBytes", "levels")}}
    images = (IMG_HIT, IMG_HIT_PACKED, IMG_ACCUM)
    keep = None
    if G > 1:
        keep = sharding.connect_multicast(r, images)
        if keep is None:
            sharding.connect_peers(r, images)
        out["exchange"] = "NVLink multicast (multimem.st)" if keep is not None else "P2P stores into IPC-mapped peer images"
    r.set_strips(*sharding.strip_set(rank, G))
    token = torch.zeros(1, device=env.dev)
    fence_keep = sharding.connect_fence(r) if keep is not None else None
    out["fence"] = "blokb_frame_fence" if fence_keep is not None else "4-byte NCCL all-reduce"

    def fence():
        if fence_keep is not None:
            r.frame_fence(stream=env.sp)
        else:
            with torch.cuda.stream(env.stream):
                sharding.frame_barrier(token)

    def p_primary():
        r.trace_primary(stream=env.sp)
        fence()

    def p_primary_packed():
        r.trace_primary(packed=True, stream=env.sp)
        fence()

    def p_path():
        fence()                         # nobody may still be reading last frame's accum when it is cleared
        r.reset_accum(stream=env.sp)
        fence()                         # every rank's clear has landed before peers store into it
        r.trace_path(0, 1, stream=env.sp)
        fence()
        r.resolve(stream=env.sp)

    _, ms, clocks = env.timed(p_primary_packed, 10, 3)
    out["primary"] = {"ms_per_frame": ms, "rays": W * H, "value": W * H / ms / 1e3, "unit": "Mrays/s",
                      "what": "8-byte packed hit records (node | face << 29, t) through the exchange, every rank ends with the whole image"}
    _, ms16, _ = env.timed(p_primary, 10, 3)
    out["primary_full_records"] = {"ms_per_frame": ms16, "rays": W * H, "value": W * H / ms16 / 1e3, "unit": "Mrays/s",
                                   "what": "16-byte records through the exchange (465 MB of NVLink ingress per GPU at 8 GPUs)"}
    out["clocks"] = clocks
    r.stats_enable(True)
    r.reset_accum(stream=env.sp)
    r.trace_path(0, 1, stream=env.sp)
    torch.cuda.synchronize()
    rays_path = env.sum(r.stats_read()[1])
    r.stats_enable(False)
    _, ms, _ = env.timed(p_path, 5, 2)
    out["path_1spp"] = {"ms_per_frame": ms, "rays": rays_path, "value": rays_path / ms / 1e3, "unit": "Mrays/s", "path_options": r.set_path_options(),
                        "what": "primary + sun shadow + one incoherent bounce ray per pixel; bounce rays sorted by direction cell + origin Morton code "
                                "before they are traced (wavefront), accum strips exchanged by the kernel's stores, resolve"}
    path_cols = {"path_1spp": colour_sample()}
    try:   # the same frame with the fused kernel (no reordering): its accum stores leave DURING the traversal instead of after it
        from blok_b200 import PATH_FUSED, PATH_WAVEFRONT
        r.set_path_options(mode=PATH_FUSED)
        _, msf, _ = env.timed(p_path, 5, 2)
        out["path_1spp_fused"] = {"ms_per_frame": msf, "rays": rays_path, "value": rays_path / msf / 1e3, "unit": "Mrays/s",
                                  "what": "BLOKB_PATH_FUSED: one thread walks camera, shadow and bounce ray back to back, bounce rays in pixel order"}
        path_cols["path_1spp_fused"] = colour_sample()
        r.set_path_options(mode=PATH_WAVEFRONT)
    except Exception as ex:
        out["path_1spp_fused"] = {"error": repr(ex)}
    # ---- parity at full size (rank 0 holds the whole hit image after the fenced dispatch)
    p_primary()
    env.barrier()
    if rank == 0:
        full = r.readback(IMG_HIT)
        try:
            if not check:
                raise InterruptedError
            from oracle import pyoracle
            lib, kind = cpu_checker()
            t0 = time.perf_counter()
            world = WorldSvoGpu.from_voxels(xyz, mid, mats)
            t_pack = time.perf_counter() - t0
            sc = pyoracle.Scene(world.globalNodes, world.globalSubChunks, world.materials)
            rays = lib.primary_rays(ubo.to_numpy())[::251]
            h, t, cnt = lib.trace(sc, rays, threads=os.cpu_count() or 1)
            res = dict(hits=(h, t), cap_hits=int(cnt[2]), kind=kind)
            par = hit_parity(full.reshape(-1)[::251], res)
            par["what"] = par["what"].replace("on the cpu_baseline sample", f"on every 251st pixel of the 8K frame (reference-format arrays packed on the host in {t_pack:.0f} s)")
            out["parity"] = par
            # the path-traced frames (three fences, the accum exchange and the resolve sit between the kernels and these pixels)
            img = lib.render(sc, ubo.to_numpy(), 0, W * H, 251, threads=os.cpu_count() or 1, want=("color",))
            pres = dict(color=img["color"].reshape(-1, 4)[::251, :3].copy(), cap_hits=int(img["counters"][2]), kind=kind,
                        sample=f"every 251st pixel of the {W}x{H} frame, 1 spp full raygen")
            del img
            for key, col in path_cols.items():
                if col is not None and isinstance(out.get(key), dict) and "error" not in out[key]:
                    out[key]["parity"] = radiance_parity(col, pres)
            del world, sc
        except InterruptedError:
            out["parity"] = {"skipped": "packing the 8192^3 scene into the reference's arrays takes ~55 s of host time: the CPU checker runs at N = 8 only; "
                                        "the heightfield properties below are checked at every N"}
        except Exception as ex:
            out["parity"] = {"error": repr(ex)}
        try:
            from tests import heightfield
            from blok_b200 import scenes
            rng = np.random.default_rng(7)
            n = 20000
            px, py = rng.integers(0, W, n), rng.integers(0, H, n)
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            from c5_run import primary_rays_np
            rays = primary_rays_np(ubo, px, py)
            img = full[py, px]
            chk = heightfield.check_hits(scenes.terrain_heights(N), 3, rays, img["node"], img["material"], img["t"], img["facePrim"] & 7, eps=5e-2)
            out["heightfield_properties"] = {k: chk[k] for k in ("bad_hit_voxel", "bad_material", "filled_before_hit", "filled_on_missed_ray", "unfinished")}
            out["heightfield_properties"]["pixels"] = n
            out["heightfield_properties"]["passed"] = bool(heightfield.passed(chk))
        except Exception as ex:
            out["heightfield_properties"] = {"error": repr(ex)}
    env.barrier()
    if fence_keep is not None:
        sharding.disconnect_fence(r, fence_keep)
    if G > 1:
        if keep is None:
            sharding.disconnect_peers(r, images)
        else:
            sharding.disconnect_multicast(r, keep)
    r.close()
    return out


_REAL_STDOUT = None


def parity_summary(extras):
    """Every parity block of the line in ~300 bytes: {where: mismatches (hit records, exact) | passed (radiance, tolerance 2e-3)}."""
    out = {}

    def walk(prefix, d):
        if not isinstance(d, dict):
            return
        p = d.get("parity")
        if isinstance(p, dict):
            if "error" in p:
                out[prefix] = "error"
            elif "skipped" in p:
                out[prefix] = "skipped"
            elif "mismatches" in p:
                out[prefix] = {"mismatches": p["mismatches"], "checked": p.get("checked_rays")}
            elif "passed" in p:
                out[prefix] = {"passed": p["passed"], "checked": p.get("checked_pixels")}
        for k in ("identical_on_every_rank",):
            if k in d:
                out[prefix + "." + k] = d[k]
        hp = d.get("heightfield_properties")
        if isinstance(hp, dict) and "passed" in hp:
            out[prefix + ".heightfield"] = hp["passed"]
        for k, v in d.items():
            if k not in ("parity", "roofline", "e2e", "clocks", "world") and isinstance(v, dict):
                walk(f"{prefix}.{k}" if prefix else k, v)
    walk("headline", {"parity": extras.get("parity")})
    walk("", {k: v for k, v in extras.items() if k != "parity"})
    return out


def emit(line: dict):
    """The one JSON line goes to the real stdout; everything else printed during the run (NCCL's version banner, library
    chatter) was redirected to stderr by main()."""
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def guarded(env, line, extras, key, fn, seconds=240.0):
    """Runs an extra that holds collectives of its own under a watchdog: a hang in it (one rank failing inside a collective) costs the
    extra, never the line."""
    def bail():
        if env.rank == 0:
            extras[key] = {"error": f"watchdog: no result within {seconds:.0f} s"}
            line["parity_summary"] = parity_summary(extras)
            emit(line)
        os._exit(0)
    wd = threading.Timer(seconds, bail)
    wd.daemon = True
    wd.start()
    try:
        res = fn()
    except Exception as ex:
        res = {"error": repr(ex)}
    wd.cancel()
    extras[key] = res


def main():
    global _REAL_STDOUT
    args = parse_args()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
        return
    env = Env()
    rank, G = env.rank, env.world_size
    name = args.workload
    if name == "c5":   # the 8192^3 workload has no reference-format scene on the GPUs: it runs as the c5 leg alone
        res = extra_c5(env)
        if rank == 0:
            pr = res.get("primary", {})
            emit({"metric": "Mrays/s", "value": pr.get("value"), "unit": "Mrays/s", "n_gpus": G, "steps": 10, "warmup": 3, "ms_per_step": pr.get("ms_per_frame"),
                  "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                  "config": {"workload": "c5: " + res["workload"]}, "extras": {"c5": res}, "parity_summary": parity_summary({"c5": res})})
        return
    ctx = open_workload(env, name)
    wl, r = ctx["wl"], ctx["renderer"]
    W, H, kind = wl["width"], wl["height"], wl["kind"]
    do_cpu = rank == 0 and G == 1 and not args.no_cpu
    do_check = rank == 0 and not args.no_cpu    # the CPU checker as the parity check: every N (cpu_baseline itself stays an N = 1 figure)
    extras = {"world": {k: ctx["stats"][k] for k in ("nodes", "subChunks", "voxels", "bricks", "compactBytes", "levels", "buildMs", "uploadMs")},
              "scene_gen_s": ctx["gen_s"]}
    tiled = G > 1 and kind == "primary"
    e2e = None
    if kind == "path":
        m = measure_path(env, ctx, args.steps, args.warmup)
        rays_total, ms, e2e = m["rays_total"], m["ms"], m["e2e"]
        scaling = "strong" if G > 1 else "weak"
        par = (f"{G} GPU(s), scene replicated, the frame's {wl['spp']} spp sharded over the ranks, accumulation buffers reduced by "
               "blokb_reduce_accum (NCCL reduce-scatter + resolve + all-gather)")
    elif tiled:
        # the records that cross NVLink are the 8-byte packed ones (node | face << 29, bits of t: every hit ID and the distance; material and
        # gl_PrimitiveID are functions of the node): an all-to-all of 16-byte records is bound by NVLink ingress at 7/8 of the image per GPU
        # Same records at every N: the headline exchanges the full 16-byte blokb_hit records the N = 1 headline, the e2e figure and the CPU
        # arm produce; the 8-byte packed exchange (half the NVLink bytes) is extras.tiled_packed_records, to be set against
        # extras.packed_records of the N = 1 line.
        m = measure_tiled(env, ctx, args.steps, args.warmup, packed=False)
        m["rays_rank"] = W * H / G
        rays_total, ms = W * H, m["ms"]
        scaling = "strong"
        par = (f"ONE frame tiled over {G} GPUs in interleaved 8-row strips, scene replicated; exchange fused into the traversal kernel: {m['exchange']} "
               f"of the full 16-byte hit records (blokb_hit, as at N = 1), every rank ends with the whole image; frame fence = {m['fence']}")
        tiled_sample = m.pop("assembled_sample")
        extras["tiled"] = {"identical_on_every_rank": m["identical"], "share_only_ms": m["share_only_ms"], "trace_only_ms": m["trace_only_ms"],
                           "split": "share_only_ms = this rank's strips, stores local (launch + wave tail of 1/G of the frame); trace_only_ms = the same with the "
                                    "peer / multicast stores; ms_per_step = that plus the frame fence. All max over ranks, warm L2",
                           "exchange": m["exchange"], "fence": m["fence"],
                           "nvlink_ingress_bytes_per_gpu": m["nvlink_ingress_bytes_per_gpu"], "record_bytes": m["record_bytes"]}
        from blok_b200 import sharding
        r.set_strips(*sharding.strip_set(rank, G))
        e2e = e2e_pipeline(env, ctx, args.steps)
        r.set_strips(0, 1)
    else:
        m = measure_single(env, ctx, args.steps, args.warmup)
        rays_total, ms = env.sum(m["rays_rank"]), m["ms"]
        scaling = "weak"
        par = f"{G} GPU(s), scene replicated, one frame per GPU per step"
        if kind == "primary":
            e2e = e2e_pipeline(env, ctx, args.steps)
    value = rays_total / (ms * 1e-3) / 1e6
    if "ms_warm_l2" in m:
        extras["warm_l2"] = {"value": rays_total / (m["ms_warm_l2"] * 1e-3) / 1e6, "ms_per_step": m["ms_warm_l2"],
                             "what": "same frames back to back without the L2 flush (steady-state render loop)"}

    # ---- CPU leg (rank 0, N = 1): the checker timed on a strided sample of the same frame; doubles as the parity check
    cpu, roof = None, None
    res = None
    if do_check and not do_cpu:
        # N > 1: rank 0 checks what the exchange / the reduction delivered against the CPU checker (same strided sample as at N = 1)
        from blok_b200 import IMG_COLOR
        res = run_cpu(ctx["world"], wl, ctx["cam"])
        if tiled:
            extras["tiled"]["parity"] = hit_parity(tiled_sample, res)
            extras["tiled"]["parity"]["what"] = "the ASSEMBLED image on rank 0 (own strips + the peers' strips that arrived over NVLink): " + extras["tiled"]["parity"]["what"]
            extras["parity"] = extras["tiled"]["parity"]
        elif kind == "path":
            extras["parity"] = radiance_parity(r.readback(IMG_COLOR).reshape(-1, 4)[::res["stride"], :3], res)
            extras["parity"]["what"] = f"the REDUCED image of the {G} ranks' sample shards (blokb_reduce_accum over NCCL): " + extras["parity"]["what"]
    if kind == "path" and m.get("mc"):
        extras["multicast_reduce"] = mc_block(m, res if not do_cpu else None)
    if do_cpu:
        from blok_b200 import IMG_COLOR, IMG_HIT, IMG_SHADOW
        res = run_cpu(ctx["world"], wl, ctx["cam"])
        r.set_frame(ctx["ubo"])
        if kind == "path":
            extras["parity"] = radiance_parity(r.readback(IMG_COLOR).reshape(-1, 4)[::res["stride"], :3], res)
        else:
            r.trace_primary(shadow=(kind == "primary_shadow"))
            sh = r.readback(IMG_SHADOW).reshape(-1)[::CPU_SAMPLE_STRIDE] if kind == "primary_shadow" else None
            extras["parity"] = hit_parity(r.readback(IMG_HIT).reshape(-1)[::CPU_SAMPLE_STRIDE], res, sh)
        cpu = {"value": res["value"], "unit": "Mrays/s", "cores": res["cores"], "kind": res["kind"], "sample": res["sample"], "seconds": res["seconds"]}
        V, A, va_src = res["V"], res["A"], f"counted by the CPU checker ({res['kind']}) on the cpu_baseline sample in this run"
    else:
        cc = committed_counts(name)
        V, A, va_src = (cc["V"], cc["A"], "blok_b200/workload_counts.json (CPU checker counts of the same sample, committed)") if cc else (None, None, None)
    if V is not None:
        rpp = rays_total / (W * H) / (G if (kind != "path" and not tiled) else 1)
        roof = roofline_for(name, kind, m["rays_rank"], m["ms_rank"], V, A, rpp, va_src, m["clocks"], env.sms,
                            share=(1.0 / G if (tiled or kind == "path") else 1.0))

    line = None
    if rank == 0:
        line = {"metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": G, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms, "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"{name}: {wl['desc']}", "rays_per_step": rays_total, "rays_per_step_per_gpu": m["rays_rank"],
                           "l2": "flushed between timed steps (256 MiB memset outside the event pairs); per-step output 132.7 MB also exceeds the 126 MB L2",
                           "timing": "CUDA events on the launching stream around each step, summed; max over ranks", "parallelism": par},
                "e2e": e2e, "gpu_launches": m["launches"], "clocks": m["clocks"], "roofline": roof, "cpu_baseline": cpu, "extras": extras}
    else:
        line = {}

    if not args.no_extras:
        if rank == 0 and G == 1:
            for key, fn in (("device_build", extra_device_build), ("frame", extra_full_frame)):
                try:
                    extras[key] = fn(env, ctx)
                except Exception as ex:  # an extra must never cost the headline line
                    extras[key] = {"error": repr(ex)}
            if kind == "primary":
                try:   # the kernel the N > 1 headline runs on each rank (8-byte packed records), whole frame on one GPU
                    _, pk_ms, _ = env.timed(lambda: r.trace_primary(packed=True, stream=env.sp), max(10, args.steps // 2), 3)
                    extras["packed_records"] = {"ms_per_step": pk_ms, "value": W * H / pk_ms / 1e3, "unit": "Mrays/s",
                                                "what": "k_primary_packed: same traversal, 8-byte records (node | face << 29, t) — the form a tiled frame exchanges at N > 1"}
                except Exception as ex:
                    extras["packed_records"] = {"error": repr(ex)}
                try:
                    extras["e2e_packed"] = e2e_pipeline(env, ctx, args.steps, packed=True)
                    extras["e2e_packed"]["what"] += ("; 8-B records (node | face << 29, t): material and gl_PrimitiveID are looked up on the host from "
                                                     "the uploaded arrays (blokb_unpack_hits), outside the timed region — not a like-for-like e2e")
                except Exception as ex:
                    extras["e2e_packed"] = {"error": repr(ex)}
        if tiled:
            def packed_leg():
                mp = measure_tiled(env, ctx, max(10, args.steps // 2), 3, packed=True)
                out = {"what": "the same tiled frame with 8-byte packed records (node | face << 29, t; material and gl_PrimitiveID re-derived on the host by "
                               "blokb_unpack_hits) through the exchange: half the NVLink bytes. N = 1 denominator: extras.packed_records of the N = 1 line",
                       "ms_per_frame": mp["ms"], "value": W * H / mp["ms"] / 1e3, "unit": "Mrays/s", "identical_on_every_rank": mp["identical"],
                       "share_only_ms": mp["share_only_ms"], "trace_only_ms": mp["trace_only_ms"], "exchange": mp["exchange"],
                       "nvlink_ingress_bytes_per_gpu": mp["nvlink_ingress_bytes_per_gpu"]}
                if rank == 0 and res is not None:
                    out["parity"] = hit_parity(ctx["world"].unpack_hits(mp["assembled_sample"]), res)
                return out
            guarded(env, line, extras, "tiled_packed_records", packed_leg)

            def replicas_leg():
                mr = measure_single(env, ctx, 20, 3, cam_rank=rank)
                tot = env.sum(mr["rays_rank"])
                return {"what": "every rank renders its own camera's whole frame, no exchange (the frames in flight of a multi-GPU box): weak scaling",
                        "value": tot / (mr["ms"] * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": mr["ms"], "scaling": "weak"}
            guarded(env, line, extras, "replicas", replicas_leg)
        r.close()
        ctx = None
        for other, st, wu in (("c2", 20, 3), ("c4", 2, 1)):
            if other == name:
                continue
            guarded(env, line, extras, other, lambda o=other, s=st, w=wu: extra_workload(env, o, s, w, do_cpu, do_check), seconds=420.0)
        # the 8K frame at every N (a 1 -> 8 curve on a frame big enough to scale); the CPU-checker leg only at N = 8
        guarded(env, line, extras, "c5", lambda: extra_c5(env, check=(G == 8)), seconds=420.0)
    elif ctx is not None:
        r.close()
    if rank == 0:
        line["parity_summary"] = parity_summary(extras)   # LAST key: the driver's record keeps the tail of the line
        emit(line)
    if G > 1:
        env.dist.barrier()
        env.dist.destroy_process_group()


if __name__ == "__main__":
    main()
