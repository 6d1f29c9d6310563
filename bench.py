#!/usr/bin/env python
"""bench.py — Mrays/s of the voxel ray-traversal hot path on N B200s (contract: see the task statement / DESIGN.md §6).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c3|c2|c4] [--impl ours|reference] [--no-extras]

A step = one frame of the workload (default c3: 2048^3 SVO terrain, 3840x2160 primary rays = 8 294 400 rays per GPU).
  value     rays/s with the scene and frame constants already resident in HBM; CUDA events on the launching stream
            around every step, L2 flushed (256 MiB memset) between steps outside the event pairs; max over ranks.
  e2e       the same frames through the public API with HOST buffers: per step the host fills a FrameUBO (560 B, pinned),
            blokb_set_frame + blokb_trace_primary, then the hit-record image is read back into pinned host memory.
  roofline  algorithmic bytes per ray B = 16 V + 48 A + O; V, A counted by the CPU oracle on the
            cpu_baseline sample of the same frame; achieved = rays * B / kernel time; peak = MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline / --impl reference   the reference's own sources compiled in place (oracle/_ref, kind "reference") or, when
            that library is absent, the CPU restatement (kind "port"), all host threads, on a strided pixel sample.
N > 1 (torchrun): one rank per GPU, scene replicated, each rank renders its own camera's frame per step (weak scaling,
no data-path collective); extras.c4 shards 64 spp across ranks and all-reduces the accumulation buffer (NCCL).
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


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=100)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--workload", default="c3", choices=["c2", "c3", "c4", "c5", "t1"])
    p.add_argument("--no-extras", action="store_true", help="skip the secondary workloads (c2, c4) and warm-L2 numbers")
    p.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (then roofline uses cached V, A)")
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


def cpu_checker():
    """(library, kind): the reference's own sources compiled in place when present, else the restatement."""
    from oracle import pyoracle
    if os.path.exists(pyoracle.REF_SO):
        return pyoracle.RefLib(), "reference"
    if not os.path.exists(pyoracle.ORACLE_SO):
        pyoracle.build(ref=False)
    return pyoracle.OracleLib(), "port"


def cpu_sample_rays(lib, world, wl, cam, kind):
    """The bounded CPU sample: strided pixels of the workload's frame (primary rays; for primary_shadow the oracle also
    traces the sun shadow ray of every lit hit, for path it runs the full raygen on the strided pixels)."""
    from oracle import pyoracle
    sc = pyoracle.Scene(world.globalNodes, world.globalSubChunks, world.materials)
    ubo = cam.frame(wl["width"], wl["height"], 0, wl["spp"]).to_numpy()
    return sc, ubo


def run_cpu(world, wl, cam, threads=None):
    """Times the CPU checker on the strided sample. Returns dict(value Mrays/s, rays, seconds, V, A, kind, cores, sample)."""
    lib, kind = cpu_checker()
    sc, ubo = cpu_sample_rays(lib, world, wl, cam, wl["kind"])
    threads = threads or os.cpu_count() or 1
    W, H = wl["width"], wl["height"]
    if wl["kind"] == "path":
        stride = CPU_SAMPLE_STRIDE * 64
        t0 = time.perf_counter()
        img = lib.render(sc, ubo, 0, W * H, stride, threads=threads, want=("color",))
        dt = time.perf_counter() - t0
        cnt = img["counters"]
        npx = (W * H + stride - 1) // stride
        rays = None  # the oracle counts node fetches / AABB passes, not rays: filled in by the caller from the GPU's count
        return dict(seconds=dt, pixels=npx, fetches=int(cnt[0]), aabbs=int(cnt[1]), kind=kind, cores=threads, rays=rays,
                    sample=f"every {stride}th pixel of the {W}x{H} frame, {wl['spp']} spp full raygen ({npx} pixels)")
    rays = lib.primary_rays(ubo)[::CPU_SAMPLE_STRIDE]
    t0 = time.perf_counter()
    h, t, cnt = lib.trace(sc, rays, threads=threads)
    n = len(rays)
    fetch, aabb = int(cnt[0]), int(cnt[1])
    if wl["kind"] == "primary_shadow":
        srays = shadow_rays(rays, h, t)
        _, _, c2 = lib.trace(sc, srays, any_hit=True, threads=threads)
        n += len(srays)
        fetch += int(c2[0]); aabb += int(c2[1])
    dt = time.perf_counter() - t0
    return dict(seconds=dt, rays=n, value=n / dt / 1e6, V=fetch / n, A=aabb / n, kind=kind, cores=threads,
                hit_fraction=float((t >= 0).mean()), hits=(h, t),
                sample=f"every {CPU_SAMPLE_STRIDE}th pixel (row-major) of the {W}x{H} frame: {n} rays")


def shadow_rays(rays, h, t):
    """Sun shadow rays of the lit primary hits, built like raygen.rgen:280-298 (fp32, unfused)."""
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
    return np.concatenate([o[cast], np.tile(sun, (n, 1)), np.full((n, 1), 0.001, np.float32), np.full((n, 1), 1000, np.float32)], 1).astype(np.float32)


COUNTS_PATH = os.path.join(ROOT, "blok_b200", "workload_counts.json")


def committed_counts(name):
    """V, A of the workload's strided sample as counted by the CPU checker in the build container (used only when the
    cpu_baseline leg does not run: N > 1 or --no-cpu). Regenerate with `python bench.py --impl reference --write-counts`."""
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


# ------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
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
        rays += res["rays"] or 0
        dt += res["seconds"]
    if wl["kind"] == "path":
        # the oracle does not count rays; 1 primary + <=1 shadow + <=1 bounce per sample: use the committed GPU count
        cc = committed_counts(name) or {}
        rpp = cc.get("rays_per_pixel", 2.5 * wl["spp"])
        rays = res["pixels"] * rpp * args.steps
    value = rays / dt / 1e6
    line = {"impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{name}: {wl['desc']}", "step": "one bounded CPU sample of the frame: " + res["sample"]},
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": res["cores"], "kind": res["kind"], "sample": res["sample"]},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if "V" in res:
        line["counts"] = {"V": res["V"], "A": res["A"], "hit_fraction": res["hit_fraction"]}
    if getattr(args, "write_counts", False) and "V" in res:
        allc = {}
        if os.path.exists(COUNTS_PATH):
            with open(COUNTS_PATH) as f:
                allc = json.load(f)
        allc.setdefault(name, {}).update({"V": res["V"], "A": res["A"], "hit_fraction": res["hit_fraction"], "kind": res["kind"],
                                          "sample": res["sample"]})
        with open(COUNTS_PATH, "w") as f:
            json.dump(allc, f, indent=1, sort_keys=True)
    emit(line)


# ------------------------------------------------------------------------------------------ our arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    import blok_b200
    from blok_b200 import IMG_ACCUM, IMG_HIT, IMG_SHADOW, sharding, workloads

    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world_size == 1:
            return float(x)
        t = torch.tensor([float(x)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world_size == 1:
            return float(x)
        t = torch.tensor([float(x)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    stream = torch.cuda.Stream(device=dev)
    sp = stream.cuda_stream
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def flush_l2():
        with torch.cuda.stream(stream):
            flush_buf.zero_()

    def measure(name, steps, warmup, flush=True, e2e=True):
        """Device-timed and end-to-end numbers of one workload on this rank's GPU."""
        wl = workloads.WORKLOADS[name]
        kind, W, H, spp = wl["kind"], wl["width"], wl["height"], wl["spp"]
        t0 = time.perf_counter()
        world = workloads.build_world(name)
        t_gen = time.perf_counter() - t0
        r = blok_b200.Renderer(W, H, local)
        r.add_world(world)
        stats = r.world_stats()
        # weak-scaling kinds render one frame per GPU (own camera); the spp-sharded path frame is ONE frame for all ranks
        cam = workloads.camera_for(name, 0 if kind == "path" else rank)
        ubo = cam.frame(W, H, 0, spp)
        r.set_frame(ubo)
        s0, s1 = sharding.sample_range(rank, world_size, spp) if kind == "path" else (0, 1)
        accum_t = torch.as_tensor(r.device_array(IMG_ACCUM), device=dev) if kind == "path" else None

        def step():
            if kind == "primary":
                r.trace_primary(stream=sp)
            elif kind == "primary_shadow":
                r.trace_primary(shadow=True, stream=sp)
            else:
                r.reset_accum(stream=sp)
                r.trace_path(s0, s1, stream=sp)
                if world_size > 1:
                    with torch.cuda.stream(stream):
                        dist.all_reduce(accum_t, op=dist.ReduceOp.SUM)
                r.resolve(stream=sp)

        # rays per step on this rank (deterministic): primary rays + the shadow / bounce rays the kernels actually cast
        step()
        torch.cuda.synchronize()
        rays = W * H
        if kind == "primary_shadow":
            rays += int(np.count_nonzero(r.readback(IMG_SHADOW)))
        elif kind == "path":
            r.stats_enable(True)
            r.reset_accum(stream=sp)
            r.trace_path(s0, s1, stream=sp)
            torch.cuda.synchronize()
            rays = r.stats_read()[1]
            r.stats_enable(False)
        for _ in range(warmup):
            step()
        barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        l0 = r.launch_count()
        with ClockSampler(local) as clk:
            barrier()
            for a, b in ev:
                if flush:
                    flush_l2()
                a.record(stream)
                step()
                b.record(stream)
            barrier()
        launches = r.launch_count() - l0
        ms = sum(a.elapsed_time(b) for a, b in ev)
        ms_step = max_over_ranks(ms / steps)
        total_rays = sum_over_ranks(rays)
        out = dict(name=name, wl=wl, world=world, renderer=r, cam=cam, stats=stats, rays_rank=rays, rays_total=total_rays,
                   ms_per_step=ms_step, ms_per_step_rank=ms / steps, value=total_rays / (ms_step * 1e-3) / 1e6, launches=launches, clocks=clk.summary(),
                   gen_s=t_gen, ubo=ubo)
        if not flush:
            return out
        # warm-L2 variant: back-to-back frames, one event pair (what a render loop sees in steady state)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        a.record(stream)
        for _ in range(steps):
            step()
        b.record(stream)
        barrier()
        out["ms_per_step_warm_l2"] = max_over_ranks(a.elapsed_time(b) / steps)
        out["value_warm_l2"] = total_rays / (out["ms_per_step_warm_l2"] * 1e-3) / 1e6
        if e2e and kind == "primary":
            # frames in flight through the public API: per step the host builds the FrameUBO from the Camera, submits the
            # frame (trace on the context's stream, hit records copied to pinned host memory on the copy stream) and reads
            # the previous frame's records; the copy of frame i overlaps the tracing of frame i+1 (depth 2, like the
            # reference's 2 frames in flight, blok/include/renderer.hpp:185)
            def pipelined(packed):
                rec = 8 if packed else 16
                pipe = blok_b200.FramePipeline(r, W, H, depth=2, packed=packed)
                key = "nodeFace" if packed else "node"
                e2e_steps = max(3, min(steps, 50))
                for i in range(2):
                    pipe.result(pipe.submit(cam.frame(W, H, i, spp)))
                pipe.drain()
                barrier()
                t0 = time.perf_counter()
                last, seen = None, 0
                for i in range(e2e_steps):
                    tk = pipe.submit(cam.frame(W, H, i, spp))
                    if last is not None:
                        seen += int(pipe.result(last)[key][H // 2, W // 2] != 0xFFFFFFFF)
                    last = tk
                seen += int(pipe.result(last)[key][H // 2, W // 2] != 0xFFFFFFFF)
                pipe.drain()
                barrier()
                dt = max_over_ranks(time.perf_counter() - t0)
                pipe.close()
                return {"value": total_rays * e2e_steps / dt / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 560,
                        "d2h_bytes_per_step": W * H * rec, "steps": e2e_steps, "ms_per_step": dt / e2e_steps * 1e3}

            out["e2e"] = pipelined(False)
            out["e2e"]["what"] = ("FramePipeline (blokb_pipeline_*, 2 frames in flight): Camera -> FrameUBO on host, submit = primary pass + "
                                  "async copy of the 16-B hit records to pinned host memory, result(previous frame) read on the host; "
                                  "scene uploaded once beforehand (Renderer::addWorld)")
            try:  # an extra: must never cost the headline line
                out["e2e_packed"] = pipelined(True)
                out["e2e_packed"]["what"] = ("the same loop with blokb_pipeline_*_packed: 8-B records (node | face << 29, t); material and "
                                             "gl_PrimitiveID are looked up on the host from the uploaded arrays (blokb_unpack_hits), "
                                             "outside the timed region")
            except Exception as ex:
                out["e2e_packed"] = {"error": repr(ex)}
            r.set_frame(ubo)
        elif e2e:
            nbytes = W * H * 16
            host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
            e2e_steps = 2
            barrier()
            t0 = time.perf_counter()
            for i in range(e2e_steps):
                r.set_frame(ubo)
                step()
                stream.synchronize()
                r.readback_into(blok_b200.IMG_COLOR, host.data_ptr(), nbytes)
            barrier()
            dt = max_over_ranks(time.perf_counter() - t0)
            out["e2e"] = {"value": total_rays * e2e_steps / dt / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 560,
                          "d2h_bytes_per_step": nbytes, "steps": e2e_steps, "ms_per_step": dt / e2e_steps * 1e3,
                          "what": "blokb_set_frame, reset_accum, trace_path(sample shard), all-reduce, resolve, readback of RGBA32F colour"}
        return out

    return measure, dict(rank=rank, world_size=world_size, local=local, dev=dev, barrier=barrier, torch=torch, dist=dist)


def extra_device_build(name, r, env):
    """Scene upload on the device (blokb_upload_voxels) beside the host path the headline used, and the
    full-size parity property: both leave byte-identical compact arrays (incl. the reference node index of every leaf block)."""
    import hashlib
    import blok_b200
    from blok_b200 import WORLD_BLOCKFIRSTLEAF, WORLD_NODES, WORLD_VOXELMAT, workloads
    xyz, mid, mats = workloads.voxels(name)
    b = blok_b200.Renderer(64, 64, env["local"])
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


def extra_tiled_frame(name, r, env, steps=30, warmup=3):
    """N > 1: ONE frame of the workload tiled over the GPUs (strong scaling; SURVEY.md 8(e) pixel partition). Rank g renders
    the 8-row strips g, g + G, ... (blokb_set_strips) and the kernel stores every hit record into all ranks' images — through
    an NVLink multicast address when the box supports it (blokb_bind_image), else P2P stores to the peers' IPC-mapped
    images (blokb_set_peers); a 4-byte all-reduce on the dispatch stream is the frame fence. Parity: the assembled image on
    every rank is byte-identical to the rank's own undivided dispatch."""
    import blok_b200
    from blok_b200 import IMG_HIT, sharding, workloads
    torch, dist = env["torch"], env["dist"]
    rank, G, dev = env["rank"], env["world_size"], env["dev"]
    wl = workloads.WORKLOADS[name]
    W, H = wl["width"], wl["height"]
    r.set_strips(0, 1)
    r.set_frame(workloads.camera_for(name, 0).frame(W, H, 0, 1))
    r.trace_primary()
    full = r.readback(IMG_HIT).copy()
    keep = sharding.connect_multicast(r, (IMG_HIT,))
    if keep is None:
        sharding.connect_peers(r, (IMG_HIT,))
    r.set_strips(*sharding.strip_set(rank, G))
    stream = torch.cuda.Stream(device=dev)
    token = torch.zeros(1, device=dev)

    def step():
        r.trace_primary(stream=stream.cuda_stream)
        with torch.cuda.stream(stream):
            sharding.frame_barrier(token)

    for _ in range(warmup):
        step()
    env["barrier"]()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        a.record(stream)
        step()
        b.record(stream)
    env["barrier"]()
    ms = sum(a.elapsed_time(b) for a, b in ev) / steps
    t = torch.tensor([ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    same = torch.tensor([int(r.readback(IMG_HIT).tobytes() == full.tobytes())], device=dev)
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    if keep is None:
        sharding.disconnect_peers(r, (IMG_HIT,))
    else:
        sharding.disconnect_multicast(r, keep)
    r.set_strips(0, 1)
    return {"what": f"one {W}x{H} frame of {name} tiled over {G} GPUs in interleaved 8-row strips, exchange fused into the traversal kernel",
            "exchange": "NVLink multicast (multimem.st into symmetric memory)" if keep is not None else "P2P stores into IPC-mapped peer images",
            "ms_per_frame": ms, "value": W * H / ms / 1e3, "unit": "Mrays/s", "scaling": "strong", "steps": steps,
            "identical_on_every_rank": bool(int(same.item())),
            "nvlink_ingress_bytes_per_gpu": int(W * H * 16 * (G - 1) / G)}


def extra_full_frame(name, r, env):
    """The reference's whole frame (Renderer::render, blok/src/renderer_draw.cpp): raygen at the reference's 8 spp, Denoiser::denoise,
    PostProcess::process, over a slow dolly so that both temporal passes are live. Per stage, CUDA events on the launching stream."""
    import blok_b200
    from blok_b200 import Camera, PostSettings, workloads
    torch = env["torch"]
    wl = workloads.WORKLOADS[name]
    W, H = wl["width"], wl["height"]
    base = workloads.camera_for(name)
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
    out["what"] = f"{W}x{H}, 8 spp raygen + temporal/variance + 4 a-trous + TAA/tonemap + sharpen (reference defaults)"
    return out


_REAL_STDOUT = None


def emit(line: dict):
    """The one JSON line goes to the real stdout; everything else printed during the run (NCCL's version banner, library
    chatter) was redirected to stderr by main()."""
    data = (json.dumps(line) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, data)


def main():
    global _REAL_STDOUT
    args = parse_args()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
        return
    measure, env = run_ours(args)
    rank, world_size = env["rank"], env["world_size"]
    name = args.workload
    m = measure(name, args.steps, args.warmup)
    wl, r = m["wl"], m["renderer"]
    peak, peak_src = measured_peaks()

    # ---- CPU leg (rank 0, N = 1): the oracle timed on a strided sample of the same frame; doubles as a parity check
    cpu = None
    parity = None
    if rank == 0 and world_size == 1 and not args.no_cpu:
        res = run_cpu(m["world"], wl, m["cam"])
        if wl["kind"] == "path":
            cc = committed_counts(name) or {}
            rpp = m["rays_rank"] / (wl["width"] * wl["height"])
            res["rays"] = res["pixels"] * rpp
            res["value"] = res["rays"] / res["seconds"] / 1e6
            res["V"], res["A"] = res["fetches"] / res["rays"], res["aabbs"] / res["rays"]
        else:
            import blok_b200
            r.set_frame(m["ubo"])
            if wl["kind"] == "primary":
                r.trace_primary()
            else:
                r.trace_primary(shadow=True)
            got = r.readback(blok_b200.IMG_HIT).reshape(-1)[::CPU_SAMPLE_STRIDE]
            h, t = res["hits"]
            bad = int(np.count_nonzero((got["node"] != h[:, 1]) | (got["material"] != h[:, 2]) |
                                       (got["t"].view(np.uint32) != t.view(np.uint32))))
            parity = {"checked_rays": int(len(t)), "mismatches": bad, "what": "GPU hit records (node, material, bits of t) vs the CPU "
                      f"checker ({res['kind']}) on the cpu_baseline sample"}
        cpu = {"value": res["value"], "unit": "Mrays/s", "cores": res["cores"], "kind": res["kind"], "sample": res["sample"],
               "seconds": res["seconds"]}
        V, A, va_src = res["V"], res["A"], f"counted by the CPU checker ({res['kind']}) on the cpu_baseline sample in this run"
    else:
        cc = committed_counts(name)
        if cc:
            V, A, va_src = cc["V"], cc["A"], "blok_b200/workload_counts.json (CPU checker counts of the same sample, committed)"
        else:
            V = A = None
            va_src = "unavailable (no cpu leg and no committed counts)"

    rays_per_pixel = m["rays_rank"] / (wl["width"] * wl["height"]) * (world_size if wl["kind"] == "path" else 1)
    roof = None
    if V is not None:
        B = 16.0 * V + 48.0 * A + out_bytes_per_ray(wl["kind"], rays_per_pixel)
        achieved = m["rays_rank"] * B / (m["ms_per_step_rank"] * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "bytes_per_ray": B, "V": V, "A": A, "counts_source": va_src, "peak_source": peak_src,
                "kernel": {"primary": "k_primary<false>", "primary_shadow": "k_primary<true>", "path": "k_path"}[wl["kind"]],
                "note": "algorithmic bytes are defined in the reference's 16-B-node / 48-B-sub-chunk format (SURVEY.md §8(d)); the kernel "
                        "reads a compact L2-resident layout instead, so DRAM traffic (profiles/) is far below this figure"}
        tr = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tr):
            with open(tr) as f:
                t = json.load(f).get(name)
            if t:  # T = dram__bytes_read.sum + dram__bytes_write.sum of one launch (ncu --set full); the split and its source beside it
                roof["traffic"] = t["total"]
                roof["traffic_detail"] = t
                if t.get("warp_instructions") and (m.get("clocks") or {}).get("sm_mhz"):
                    # the bound this kernel actually sits under: warp instructions of one launch (smsp__inst_executed.sum of
                    # the same ncu capture; deterministic for this scene and camera) / the live kernel time, against 4 issue slots per SM
                    # per clock at the SM clock sampled during the timed region
                    import torch
                    sms = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
                    ach = t["warp_instructions"] / (m["ms_per_step_rank"] * 1e-3) / 1e9
                    pk = sms * 4 * m["clocks"]["sm_mhz"] * 1e6 / 1e9
                    roof["issue"] = {"achieved": ach, "peak": pk, "unit": "G warp-instr/s", "frac": ach / pk,
                                     "threads_per_instruction": t.get("threads_per_instruction")}

    extras = {"warm_l2": {"value": m["value_warm_l2"], "ms_per_step": m["ms_per_step_warm_l2"],
                          "what": "same frames back to back without the L2 flush (steady-state render loop)"},
              "world": {k: m["stats"][k] for k in ("nodes", "subChunks", "voxels", "bricks", "compactBytes", "levels", "buildMs", "uploadMs")},
              "scene_gen_s": m["gen_s"]}
    if parity:
        extras["parity"] = parity
    e2e = m.get("e2e")
    if "e2e_packed" in m:
        extras["e2e_packed"] = m["e2e_packed"]
    if rank == 0 and world_size == 1 and not args.no_extras:
        for key, fn in (("device_build", extra_device_build), ("frame", extra_full_frame)):
            try:
                extras[key] = fn(name, r, env)
            except Exception as ex:  # an extra must never cost the headline line
                extras[key] = {"error": repr(ex)}
    if not args.no_extras:
        for other, st, wu in (("c2", 20, 3), ("c4", 2, 1)):
            if other == name:
                continue
            try:
                o = measure(other, st, wu, flush=True, e2e=(other == "c4"))
                extras[other] = {"workload": o["wl"]["desc"], "value": o["value"], "unit": "Mrays/s", "ms_per_step": o["ms_per_step"],
                                 "rays_per_step": o["rays_total"], "value_warm_l2": o.get("value_warm_l2"),
                                 "scaling": "strong (spp sharded, accum all-reduced)" if other == "c4" else "weak"}
                if "e2e" in o:
                    extras[other]["e2e"] = o["e2e"]
                o["renderer"].close()
            except Exception as ex:  # an extra must never cost the headline line
                extras[other] = {"error": repr(ex)}
    if rank == 0:
        line = {"metric": "Mrays/s", "value": m["value"], "unit": "Mrays/s", "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic",
                "config": {"workload": f"{name}: {wl['desc']}", "rays_per_step_per_gpu": m["rays_rank"], "l2": "flushed between timed steps "
                           "(256 MiB memset outside the event pairs); per-step output 132.7 MB also exceeds the 126 MB L2",
                           "timing": "CUDA events on the launching stream around each step, summed; max over ranks",
                           "parallelism": f"{world_size} GPU(s), scene replicated, one frame per GPU per step"},
                "e2e": e2e, "gpu_launches": m["launches"], "clocks": m["clocks"], "roofline": roof, "cpu_baseline": cpu, "extras": extras}
    if world_size > 1 and not args.no_extras:
        # last, with everything else already in `line`: the one extra that adds collectives of its own. A watchdog makes sure a
        # hang in it (one rank failing inside a collective) costs the extra, never the line.
        import threading

        def bail():
            if rank == 0:
                extras["tiled_frame"] = {"error": "watchdog: no result within 180 s"}
                emit(line)
            os._exit(0)
        wd = threading.Timer(180.0, bail)
        wd.daemon = True
        wd.start()
        try:
            res = extra_tiled_frame(name, r, env)
        except Exception as ex:
            res = {"error": repr(ex)}
        wd.cancel()
        extras["tiled_frame"] = res
    r.close()
    if rank == 0:
        emit(line)
    if world_size > 1:
        env["dist"].barrier()
        env["dist"].destroy_process_group()


if __name__ == "__main__":
    main()
