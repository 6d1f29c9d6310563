"""Host side of the readback: what bounds the aggregate device->host rate when several GPUs of one box deliver frames at once?
One process per GPU (torchrun); each copies a 132.7 MB device buffer into pinned host memory `iters` times, alone and all together,
with the pinned buffer (a) wherever cudaHostAlloc puts it and (b) bound to the NUMA node of the GPU (numactl-style mbind through
libnuma when present). Prints per-rank and aggregate GB/s plus the topology the box reports.
  python -m torch.distributed.run --nproc-per-node N tools/d2h_probe.py"""
import ctypes
import os
import subprocess
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
NB = 132710400
dev = torch.empty(NB, dtype=torch.uint8, device="cuda")


def gpu_numa_node():
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        bus = pynvml.nvmlDeviceGetPciInfo(h).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        p = f"/sys/bus/pci/devices/{bus.lower()[4:] if len(bus) > 12 else bus.lower()}/numa_node"
        return int(open(p).read())
    except Exception as ex:
        return f"? ({ex!r})"


def rate(host, iters=20, together=True):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        host.copy_(dev, non_blocking=True)
    s.synchronize()
    if together and world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    with torch.cuda.stream(s):
        for _ in range(iters):
            host.copy_(dev, non_blocking=True)
    s.synchronize()
    dt = time.perf_counter() - t0
    return NB * iters / dt / 1e9


def allsum(x):
    if world == 1:
        return x
    t = torch.tensor([x], device="cuda", dtype=torch.float64)
    dist.all_reduce(t)
    return float(t.item())


node = gpu_numa_node()
host = torch.empty(NB, dtype=torch.uint8, pin_memory=True)
alone = None
for r in range(world):          # one rank at a time
    if r == rank:
        alone = rate(host, together=False)
    if world > 1:
        dist.barrier()
both = rate(host)
agg = allsum(both)
res = {"rank": rank, "gpu_numa_node": node, "alone_GBs": round(alone, 2), "together_GBs": round(both, 2)}
# NUMA-bound pinned buffer: allocate ordinary memory bound to the GPU's node (libnuma), then cudaHostRegister it
bound = None
try:
    numa = ctypes.CDLL("libnuma.so.1")
    numa.numa_alloc_onnode.restype = ctypes.c_void_p
    numa.numa_alloc_onnode.argtypes = [ctypes.c_size_t, ctypes.c_int]
    if isinstance(node, int) and node >= 0:
        p = numa.numa_alloc_onnode(NB, node)
        ctypes.memset(p, 0, NB)
        cudart = ctypes.CDLL("libcudart.so.12")
        rc = cudart.cudaHostRegister(ctypes.c_void_p(p), ctypes.c_size_t(NB), 0)
        if rc == 0:
            class V:
                __array_interface__ = {"shape": (NB,), "typestr": "|u1", "data": (p, False), "version": 3}
            import numpy as np
            arr = np.asarray(V())
            hostb = torch.from_numpy(arr)
            bound = rate(hostb)
            res["together_numa_bound_GBs"] = round(bound, 2)
        else:
            res["numa_bound"] = f"cudaHostRegister rc={rc}"
    else:
        res["numa_bound"] = "gpu numa node unknown"
except Exception as ex:
    res["numa_bound"] = repr(ex)
aggb = allsum(bound or 0.0)
out = [None] * world
if world > 1:
    dist.all_gather_object(out, res)
else:
    out = [res]
if rank == 0:
    print({"aggregate_together_GBs": round(agg, 2), "aggregate_numa_bound_GBs": round(aggb, 2), "ranks": out})
    for cmd in (["nvidia-smi", "topo", "-m"], ["lscpu"], ["numactl", "-H"]):
        try:
            print("$", " ".join(cmd)); print(subprocess.run(cmd, capture_output=True, text=True, timeout=30).stdout[:6000])
        except Exception as ex:
            print(cmd, "failed:", ex)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
