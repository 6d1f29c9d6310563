"""Probe: does this box give torch symmetric memory an NVLink multicast address? (run under torchrun, >= 2 ranks)
Prints every rank's peer buffer pointers and multicast pointer; sharding.connect_multicast relies on the latter."""
import os, torch, torch.distributed as dist
rank=int(os.environ["RANK"]); local=int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev=torch.device("cuda",local)
dist.init_process_group("nccl", device_id=dev)
import torch.distributed._symmetric_memory as symm_mem
try:
    t = symm_mem.empty(1<<20, dtype=torch.int32, device=dev)
    h = symm_mem.rendezvous(t, dist.group.WORLD.group_name if hasattr(dist.group.WORLD,'group_name') else dist.group.WORLD)
    print(rank, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "mc", hex(h.multicast_ptr), "world", h.world_size, flush=True)
    print([a for a in dir(h) if not a.startswith('_')], flush=True)
except Exception as e:
    import traceback; traceback.print_exc()
dist.barrier(); dist.destroy_process_group()
