"""CPU, world_size 2 over gloo: the spp-sharded path (each rank renders its sample range through the host emulation of
the device functions, then the accumulation buffers are reduced) equals the single-rank frame."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, gold, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from blok_b200 import WorldSvoGpu, sharding
    from tests.host_emul import emu
    z = np.load(gold)
    w = WorldSvoGpu(z["nodes"], z["subs"], z["mats"])
    e = emu.load().world(w)
    spp = 6
    s0, s1 = sharding.sample_range(rank, world, spp)
    part = torch.from_numpy(e.path(z["ubo_r"], s0, s1, threads=2))
    for mode in ("allreduce", "reduce_scatter"):
        acc = part.clone()
        if mode == "allreduce":
            color = sharding.resolve(sharding.allreduce_accum(acc))
        else:
            color = sharding.reduce_scatter_resolve(acc)
        np.save(os.path.join(out_dir, f"{mode}_{rank}.npy"), color.numpy())
    if rank == 0:
        full = torch.from_numpy(e.path(z["ubo_r"], 0, spp, threads=2))
        np.save(os.path.join(out_dir, "single.npy"), sharding.resolve(full).numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_spp_sharding_world2_equals_single(tmp_path):
    gold = os.path.join(ROOT, "tests", "golden", "castle.npz")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, gold, str(tmp_path)), nprocs=2, join=True)
    single = np.load(tmp_path / "single.npy")
    for mode in ("allreduce", "reduce_scatter"):
        a, b = np.load(tmp_path / f"{mode}_0.npy"), np.load(tmp_path / f"{mode}_1.npy")
        assert np.array_equal(a, b)  # every rank ends with the same frame
        np.testing.assert_allclose(a, single, rtol=1e-5, atol=1e-6)  # fp32 summation order only
