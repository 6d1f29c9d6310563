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


def _strip_worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from blok_b200 import sharding
    for h, w in ((64, 5), (72, 3), (13, 4)):  # strips divisible by world, odd strip count, ragged last strip
        full = torch.arange(h * w * 4, dtype=torch.int32).reshape(h, w, 4)
        owner = (torch.arange(h) // sharding.STRIP_ROWS) % world
        first, stride = sharding.strip_set(rank, world)
        assert (first, stride) == (rank, world)
        mine = torch.full_like(full, -1)           # what this rank's kernels wrote: its strips only, at frame positions
        mine[owner == rank] = full[owner == rank]
        assert sharding.strips_owned(rank, world, h) == len(torch.unique(torch.arange(h)[owner == rank] // sharding.STRIP_ROWS))
        got = sharding.gather_strips(mine)
        np.save(os.path.join(out_dir, f"strips_{h}_{rank}.npy"), got.numpy())
        if rank == 0:
            np.save(os.path.join(out_dir, f"strips_{h}_full.npy"), full.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_gather_strips_world2(tmp_path):
    port = 31500 + (os.getpid() % 2000)
    mp.spawn(_strip_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    for h in (64, 72, 13):
        full = np.load(tmp_path / f"strips_{h}_full.npy")
        for rank in range(2):
            assert np.array_equal(np.load(tmp_path / f"strips_{h}_{rank}.npy"), full)


def test_strip_sets_partition_the_frame():
    from blok_b200 import sharding
    for h in (1, 8, 9, 2160, 4320):
        strips = (h + 7) // 8
        for G in (1, 2, 3, 4, 8):
            assert sum(sharding.strips_owned(g, G, h) for g in range(G)) == strips
