"""The cross-GPU reduction entry of the C-ABI (blokb_comm_unique_id / blokb_comm_init / blokb_reduce_accum / blokb_reduce_slice /
blokb_readback_slice): sample-sharded path tracing without Python in the frame loop. One GPU: the reduction is the identity
(= blokb_resolve). Two GPUs (run with `gpurun --gpus 2`; skipped otherwise): NCCL reduce-scatter + resolve (+ all-gather), and — where
the box has NVLink multicast — the multimem.ld_reduce kernel over symmetric memory, both against the undivided single-GPU render;
also the packed 8-byte records through the tiled-frame exchange."""
import os

import numpy as np
import pytest

import blok_b200
from blok_b200 import (IMG_ACCUM, IMG_COLOR, IMG_HIT, IMG_HIT_PACKED, REDUCE_ALL, REDUCE_SLICE, WorldSvoGpu, sharding)

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "castle.npz")
SPP = 6


def test_reduce_on_one_gpu_is_resolve(renderer_factory):
    z = np.load(GOLD)
    r = renderer_factory()
    r.add_world(WorldSvoGpu(z["nodes"], z["subs"], z["mats"]))
    ubo = blok_b200.FrameUBO.from_bytes(z["ubo_r"])
    ubo.sampleCount = SPP
    r.set_frame(ubo)
    assert r.comm_rank() == (0, 1) and r.reduce_slice() == (0, 64 * 64)
    r.reset_accum(); r.trace_path(0, SPP); r.resolve()
    want = r.readback(IMG_COLOR).copy()
    r.write_image(IMG_COLOR, np.zeros_like(want))
    r.reduce_accum(REDUCE_ALL)
    assert r.readback(IMG_COLOR).tobytes() == want.tobytes()
    r.reduce_accum(REDUCE_SLICE)
    assert r.readback_slice(IMG_COLOR).tobytes() == want.tobytes()
    with pytest.raises(blok_b200.BlokbError):
        r.reduce_accum(5)


def test_packed_primary_records_match_the_full_ones(renderer_factory):
    z = np.load(GOLD)
    r = renderer_factory()
    w = WorldSvoGpu(z["nodes"], z["subs"], z["mats"])
    r.add_world(w)
    r.set_frame(z["ubo_s"])
    r.trace_primary()
    full = r.readback(IMG_HIT)
    r.trace_primary(packed=True)
    pk = r.readback(IMG_HIT_PACKED)
    assert np.array_equal(w.unpack_hits(pk).view(np.uint8), full.view(np.uint8))
    # strips of the packed image compose like the others; blokb_readback_strips delivers only the strips of the current set
    host = np.zeros_like(pk)
    for g in range(3):
        r.set_strips(g, 3)
        r.trace_primary(packed=True)
        r.readback_strips_into(IMG_HIT_PACKED, host.ctypes.data, host.nbytes)
    r.set_strips(0, 1)
    assert host.tobytes() == pk.tobytes()


def _worker(rank, world, port, out_dir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    z = np.load(GOLD)
    W, H = 97, 61                                   # 5917 pixels: odd, so the slices are ceil(n / 2) and the last one is short
    r = blok_b200.Renderer(W, H, rank)
    r.add_world(WorldSvoGpu(z["nodes"], z["subs"], z["mats"]))
    cam = blok_b200.Camera((-12.0, 14.0, -10.0), 40.0, -20.0)
    r.set_frame(cam.frame(W, H, 2, SPP))
    r.reset_accum(); r.trace_path(0, SPP); r.resolve()
    want = r.readback(IMG_COLOR).copy()             # the undivided frame on this GPU
    status = []
    sharding.connect_comm(r)
    status.append(r.comm_rank() == (rank, world))
    s0, s1 = sharding.sample_range(rank, world, SPP)

    def shard_frame():
        r.reset_accum(); r.trace_path(s0, s1)

    def close(a, b):
        return bool(np.allclose(a, b, rtol=1e-5, atol=1e-6))
    # NCCL reduce-scatter + resolve + all-gather
    shard_frame(); r.reduce_accum(REDUCE_ALL)
    status.append(close(r.readback(IMG_COLOR), want))
    first, count = r.reduce_slice()
    per = -(-W * H // world)
    status.append((first, count) == (min(W * H, per * rank), min(W * H, per * (rank + 1)) - min(W * H, per * rank)))
    r.write_image(IMG_COLOR, np.zeros_like(want))
    shard_frame(); r.reduce_accum(REDUCE_SLICE)
    status.append(close(r.readback_slice(IMG_COLOR), want.reshape(-1, 4)[first:first + count]))
    # accum still holds the LOCAL sums: a later trace_path keeps accumulating
    status.append(bool(np.all(r.readback(IMG_ACCUM)[..., 3] == float(s1 - s0))))
    # in-switch reduction over symmetric memory (NVLink multicast), colour broadcast by multicast stores
    keep = sharding.connect_multicast(r, (IMG_ACCUM, IMG_COLOR), reduce_only=(IMG_ACCUM,))
    mc = keep is not None
    if mc:
        shard_frame(); r.reduce_accum(REDUCE_ALL)
        r.sync(); dist.barrier()
        status.append(close(r.readback(IMG_COLOR), want))
        shard_frame(); r.reduce_accum(REDUCE_SLICE)
        r.sync(); dist.barrier()
        status.append(close(r.readback_slice(IMG_COLOR), want.reshape(-1, 4)[first:first + count]))
        sharding.disconnect_multicast(r, keep)
    # packed records through the tiled-frame exchange
    r.trace_primary(packed=True)
    full_pk = r.readback(IMG_HIT_PACKED).copy()
    keep = sharding.connect_multicast(r, (IMG_HIT_PACKED,))
    if keep is None:
        sharding.connect_peers(r, (IMG_HIT_PACKED,))
    r.set_strips(*sharding.strip_set(rank, world))
    r.trace_primary(packed=True)
    r.sync(); dist.barrier()
    status.append(r.readback(IMG_HIT_PACKED).tobytes() == full_pk.tobytes())
    # the library's own frame fence (multimem.red counter + acquire spin) in place of the barrier: after it, every rank's strips are here
    fk = sharding.connect_fence(r) if keep is not None else None
    if fk is not None:
        import time
        for frame in range(5):
            r.write_image(IMG_HIT_PACKED, np.zeros((H, W), blok_b200.PACKED_DT))   # own copy cleared ...
            r.frame_fence()                                                        # ... on every rank before anyone stores into it
            if rank == frame % world:
                time.sleep(0.05)                                                   # ranks out of step: the fence has to wait for the late one
            r.trace_primary(packed=True)
            r.frame_fence()
            status.append(r.readback(IMG_HIT_PACKED).tobytes() == full_pk.tobytes())
            r.frame_fence()      # nobody clears the image before everyone has read it
        r.fence_status()
        # a rank that fences alone times out after ~1 s; from then on its fences are refused (not silently passed) until the fence is
        # bound again on every rank, which restarts the count
        if rank == 0:
            r.frame_fence(); r.sync()
            for call in (r.frame_fence, r.fence_status):
                try:
                    call(); status.append(False)
                except blok_b200.BlokbError:
                    status.append(True)
        dist.barrier()
        sharding.disconnect_fence(r, fk)
        fk = sharding.connect_fence(r)          # fresh words, epoch 0: its own three proving rounds must pass again
        status.append(fk is not None)
        r.frame_fence(); r.fence_status()
        sharding.disconnect_fence(r, fk)
    if keep is None:
        sharding.disconnect_peers(r, (IMG_HIT_PACKED,))
    else:
        sharding.disconnect_multicast(r, keep)
    r.set_strips(0, 1)
    with open(os.path.join(out_dir, f"ok_{rank}"), "w") as f:
        f.write(("mc:" if mc else "nomc:") + "".join("1" if s else "0" for s in status))
    r.close()
    dist.barrier()
    dist.destroy_process_group()


def test_reduce_accum_two_gpus(tmp_path):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 37500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = [(tmp_path / f"ok_{k}").read_text() for k in range(2)]
    for g in got:
        tag, bits = g.split(":")
        assert set(bits) == {"1"}, got
