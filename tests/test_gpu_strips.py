"""GPU: the tile scheduler (blokb_set_strips). Any partition of a frame's 8-row strips over dispatches — the way one frame
is split across GPUs (SURVEY.md 8(e), configs[4]) — writes exactly the bytes of the undivided dispatch."""
import os

import numpy as np
import pytest

import blok_b200
from blok_b200 import (IMG_ACCUM, IMG_ALBEDOMETAL, IMG_COLOR, IMG_HIT, IMG_MOTION, IMG_NORMALROUGH, IMG_SHADOW, IMG_WORLDPOS,
                       WorldSvoGpu, sharding)

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "castle.npz")


def _setup(renderer_factory):
    z = np.load(GOLD)
    r = renderer_factory()
    r.add_world(WorldSvoGpu(z["nodes"], z["subs"], z["mats"]))
    return z, r


@pytest.mark.parametrize("G", [2, 3, 8])
def test_primary_and_shadow_strips_compose_to_the_full_frame(renderer_factory, G):
    z, r = _setup(renderer_factory)
    r.set_frame(z["ubo_s"])
    r.trace_primary(shadow=True)
    full_hit, full_sh = r.readback(IMG_HIT).copy(), r.readback(IMG_SHADOW).copy()
    h = full_hit.shape[0]
    acc_hit, acc_sh = np.zeros_like(full_hit), np.zeros_like(full_sh)
    owner = (np.arange(h) // sharding.STRIP_ROWS) % G
    for g in range(G):
        r.set_strips(*sharding.strip_set(g, G))
        r.trace_primary(shadow=True)
        hit, sh = r.readback(IMG_HIT), r.readback(IMG_SHADOW)
        acc_hit[owner == g] = hit[owner == g]
        acc_sh[owner == g] = sh[owner == g]
        assert sharding.strips_owned(g, G, h) == len(np.unique(np.arange(h)[owner == g] // sharding.STRIP_ROWS))
    r.set_strips(0, 1)
    assert acc_hit.tobytes() == full_hit.tobytes()
    assert acc_sh.tobytes() == full_sh.tobytes()


def test_strip_dispatch_touches_only_its_rows(renderer_factory):
    z, r = _setup(renderer_factory)
    r.set_frame(z["ubo_s"])
    r.trace_primary()
    full = r.readback(IMG_HIT).copy()
    marker = np.zeros_like(full)
    marker["node"] = 0x12345678
    r.write_image(IMG_HIT, marker)
    r.set_strips(1, 3)
    r.trace_primary()
    got = r.readback(IMG_HIT)
    mine = ((np.arange(full.shape[0]) // sharding.STRIP_ROWS) % 3) == 1
    assert got[mine].tobytes() == full[mine].tobytes()
    assert np.all(got[~mine]["node"] == 0x12345678)
    # a strip set beyond the frame is an empty dispatch, not an error
    r.set_strips(10_000, 3)
    r.trace_primary()
    r.sync()
    with pytest.raises(blok_b200.BlokbError):
        r.set_strips(0, 0)


def test_render_and_path_strips_compose(renderer_factory):
    z, r = _setup(renderer_factory)
    r.set_frame(z["ubo_r"])
    ids = (IMG_COLOR, IMG_WORLDPOS, IMG_NORMALROUGH, IMG_ALBEDOMETAL, IMG_MOTION)
    r.render()
    full = {i: r.readback(i).copy() for i in ids}
    r.reset_accum()
    r.trace_path(0, 3)
    full_acc = r.readback(IMG_ACCUM).copy()
    G = 4
    h = full_acc.shape[0]
    owner = (np.arange(h) // sharding.STRIP_ROWS) % G
    parts = {i: np.zeros_like(full[i]) for i in ids}
    acc = np.zeros_like(full_acc)
    r.reset_accum()
    for g in range(G):
        r.set_strips(g, G)
        r.render()
        for i in ids:
            parts[i][owner == g] = r.readback(i)[owner == g]
        r.trace_path(0, 3)
    acc = r.readback(IMG_ACCUM)  # each strip accumulated exactly once over the G dispatches
    for i in ids:
        assert parts[i].tobytes() == full[i].tobytes()
    assert acc.tobytes() == full_acc.tobytes()


def _peer_worker(rank, world, port, out_dir):
    """Two processes on cuda:0 (CUDA IPC works between processes on one device too): rank g renders strips g, g + 2, ...
    and the kernel stores them into BOTH processes' images; after a barrier each holds the full frame."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    z = np.load(GOLD)
    r = blok_b200.Renderer(64, 64, 0)
    r.add_world(WorldSvoGpu(z["nodes"], z["subs"], z["mats"]))
    r.set_frame(z["ubo_r"])
    images = (IMG_HIT, IMG_SHADOW, IMG_ACCUM)
    r.trace_primary(shadow=True)
    r.reset_accum()
    r.trace_path(0, 2)
    full = {i: r.readback(i).copy() for i in images}
    for i in images:
        r.write_image(i, np.zeros_like(full[i]))
    sharding.connect_peers(r, images)
    r.set_strips(*sharding.strip_set(rank, world))
    r.trace_primary(shadow=True)
    r.trace_path(0, 2)       # accum was zeroed on both sides: each strip accumulated once by its owner, replicated to the peer
    r.sync()
    dist.barrier()
    ok = all(r.readback(i).tobytes() == full[i].tobytes() for i in images)
    sharding.disconnect_peers(r, images)
    r.set_strips(0, 1)
    r.trace_primary(shadow=True)   # unmapped again: plain local dispatch still works
    ok = ok and r.readback(IMG_HIT).tobytes() == full[IMG_HIT].tobytes()
    with open(os.path.join(out_dir, f"ok_{rank}"), "w") as f:
        f.write("1" if ok else "0")
    r.close()
    dist.barrier()
    dist.destroy_process_group()


def test_fused_exchange_between_two_processes(tmp_path):
    import torch.multiprocessing as mp
    port = 33500 + (os.getpid() % 2000)
    mp.spawn(_peer_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok_0").read_text() == "1" and (tmp_path / "ok_1").read_text() == "1"


def test_bind_image_substitutes_external_memory(renderer_factory):
    """blokb_bind_image without a multicast address: the kernels write into caller-owned memory, readback follows it, and
    unbinding restores the context's own buffer."""
    import torch
    z, r = _setup(renderer_factory)
    r.set_frame(z["ubo_s"])
    r.trace_primary()
    full = r.readback(IMG_HIT).copy()
    ext = torch.zeros(full.nbytes, dtype=torch.uint8, device="cuda:0")
    r.bind_image(IMG_HIT, ext.data_ptr(), None)
    r.trace_primary()
    r.sync()
    assert ext.cpu().numpy().tobytes() == full.tobytes()
    assert r.readback(IMG_HIT).tobytes() == full.tobytes()
    r.bind_image(IMG_HIT)
    ext.zero_()
    r.trace_primary()
    assert r.readback(IMG_HIT).tobytes() == full.tobytes() and int(ext.sum()) == 0
    with pytest.raises(blok_b200.BlokbError):
        r.bind_image(IMG_HIT, ext.data_ptr() + 4, None)   # misaligned
    with pytest.raises(blok_b200.BlokbError):
        r.bind_image(blok_b200.IMG_WORLDPOS, ext.data_ptr(), None)     # not an exchangeable image


def _mc_worker(rank, world, port, out_dir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    z = np.load(GOLD)
    r = blok_b200.Renderer(64, 64, rank)
    r.add_world(WorldSvoGpu(z["nodes"], z["subs"], z["mats"]))
    r.set_frame(z["ubo_r"])
    images = (IMG_HIT, IMG_SHADOW, IMG_ACCUM)
    r.trace_primary(shadow=True)
    r.reset_accum()
    r.trace_path(0, 2)
    full = {i: r.readback(i).copy() for i in images}
    keep = sharding.connect_multicast(r, images)
    status = "nomc"
    if keep is not None:
        r.set_strips(*sharding.strip_set(rank, world))
        r.trace_primary(shadow=True)
        r.trace_path(0, 2)           # the symmetric images start zeroed
        r.sync()
        dist.barrier()
        status = "1" if all(r.readback(i).tobytes() == full[i].tobytes() for i in images) else "0"
        sharding.disconnect_multicast(r, keep)
        r.set_strips(0, 1)
        r.trace_primary(shadow=True)
        if r.readback(IMG_HIT).tobytes() != full[IMG_HIT].tobytes():
            status = "0"
    with open(os.path.join(out_dir, f"ok_{rank}"), "w") as f:
        f.write(status)
    r.close()
    dist.barrier()
    dist.destroy_process_group()


def test_multicast_exchange_two_gpus(tmp_path):
    """Needs two GPUs on an NVSwitch (run with `gpurun --gpus 2`); skipped on the single-GPU box."""
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 35500 + (os.getpid() % 2000)
    mp.spawn(_mc_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    got = [(tmp_path / f"ok_{k}").read_text() for k in range(2)]
    if got == ["nomc", "nomc"]:
        pytest.skip("no NVLink multicast support on this box")
    assert got == ["1", "1"]
