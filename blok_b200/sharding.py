"""Partition of one frame across GPUs. Rays are independent and the scene is replicated, so the
only exchange is the reduce / gather of per-GPU accumulation buffers over torch.distributed (NCCL on GPUs, gloo in
the CPU tests).

  sample_range(rank, world, spp)  path tracing: GPU g renders sampleIdx in [spp*g/G, spp*(g+1)/G). With raygen.rgen's
                                  per-(pixel, frame, sample) seeding (raygen.rgen:92-99) the union of samples is exactly
                                  the single-GPU sample set; only the fp32 summation order differs.
  row_band(rank, world, height)   primary rays: contiguous 8-row-aligned bands (a kernel block is 32x8 pixels).
  strip_set(rank, world)          primary / G-buffer passes of ONE frame: GPU g renders the 8-row strips g, g + G, ... of
                                  the full frame (Renderer.set_strips) — interleaved so sky and ground rows are shared
                                  evenly; gather_strips all-gathers the strips so every rank ends with the full image.
  connect_peers(renderer, images) the fused form of that exchange: ranks swap CUDA IPC handles of their hit / shadow / accum
                                  images and map each other's buffers (Renderer.set_peers); the traversal kernels then store
                                  every record into all ranks' images over NVLink, and frame_barrier() — one tiny all-reduce
                                  on the dispatch stream — is the only collective left in the frame.
  allreduce_accum(t)              sum of (sum.rgb, spp) buffers; every rank ends with the full frame.
  reduce_scatter_resolve(...)     reduce-scatter of the accumulation buffer, divide-by-spp + firefly clamp on the owned
                                  slice, all-gather of the resolved slices (1/G of the all-reduce's resolve work).
"""
from typing import Tuple


def sample_range(rank: int, world: int, spp: int) -> Tuple[int, int]:
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return (spp * rank) // world, (spp * (rank + 1)) // world


def row_band(rank: int, world: int, height: int, align: int = 8) -> Tuple[int, int]:
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    blocks = (height + align - 1) // align
    b0, b1 = (blocks * rank) // world, (blocks * (rank + 1)) // world
    return min(b0 * align, height), min(b1 * align, height)


STRIP_ROWS = 8  # one row of 32x8-pixel kernel blocks (csrc/kernels.cu)


def strip_set(rank: int, world: int) -> Tuple[int, int]:
    """(first, stride) for Renderer.set_strips: rank g of G owns strips g, g + G, g + 2G, ..."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    return rank, world


def strips_owned(rank: int, world: int, height: int) -> int:
    strips = (height + STRIP_ROWS - 1) // STRIP_ROWS
    return (strips - rank + world - 1) // world if rank < strips else 0


def gather_strips(image, group=None):
    """image: torch tensor (h, w, ...) holding this rank's strips at their full-frame positions (what the kernels wrote
    after set_strips(rank, world)). All-gathers the strips; returns the full (h, w, ...) image, on every rank.
    One exchange step: pack own strips (1/G of the image) -> all_gather_into_tensor -> un-interleave."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return image
    G, g = dist.get_world_size(group), dist.get_rank(group)
    h = image.shape[0]
    strips = (h + STRIP_ROWS - 1) // STRIP_ROWS
    per = (strips + G - 1) // G                      # strips per rank, padded
    hp = per * G * STRIP_ROWS                        # padded height
    rest = tuple(image.shape[1:])
    if hp != h:
        padded = torch.zeros((hp,) + rest, dtype=image.dtype, device=image.device)
        padded[:h] = image
    else:
        padded = image
    v = padded.reshape((per, G, STRIP_ROWS) + rest)  # strip s = per-index s // G, owner s % G
    mine = v[:, g].contiguous()
    out = torch.empty((G,) + tuple(mine.shape), dtype=image.dtype, device=image.device)
    dist.all_gather_into_tensor(out.view(-1), mine.view(-1), group=group)
    full = out.transpose(0, 1).reshape((hp,) + rest)[:h]
    return full.contiguous()


def connect_peers(renderer, images, group=None):
    """Map every rank's `images` (IMG_HIT / IMG_SHADOW / IMG_ACCUM) into this rank's renderer. Call after set_frame (the
    images exist) on every rank; the renderers must keep their resolution while connected."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    G, g = dist.get_world_size(group), dist.get_rank(group)
    for img in images:
        handles = [None] * G
        dist.all_gather_object(handles, renderer.ipc_export(img), group=group)
        renderer.set_peers(img, handles, g)
    dist.barrier(group=group)


def connect_multicast(renderer, images, group=None, reduce_only=()):
    """The fused exchange through NVLink multicast: each of `images` is re-homed in torch symmetric memory bound to a
    multicast object (Renderer.bind_image); kernels then store every record once and the NVSwitch replicates it into all
    ranks' copies. Returns the keep-alive list [(image, tensor, handle)] (pass it to disconnect_multicast), or None when
    the platform has no multicast support — callers fall back to connect_peers. Images listed in reduce_only (IMG_ACCUM of a
    sample-sharded frame) keep their dispatch stores local: the multicast address only feeds Renderer.reduce_accum."""
    import torch
    import torch.distributed as dist
    import torch.distributed._symmetric_memory as symm_mem
    from .api import IMG_ACCUM, IMG_HIT, IMG_SHADOW
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return None
    b = renderer.buffers()
    n = int(b.width) * int(b.height)
    from .api import IMG_COLOR, IMG_HIT_PACKED
    per_pixel = {IMG_HIT: 16, IMG_SHADOW: 1, IMG_ACCUM: 16, IMG_HIT_PACKED: 8, IMG_COLOR: 16}
    pg = group if group is not None else dist.group.WORLD
    name = pg.group_name if hasattr(pg, "group_name") else pg
    dev = torch.device("cuda", torch.cuda.current_device())
    keep = []
    for img in images:
        nbytes = (n * per_pixel[img] + 15) // 16 * 16
        t = symm_mem.empty(nbytes, dtype=torch.uint8, device=dev)
        h = symm_mem.rendezvous(t, name)
        if not int(h.multicast_ptr):
            for im, _, _ in keep:
                renderer.bind_image(im)
            return None
        t.zero_()
        keep.append((img, t, h))
    torch.cuda.synchronize()
    dist.barrier(group=group)
    for img, t, h in keep:
        renderer.bind_image(img, t.data_ptr(), int(h.multicast_ptr), reduce_only=img in reduce_only)
    dist.barrier(group=group)
    return keep


def connect_comm(renderer, group=None):
    """Gives the renderer's context its own NCCL communicator over the ranks of `group` (blokb_comm_init): rank 0 makes the
    ncclUniqueId, torch.distributed only carries its 128 bytes to the other ranks. After this, Renderer.reduce_accum needs no
    Python in the frame loop — exactly what a C++ host does with blokb_comm_unique_id / blokb_comm_init."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    rank, size = dist.get_rank(group), dist.get_world_size(group)
    box = [renderer.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    renderer.comm_init(box[0], rank, size)


def connect_fence(renderer, group=None):
    """The library's own frame fence (Renderer.frame_fence) instead of a collective call: a zero-filled 16-byte word pair in
    symmetric memory bound to a multicast object, shared by the ranks of `group`. Returns the keep-alive (tensor, handle), or None
    when the platform has no multicast support (callers keep using frame_barrier)."""
    import torch
    import torch.distributed as dist
    import torch.distributed._symmetric_memory as symm_mem
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return None
    pg = group if group is not None else dist.group.WORLD
    name = pg.group_name if hasattr(pg, "group_name") else pg
    t = symm_mem.empty(16, dtype=torch.uint8, device=torch.device("cuda", torch.cuda.current_device()))
    h = symm_mem.rendezvous(t, name)
    if not int(h.multicast_ptr):
        return None
    t.zero_()
    torch.cuda.synchronize()
    dist.barrier(group=group)
    renderer.bind_fence(t.data_ptr(), int(h.multicast_ptr), dist.get_world_size(group))
    dist.barrier(group=group)
    # prove it on this box before anyone relies on it: three rounds, then the error word; all ranks agree on the verdict
    ok = 1
    try:
        for _ in range(3):
            renderer.frame_fence()
        renderer.fence_status()
    except Exception:
        ok = 0
    flag = torch.tensor([ok], device=t.device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    if int(flag.item()) == 0:
        torch.cuda.synchronize()
        renderer.bind_fence()
        dist.barrier(group=group)
        return None
    return (t, h)


def disconnect_fence(renderer, keep, group=None):
    import torch
    import torch.distributed as dist
    torch.cuda.synchronize()
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.barrier(group=group)
    renderer.bind_fence()
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.barrier(group=group)


def disconnect_multicast(renderer, keep, group=None):
    import torch
    import torch.distributed as dist
    torch.cuda.synchronize()
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.barrier(group=group)
    for img, _, _ in keep or []:
        renderer.bind_image(img)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.barrier(group=group)


def disconnect_peers(renderer, images, group=None):
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.barrier(group=group)     # nobody may still be writing into a buffer that is about to be unmapped / freed
    for img in images:
        renderer.set_peers(img, None, 0)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.barrier(group=group)


def frame_barrier(token, group=None):
    """After a fused-exchange dispatch: `token` is a small tensor on the dispatching device, all-reduced on the CURRENT
    torch stream — when it completes, every rank's kernel has finished, so every rank's image is complete. Also the
    fence before the next frame may overwrite the images."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(token, group=group)
    return token


def allreduce_accum(accum, group=None):
    """accum: torch tensor (h, w, 4) float32 on this rank's device; summed in place across ranks."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(accum, op=dist.ReduceOp.SUM, group=group)
    return accum


def resolve(accum):
    """color = sum.rgb / spp with raygen.rgen's firefly clamp (:383-392), as torch ops (used on the reduced slice)."""
    import torch
    c = accum[..., :3] / accum[..., 3:4]
    mv = c.max(dim=-1, keepdim=True).values
    c = torch.where(mv > 100.0, c * (100.0 / mv), c)
    return torch.cat([c, torch.ones_like(c[..., :1])], dim=-1)


def reduce_scatter_resolve(accum, group=None):
    """accum (h, w, 4) f32, h*w divisible by world. Returns the resolved full frame (h, w, 4) on every rank."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return resolve(accum)
    G = dist.get_world_size(group)
    flat = accum.reshape(-1, 4)
    n = flat.shape[0]
    if n % G:
        dist.all_reduce(accum, op=dist.ReduceOp.SUM, group=group)
        return resolve(accum)
    mine = torch.empty((n // G, 4), dtype=flat.dtype, device=flat.device)
    dist.reduce_scatter_tensor(mine, flat.contiguous(), op=dist.ReduceOp.SUM, group=group)
    res = resolve(mine)
    out = torch.empty((n, 4), dtype=flat.dtype, device=flat.device)
    dist.all_gather_into_tensor(out, res.contiguous(), group=group)
    return out.reshape(accum.shape)
