"""Named benchmark workloads = BASELINE.json's configs at full size; all synthetic and deterministic.

  c2  512^3 dense voxelised "mesh" (union of 40 ellipsoids filling the box), 1920x1080, primary + sun shadow rays, camera
      at (-0.4N, 0.6N, -0.4N) looking at the centre; the fov (40 degrees) frames the 512^3 box: 64 % of
      the primary rays hit                                                   (BASELINE configs[1])
  c3  2048^3 heightfield-shell terrain, 3840x2160, primary rays              (configs[2])  <- bench.py headline
  c4  the c3 scene path traced: 64 spp, <=2 bounces + 1 shadow ray, spp sharded across GPUs   (configs[3])
  c5  8192^3 terrain, 7680x4320, 1 spp path (primary + shadow + incoherent bounce)            (configs[4]; ~6 GB of
      reference-format input, minutes of host-side generation: run by hand, `python tools/prof.py c5`)
  t1  96^3 terrain, 256x144 — CPU-test / smoke scale of the same shapes

Every workload gives {scene generator, N, width, height, camera, kind, spp}. Camera poses follow §8(d): the c3/c4
camera sits at (N/2, 0.6N, -0.2N) looking along +z, pitched -25 degrees.
"""
from . import scenes
from .api import Camera, WorldSvoGpu

WORKLOADS = {
    "c2": dict(scene="dense_mesh", N=512, width=1920, height=1080, kind="primary_shadow", spp=1,
               cam=lambda N: _look_at((-0.4 * N, 0.6 * N, -0.4 * N), (0.5 * N, 0.5 * N, 0.5 * N), 40.0),
               desc="512^3 dense voxelized mesh (40 ellipsoids), 1920x1080 primary + sun shadow rays, camera (-0.4N, 0.6N, -0.4N) -> centre"),
    "c3": dict(scene="terrain", N=2048, width=3840, height=2160, kind="primary", spp=1,
               cam=lambda N: Camera((N / 2, 0.6 * N, -0.2 * N), 90.0, -25.0),
               desc="2048^3 SVO terrain shell, 3840x2160 primary rays"),
    "c4": dict(scene="terrain", N=2048, width=3840, height=2160, kind="path", spp=64,
               cam=lambda N: Camera((N / 2, 0.6 * N, -0.2 * N), 90.0, -25.0),
               desc="2048^3 SVO terrain shell, 3840x2160 path traced, 64 spp, 2 bounces + sun shadow ray"),
    "c5": dict(scene="terrain", N=8192, width=7680, height=4320, kind="path", spp=1,
               cam=lambda N: Camera((N / 2, 0.6 * N, -0.2 * N), 90.0, -25.0),
               desc="8192^3 procedural terrain shell, 7680x4320, primary + sun shadow + one incoherent bounce ray per pixel"),
    "t1": dict(scene="terrain", N=96, width=256, height=144, kind="primary", spp=1,
               cam=lambda N: Camera((N / 2, 0.7 * N, -0.3 * N), 90.0, -25.0),
               desc="96^3 terrain shell, 256x144 primary rays (test scale)"),
}


def _look_at(pos, target, fov=60.0) -> Camera:
    """Camera at pos whose forward axis (camera.hpp:42-50: yaw about +y from +x, pitch up) points at target."""
    import math
    d = [t - p for t, p in zip(target, pos)]
    return Camera(pos, math.degrees(math.atan2(d[2], d[0])), math.degrees(math.atan2(d[1], math.hypot(d[0], d[2]))), fov)


def voxels(name: str):
    """(xyz, material ids, materials) of the workload's voxel set — what ChunkManager::setVoxelMaterial would be fed."""
    w = WORKLOADS[name]
    return getattr(scenes, w["scene"])(w["N"])


def build_world(name: str) -> WorldSvoGpu:
    """Voxel set of the workload -> the reference's upload arrays (host C++ restatement of the reference builder)."""
    xyz, mid, mats = voxels(name)
    return WorldSvoGpu.from_voxels(xyz, mid, mats)


def camera_for(name: str, rank: int = 0) -> Camera:
    """Rank r's camera: the §8(d) pose, translated along +x by 16 voxels per rank (weak scaling renders one frame per
    GPU per step — the frames-in-flight of an 8-GPU box — so each rank needs its own view of the same scene)."""
    w = WORKLOADS[name]
    c = w["cam"](w["N"])
    if rank:
        c.position = (c.position[0] + 16.0 * rank, c.position[1], c.position[2])
    return c


def rays_per_step(name: str) -> int:
    """Primary rays only (shadow / bounce rays are counted on the device for the kinds that cast them)."""
    w = WORKLOADS[name]
    return w["width"] * w["height"]
