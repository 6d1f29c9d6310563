"""TEST INFRASTRUCTURE ONLY — ctypes loaders for the two CPU checkers.

  RefLib     oracle/_ref/libblokref.so     the reference's own CPU scene build + RT shader sources
                                           (built in this container from /root/reference by oracle/Makefile)
  OracleLib  oracle/_build/libblokoracle.so our independent CPU restatement (oracle/restate.cpp)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
The product (blok_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libblokref.so")
ORACLE_SO = os.path.join(HERE, "_build", "libblokoracle.so")

NODE_DT = np.dtype([("childMask", "<u4"), ("firstChild", "<u4"), ("materialId", "<u4"), ("occupancy", "<f4")])
SUB_DT = np.dtype([("nodeOffset", "<u4"), ("rootNodeIndex", "<u4"), ("nodeCount", "<u4"), ("startDepth", "<u4"),
                   ("worldMin", "<f4", 3), ("subChunkSize", "<f4"), ("worldMax", "<f4", 3), ("pad0", "<f4")])
MAT_DT = np.dtype([("albedo", "<f4", 3), ("flags", "<u4"), ("emission", "<f4", 3), ("ior", "<f4")])
assert NODE_DT.itemsize == 16 and SUB_DT.itemsize == 48 and MAT_DT.itemsize == 32

_vp = C.c_void_p
_u64 = C.c_uint64


def build(ref: bool = True) -> None:
    """Compile the checkers (restatement always; _ref only where /root/reference is mounted)."""
    subprocess.check_call(["make", "-s", "-C", HERE, "restate"])
    if ref and os.path.exists("/root/reference/blok/src/svo.cpp"):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_vp)


def fnv1a64_words(words: np.ndarray) -> str:
    """FNV-1a-64 over the little-endian bytes of a u32 array (SURVEY Appendix B.4 hash)."""
    data = np.ascontiguousarray(words, dtype="<u4").tobytes()
    h = 0xCBF29CE484222325
    # chunked pure-python would be slow for 4 MB; use the byte-serial definition via numpy-free loop in C speed
    for b in data:
        h = ((h ^ b) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return f"{h:016x}"


class Scene:
    """Packed world arrays in the reference's upload format (WorldSvoGpu, resources.hpp:195-210)."""

    def __init__(self, nodes, subs, mats):
        self.nodes = np.ascontiguousarray(nodes, dtype=NODE_DT)
        self.subs = np.ascontiguousarray(subs, dtype=SUB_DT)
        self.mats = np.ascontiguousarray(mats, dtype=MAT_DT)

    def save(self, path):
        np.savez_compressed(path, nodes=self.nodes, subs=self.subs, mats=self.mats)

    @staticmethod
    def load(path):
        z = np.load(path)
        return Scene(z["nodes"], z["subs"], z["mats"])


class _TraceMixin:
    """Shared call shapes: both libraries export <p>_trace / <p>_render with identical signatures."""

    def _setup_trace(self, prefix):
        lib = self.lib
        self._trace = getattr(lib, prefix + "_trace")
        self._trace.argtypes = [_vp, _u64, _vp, _u64, _vp, _u64, C.c_int, _vp, _vp, _vp, C.c_int]
        self._trace.restype = C.c_int
        self._render = getattr(lib, prefix + "_render")
        self._render.argtypes = [_vp, _u64, _vp, _u64, _vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _u64, _u64, _u64, _vp,
                                 C.c_int]
        self._render.restype = C.c_int

    def trace(self, scene: Scene, rays: np.ndarray, any_hit=False, threads=1):
        """rays: (n,8) f32 {o,d,tmin,tmax}. Returns hit (n,4) u32 {prim,node,mat,face}, t (n,) f32, counters (3,) u64."""
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = rays.shape[0]
        hit = np.empty((n, 4), np.uint32)
        t = np.empty(n, np.float32)
        cnt = np.zeros(3, np.uint64)
        rc = self._trace(_ptr(scene.nodes), len(scene.nodes), _ptr(scene.subs), len(scene.subs), _ptr(rays), n,
                         int(any_hit), _ptr(hit), _ptr(t), _ptr(cnt), threads)
        assert rc == 0
        return hit, t, cnt

    def render(self, scene: Scene, ubo: np.ndarray, pix_begin=0, pix_end=None, pix_step=1, threads=1, want=("color", "worldPos", "normalRough", "albedoMetal", "motion")):
        """Full raygen over a pixel subset. Returns dict of (h,w,4) f32 images (untouched pixels are NaN) + counters."""
        u32 = np.frombuffer(ubo.tobytes(), dtype="<u4")
        w, h = int(u32[488 // 4]), int(u32[492 // 4])
        if pix_end is None:
            pix_end = w * h
        imgs = {k: (np.full((h, w, 4), np.nan, np.float32) if k in want else None)
                for k in ("color", "worldPos", "normalRough", "albedoMetal", "motion")}
        cnt = np.zeros(3, np.uint64)
        ubo = np.ascontiguousarray(ubo)
        rc = self._render(_ptr(scene.nodes), len(scene.nodes), _ptr(scene.subs), len(scene.subs), _ptr(scene.mats),
                          len(scene.mats), _ptr(ubo), _ptr(imgs["color"]), _ptr(imgs["worldPos"]),
                          _ptr(imgs["normalRough"]), _ptr(imgs["albedoMetal"]), _ptr(imgs["motion"]), pix_begin,
                          pix_end, pix_step, _ptr(cnt), threads)
        assert rc == 0
        imgs["counters"] = cnt
        return imgs


class RefLib(_TraceMixin):
    def __init__(self, path=REF_SO):
        if not os.path.exists(path):
            raise FileNotFoundError(path + " (build with `make -C oracle ref`; needs /root/reference)")
        self.lib = lib = C.CDLL(path)
        lib.ref_scene_from_vox.argtypes = [C.c_char_p, C.c_float, C.c_float, C.c_float, C.c_uint32]
        lib.ref_scene_from_vox.restype = _vp
        lib.ref_scene_from_voxels.argtypes = [_vp, _vp, C.c_size_t, _vp, C.c_size_t]
        lib.ref_scene_from_voxels.restype = _vp
        lib.ref_scene_counts.argtypes = [_vp, _vp, _vp, _vp]
        lib.ref_scene_copy.argtypes = [_vp, _vp, _vp, _vp]
        lib.ref_scene_voxel_material.argtypes = [_vp, C.c_int32, C.c_int32, C.c_int32]
        lib.ref_scene_voxel_material.restype = C.c_uint32
        lib.ref_scene_free.argtypes = [_vp]
        lib.ref_make_frame.argtypes = [_vp, C.c_float, C.c_float, C.c_float, C.c_uint32, C.c_uint32, C.c_uint32,
                                       C.c_uint32, _vp]
        lib.ref_hardware_threads.restype = C.c_int
        lib.ref_primary_rays.argtypes = [_vp, _vp]
        self._setup_trace("ref")

    def _take(self, h):
        if not h:
            raise RuntimeError("reference scene build failed")
        n = (C.c_uint64 * 3)()
        self.lib.ref_scene_counts(h, C.byref(n, 0), C.byref(n, 8), C.byref(n, 16))
        nodes = np.empty(n[0], NODE_DT)
        subs = np.empty(n[1], SUB_DT)
        mats = np.empty(n[2], MAT_DT)
        self.lib.ref_scene_copy(h, _ptr(nodes), _ptr(subs), _ptr(mats))
        return Scene(nodes, subs, mats)

    def scene_from_vox(self, path, offset=(0, 0, 0), model=0, keep=False):
        h = self.lib.ref_scene_from_vox(path.encode(), *[float(v) for v in offset], model)
        s = self._take(h)
        if keep:
            s.handle = h
        else:
            self.lib.ref_scene_free(h)
        return s

    def scene_from_voxels(self, xyz, mat, mats, keep=False):
        xyz = np.ascontiguousarray(xyz, np.int32)
        mat = np.ascontiguousarray(mat, np.uint32)
        mats = np.ascontiguousarray(mats, MAT_DT)
        h = self.lib.ref_scene_from_voxels(_ptr(xyz), _ptr(mat), len(mat), _ptr(mats), len(mats))
        s = self._take(h)
        if keep:
            s.handle = h
        else:
            self.lib.ref_scene_free(h)
        return s

    def voxel_material(self, scene, x, y, z):
        return self.lib.ref_scene_voxel_material(scene.handle, int(x), int(y), int(z))

    def make_frame(self, pos, yaw, pitch, w, h, fov=60.0, frame_count=0, sample_count=1):
        out = np.zeros(560, np.uint8)
        p = np.asarray(pos, np.float32)
        self.lib.ref_make_frame(_ptr(p), yaw, pitch, fov, w, h, frame_count, sample_count, _ptr(out))
        return out


    def primary_rays(self, ubo):
        u32 = np.frombuffer(ubo.tobytes(), dtype="<u4")
        w, h = int(u32[488 // 4]), int(u32[492 // 4])
        rays = np.empty((w * h, 8), np.float32)
        ubo = np.ascontiguousarray(ubo)
        self.lib.ref_primary_rays(_ptr(ubo), _ptr(rays))
        return rays


class OracleLib(_TraceMixin):
    def __init__(self, path=ORACLE_SO):
        if not os.path.exists(path):
            raise FileNotFoundError(path + " (build with `make -C oracle restate`)")
        self.lib = C.CDLL(path)
        self._setup_trace("orc")
        self.lib.orc_primary_rays.argtypes = [_vp, _vp]
        self.lib.orc_fnv1a64.argtypes = [_vp, _u64]
        self.lib.orc_fnv1a64.restype = _u64
        self.lib.orc_hardware_threads.restype = C.c_int

    def primary_rays(self, ubo):
        u32 = np.frombuffer(ubo.tobytes(), dtype="<u4")
        w, h = int(u32[488 // 4]), int(u32[492 // 4])
        rays = np.empty((w * h, 8), np.float32)
        ubo = np.ascontiguousarray(ubo)
        self.lib.orc_primary_rays(_ptr(ubo), _ptr(rays))
        return rays

    def fnv(self, arr) -> str:
        a = np.ascontiguousarray(arr)
        return f"{self.lib.orc_fnv1a64(_ptr(a), a.nbytes):016x}"

    def threads(self):
        return self.lib.orc_hardware_threads()
