"""TEST INFRASTRUCTURE ONLY — loader for tests/host_emul/libemul.so (CPU build of the product's device functions)."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
SO = os.path.join(HERE, "libemul.so")
_vp = C.c_void_p


def build():
    srcs = [os.path.join(HERE, "emul.cpp"), os.path.join(ROOT, "blok_b200", "csrc", "scene_build.cpp")]
    deps = srcs + [os.path.join(ROOT, "blok_b200", "csrc", f) for f in ("traverse.cuh", "shade.cuh", "scene.h", "hd.h", "scene_build.h", "kernels.h")]
    if os.path.exists(SO) and all(os.path.getmtime(SO) >= os.path.getmtime(d) for d in deps):
        return
    cuda_inc = "/usr/local/cuda/include"
    subprocess.check_call(["g++", "-std=c++20", "-O2", "-ffp-contract=off", "-fPIC", "-pthread", "-shared", "-w", "-I" + cuda_inc, *srcs,
                           "-o", SO, "/usr/lib/x86_64-linux-gnu/libstdc++.so.6", "-lm"])


class Emu:
    def __init__(self, lib, world):
        self.lib = lib
        err = C.create_string_buffer(512)
        rc = C.c_int(0)
        n, s, m = world.globalNodes, world.globalSubChunks, world.materials
        self.h = lib.emu_create(n.ctypes.data_as(_vp), len(n), s.ctypes.data_as(_vp), len(s), m.ctypes.data_as(_vp), len(m), err, 512, C.byref(rc))
        self.rc, self.err = rc.value, err.value.decode()

    def ok(self):
        return bool(self.h)

    def trace(self, rays, threads=8, locate=True):
        """locate: start-cell strategy of Walker::init (True = secondary-ray style, False = primary-ray style)."""
        from blok_b200.api import HIT_DT
        rays = np.ascontiguousarray(rays, np.float32)
        out = np.empty(len(rays), HIT_DT)
        steps = self.lib.emu_trace(self.h, rays.ctypes.data_as(_vp), len(rays), out.ctypes.data_as(_vp), threads, int(locate))
        return out, steps

    def path(self, ubo, s_begin, s_end, threads=8):
        from blok_b200.api import FrameUBO
        f = ubo if isinstance(ubo, FrameUBO) else FrameUBO.from_bytes(ubo)
        out = np.empty((f.screenHeight, f.screenWidth, 4), np.float32)
        self.lib.emu_path(self.h, C.byref(f), s_begin, s_end, out.ctypes.data_as(_vp), threads)
        return out

    def stats(self):
        st = np.zeros(6, np.uint64)
        self.lib.emu_stats(self.h, st.ctypes.data_as(_vp))
        return dict(zip(("bricks", "blocks", "voxels", "compactBytes", "levels", "matWide"), (int(v) for v in st)))

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.emu_destroy(self.h)


class EmuLib:
    def __init__(self):
        build()
        L = C.CDLL(SO)
        L.emu_create.restype = _vp
        L.emu_create.argtypes = [_vp, C.c_size_t, _vp, C.c_size_t, _vp, C.c_size_t, C.c_char_p, C.c_size_t, _vp]
        L.emu_destroy.argtypes = [_vp]
        L.emu_trace.restype = C.c_uint64
        L.emu_trace.argtypes = [_vp, _vp, C.c_size_t, _vp, C.c_int, C.c_int]
        L.emu_path.argtypes = [_vp, _vp, C.c_uint32, C.c_uint32, _vp, C.c_int]
        L.emu_stats.argtypes = [_vp, _vp]
        self.L = L

    def world(self, world):
        return Emu(self.L, world)


def load():
    return EmuLib()
