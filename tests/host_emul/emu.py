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
    deps = srcs + [os.path.join(ROOT, "blok_b200", "csrc", f) for f in ("traverse.cuh", "shade.cuh", "post.cuh", "post.h", "scene.h", "hd.h", "scene_build.h", "kernels.h")]
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

    def trace_chain(self, first, second, locate_first=False):
        """second[i] traced right after first[i] on the same node stack (WalkResume). Returns (records of second, #resumed)."""
        from blok_b200.api import HIT_DT
        first, second = np.ascontiguousarray(first, np.float32), np.ascontiguousarray(second, np.float32)
        assert first.shape == second.shape
        out = np.empty(len(first), HIT_DT)
        self.lib.emu_trace_chain.argtypes = [_vp, _vp, _vp, C.c_size_t, _vp, C.c_int]
        self.lib.emu_trace_chain.restype = C.c_uint64
        n = self.lib.emu_trace_chain(self.h, first.ctypes.data_as(_vp), second.ctypes.data_as(_vp), len(first), out.ctypes.data_as(_vp), int(locate_first))
        return out, int(n)

    def warp_vote(self, rays, num, den, locate=False):
        """The vote-scheduled what-if loop (emu_warp_vote): 12 counters, [11] = iterations."""
        rays = np.ascontiguousarray(rays, np.float32)
        out = np.zeros(12, np.uint64)
        self.lib.emu_warp_vote.argtypes = [_vp, _vp, C.c_size_t, C.c_int, C.c_int, C.c_int, _vp]
        self.lib.emu_warp_vote.restype = None
        self.lib.emu_warp_vote(self.h, rays.ctypes.data_as(_vp), len(rays) // 32, int(locate), int(num), int(den), out.ctypes.data_as(_vp))
        return out

    def walk_profile(self, rays, locate=True):
        """Per child-cell size (row index = sh / 2: 0 voxels, 1 bricks, ...): descents, descents that needed plane tests,
        steps, climbs landing there. Returns (counts[4, kMaxLevels + 1], hits). Single-threaded."""
        rays = np.ascontiguousarray(rays, np.float32)
        prof = np.zeros((4, 13), np.uint64)  # kMaxLevels = 12 (scene.h)
        self.lib.emu_walk_profile.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp]
        self.lib.emu_walk_profile.restype = C.c_uint64
        hits = self.lib.emu_walk_profile(self.h, rays.ctypes.data_as(_vp), len(rays), int(locate), prof.ctypes.data_as(_vp))
        return prof, int(hits)

    def warp_profile(self, rays, locate=True, max_desc_passes=0):
        """32 Walkers in lock step per warp (rays[32 * w + l] = lane l of warp w): {warp issues, lane executions} of
        descend, select-in-descend, step, widen-in-step, climb-in-step, and the warp count (11 counters)."""
        rays = np.ascontiguousarray(rays, np.float32)
        assert len(rays) % 32 == 0
        out = np.zeros(11, np.uint64)
        self.lib.emu_warp_profile.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp, C.c_int]
        self.lib.emu_warp_profile.restype = None
        self.lib.emu_warp_profile(self.h, rays.ctypes.data_as(_vp), len(rays) // 32, int(locate), out.ctypes.data_as(_vp), int(max_desc_passes))
        return out

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


class _TemporalIO(C.Structure):
    _fields_ = [(n, _vp) for n in ("color", "worldPos", "normalRough", "motion", "prevHistory", "prevMoments", "prevLength", "prevWorldPos",
                                   "prevNormal", "outHistory", "outMoments", "outLength", "outVariance")]


class EmuPostChain:
    """The product's denoise / postprocess / end_frame sequence (blok_b200/csrc/api.cpp) with the kernels replaced by host
    loops over the same per-pixel functions (post.cuh). CPU-only logic check against the reference's compute shaders."""

    def __init__(self, lib, w, h, settings=None, threads=8):
        from blok_b200.api import PostSettings
        lib = getattr(lib, "L", lib)
        self.lib, self.w, self.h, self.threads = lib, w, h, threads
        self.s = settings or PostSettings()
        lib.emu_post_frame_size.restype = C.c_size_t
        lib.emu_post_frame.argtypes = [_vp, _vp, _vp]
        lib.emu_post_temporal_variance.argtypes = [_vp, _vp, C.c_int]
        lib.emu_post_atrous.argtypes = [_vp, C.c_int, _vp, _vp, _vp, _vp, _vp, C.c_int]
        lib.emu_post_taa_tonemap.argtypes = [_vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int]
        lib.emu_post_sharpen.argtypes = [_vp, _vp, _vp, C.c_int]
        z = lambda n, dt: np.zeros((h, w, n), dt)
        self.histColor = [z(4, np.float32), z(4, np.float32)]
        self.histMoments = [z(2, np.float32), z(2, np.float32)]
        self.histLen = [z(1, np.uint16), z(1, np.uint16)]
        self.prevWorldPos, self.prevNormal = z(4, np.float32), z(4, np.uint16)
        self.variance, self.ping, self.pong = z(1, np.float32), z(4, np.float32), z(4, np.float32)
        self.taaHist = [z(4, np.float32), z(4, np.float32)]
        self.taaOut, self.tonemapOut, self.sharpenOut = z(4, np.float32), z(4, np.uint8), z(4, np.uint8)
        self.i = self.j = 0

    def frame(self, ubo, color, worldPos, normalRough16, motion16):
        from blok_b200.api import FrameUBO
        L, s, nt = self.lib, self.s, self.threads
        f = ubo if isinstance(ubo, FrameUBO) else FrameUBO.from_bytes(ubo)
        P = C.create_string_buffer(L.emu_post_frame_size())
        L.emu_post_frame(C.byref(f), C.byref(s), P)
        p = lambda a: a.ctypes.data_as(_vp)
        gc, gw = np.ascontiguousarray(color, np.float32), np.ascontiguousarray(worldPos, np.float32)
        gn, gm = np.ascontiguousarray(normalRough16, np.uint16), np.ascontiguousarray(motion16, np.uint16)
        i, q = self.i, 1 - self.i
        io = _TemporalIO(p(gc), p(gw), p(gn), p(gm), p(self.histColor[q]), p(self.histMoments[q]), p(self.histLen[q]), p(self.prevWorldPos),
                         p(self.prevNormal), p(self.histColor[i]), p(self.histMoments[i]), p(self.histLen[i]), p(self.variance))
        L.emu_post_temporal_variance(P, C.byref(io), nt)
        src = self.histColor[i]
        for it in range(s.atrousIterations):
            dst = self.ping if it % 2 == 0 else self.pong
            L.emu_post_atrous(P, 1 << it, p(src), p(self.variance), p(gw), p(gn), p(dst), nt)
            src = dst
        denoised = src
        j, r = self.j, 1 - self.j
        L.emu_post_taa_tonemap(P, int(s.enableTAA), int(s.enableTonemapping), p(denoised), p(self.taaHist[r]), p(gm), p(self.taaHist[j]),
                               p(self.taaOut), p(self.tonemapOut), nt)
        final = self.tonemapOut if s.enableTonemapping else self.taaOut
        if s.enableTonemapping and s.enableSharpening:
            L.emu_post_sharpen(P, p(self.tonemapOut), p(self.sharpenOut), nt)
            final = self.sharpenOut
        out = {"historyColor": self.histColor[i], "moments": self.histMoments[i], "historyLength": self.histLen[i][..., 0],
               "variance": self.variance[..., 0], "denoised": denoised, "taaHistory": self.taaHist[j], "tonemapped": self.tonemapOut, "final": final}
        # end_frame: geometry images change roles (the product swaps pointers; here the caller owns gw/gn, so keep references)
        self.prevWorldPos, self.prevNormal = gw, gn
        self.i, self.j = q, r
        return out
