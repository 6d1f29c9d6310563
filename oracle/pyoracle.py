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
        """rays: (n,8) f32 {o,d,tmin,tmax}. Returns hit (n,4) u32 {prim,node,mat,face}, t (n,) f32, counters (4,) u64 {node fetches, AABB passes, MAX_ITER cap hits, rays traced}."""
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = rays.shape[0]
        hit = np.empty((n, 4), np.uint32)
        t = np.empty(n, np.float32)
        cnt = np.zeros(4, np.uint64)
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
        cnt = np.zeros(4, np.uint64)
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
        lib.ref_scene_from_voxels_d.argtypes = [_vp, _vp, _vp, C.c_size_t, _vp, C.c_size_t]
        lib.ref_scene_from_voxels_d.restype = _vp
        lib.ref_scene_counts.argtypes = [_vp, _vp, _vp, _vp]
        lib.ref_scene_copy.argtypes = [_vp, _vp, _vp, _vp]
        lib.ref_scene_voxel_material.argtypes = [_vp, C.c_int32, C.c_int32, C.c_int32]
        lib.ref_scene_voxel_material.restype = C.c_uint32
        lib.ref_scene_free.argtypes = [_vp]
        lib.ref_make_frame.argtypes = [_vp, C.c_float, C.c_float, C.c_float, C.c_uint32, C.c_uint32, C.c_uint32,
                                       C.c_uint32, _vp]
        lib.ref_make_frame_seq.argtypes = [_vp, C.c_float, C.c_float, C.c_float, _vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                           C.c_int32, _vp]
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

    def scene_from_writes(self, xyz, mat, density, mats):
        """ChunkManager::setVoxelMaterial(p, mat, density) per entry, in order (density 0 erases), then rebuild + pack."""
        xyz = np.ascontiguousarray(xyz, np.int32)
        mat = np.ascontiguousarray(mat, np.uint32)
        density = np.ascontiguousarray(density, np.float32)
        mats = np.ascontiguousarray(mats, MAT_DT)
        h = self.lib.ref_scene_from_voxels_d(_ptr(xyz), _ptr(mat), _ptr(density), len(mat), _ptr(mats), len(mats))
        s = self._take(h)
        self.lib.ref_scene_free(h)
        return s

    def voxel_material(self, scene, x, y, z):
        return self.lib.ref_scene_voxel_material(scene.handle, int(x), int(y), int(z))

    def make_frame(self, pos, yaw, pitch, w, h, fov=60.0, frame_count=0, sample_count=1):
        out = np.zeros(560, np.uint8)
        p = np.asarray(pos, np.float32)
        self.lib.ref_make_frame(_ptr(p), yaw, pitch, fov, w, h, frame_count, sample_count, _ptr(out))
        return out


    def make_frame_seq(self, cam, prev_cam, w, h, frame_count, sample_count=8, jitter_index=-1):
        """cam / prev_cam: (pos3, yaw, pitch[, fov]); prev_cam None on the first frame. Renderer::render's FrameUBO for frame n."""
        def flat(c):
            return np.array([*c[0], c[1], c[2], c[3] if len(c) > 3 else 60.0], np.float32)
        a = flat(cam)
        b = None if prev_cam is None else flat(prev_cam)
        out = np.zeros(560, np.uint8)
        self.lib.ref_make_frame_seq(_ptr(a), float(a[3]), float(a[4]), float(a[5]), _ptr(b), w, h, frame_count, sample_count, jitter_index, _ptr(out))
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


# ------------------------------------------------------------------------------------------------------------
# Denoiser + post-process chain on the CPU checker's compute-shader entry point.
# ------------------------------------------------------------------------------------------------------------
FMT_RGBA32F, FMT_RG32F, FMT_R32F, FMT_RGBA16F, FMT_RG16F, FMT_R16F, FMT_RGBA8 = range(7)
_FMT_NP = {FMT_RGBA32F: (np.float32, 4), FMT_RG32F: (np.float32, 2), FMT_R32F: (np.float32, 1),
           FMT_RGBA16F: (np.uint16, 4), FMT_RG16F: (np.uint16, 2), FMT_R16F: (np.uint16, 1), FMT_RGBA8: (np.uint8, 4)}


class _ImageDesc(C.Structure):
    _fields_ = [("data", _vp), ("width", C.c_int32), ("height", C.c_int32), ("fmt", C.c_int32), ("binding", C.c_int32)]


class PostSettings:
    """Denoiser::Settings::atrousIterations (renderer_denoising.hpp:71) + PostProcess::Settings (renderer_postprocess.hpp:102-119).
    The other Denoiser settings travel inside the FrameUBO (fillFrameUBO, renderer_denoising.cpp:686-703)."""

    def __init__(self, **kw):
        self.atrousIterations = 4
        self.enableTAA, self.feedbackMin, self.feedbackMax = True, 0.93, 0.98
        self.enableTonemapping, self.exposure, self.saturationBoost, self.tonemapOperator, self.whitePoint = True, 1.0, 1.15, 1, 0.0
        self.enableSharpening, self.sharpenStrength = True, 0.5
        self.jitter = (0.0, 0.0)  # TAAPushConstants.jitterX/Y (unused by taa.comp's arithmetic)
        for k, v in kw.items():
            assert hasattr(self, k), k
            setattr(self, k, v)


class PostChain:
    """Runs the reference's per-frame pass sequence — Denoiser::denoise (renderer_denoising.cpp:714-769),
    copyCurrentGeometryToHistory (:833-861), PostProcess::process (renderer_postprocess.cpp:505-558), the two
    swapHistoryBuffers (renderer_draw.cpp:327-328) — with the image-to-binding wiring of the two updateDescriptorSets
    (renderer_denoising.cpp:432-556, renderer_postprocess.cpp:330-405). History images start zero-filled (Vulkan leaves
    them undefined; frame 0 never lets them through: frameCount == 0 disables both reprojections)."""

    def __init__(self, lib, prefix, w, h, threads=1, settings=None):
        self.run = getattr(lib, prefix + "_comp_run")
        self.run.argtypes = [C.c_char_p, _vp, C.c_size_t, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int]
        self.run.restype = C.c_int
        self.w, self.h, self.threads = w, h, threads
        self.s = settings or PostSettings()
        z = self._img
        self.histColor = [z(FMT_RGBA32F), z(FMT_RGBA32F)]
        self.histMoments = [z(FMT_RG32F), z(FMT_RG32F)]
        self.histLen = [z(FMT_R16F), z(FMT_R16F)]
        self.histWorldPos = [z(FMT_RGBA32F), z(FMT_RGBA32F)]
        self.histNormal = [z(FMT_RGBA16F), z(FMT_RGBA16F)]
        self.variance, self.ping, self.pong = z(FMT_R32F), z(FMT_RGBA32F), z(FMT_RGBA32F)
        self.taaHist = [z(FMT_RGBA32F), z(FMT_RGBA32F)]
        self.taaOut, self.tonemapOut, self.sharpenOut = z(FMT_RGBA32F), z(FMT_RGBA8), z(FMT_RGBA8)
        self.i = 0  # DenoiserGBuffer::historyIndex
        self.j = 0  # PostProcessBuffers::historyIndex

    def _img(self, fmt):
        dt, n = _FMT_NP[fmt]
        a = np.zeros((self.h, self.w, n), dt)
        return (a, fmt)

    def _dispatch(self, shader, ubo, pc, binds):
        arr = (_ImageDesc * len(binds))()
        for k, (b, (a, fmt)) in enumerate(binds.items()):
            assert a.flags["C_CONTIGUOUS"] and a.shape[:2] == (self.h, self.w)
            arr[k] = _ImageDesc(a.ctypes.data, self.w, self.h, fmt, b)
        pcb = None if pc is None else np.frombuffer(pc, np.uint8).copy()
        rc = self.run(shader.encode(), _ptr(ubo), ubo.nbytes, _ptr(pcb), C.cast(arr, _vp), len(binds), self.w, self.h, self.threads)
        assert rc == 0, (shader, rc)

    def frame(self, ubo, color, worldPos, normalRough16, motion16):
        """color/worldPos (h,w,4) f32, normalRough16 (h,w,4) u16 (binary16 bits), motion16 (h,w,2) u16. Returns the images a
        frame produces (views of internal buffers; copy them to keep across frames)."""
        import struct
        s = self.s
        assert s.atrousIterations >= 1
        ubo = np.ascontiguousarray(ubo, np.uint8)
        fl = np.frombuffer(ubo.tobytes(), "<f4")
        gc = (np.ascontiguousarray(color, np.float32), FMT_RGBA32F)
        gw = (np.ascontiguousarray(worldPos, np.float32), FMT_RGBA32F)
        gn = (np.ascontiguousarray(normalRough16, np.uint16), FMT_RGBA16F)
        gm = (np.ascontiguousarray(motion16, np.uint16), FMT_RG16F)
        i, p = self.i, 1 - self.i
        self._dispatch("temporal", ubo, None, {0: gc, 1: gw, 2: gn, 3: gm, 4: self.histColor[p], 5: self.histMoments[p],
                                               6: self.histLen[p], 7: self.histWorldPos[p], 8: self.histNormal[p],
                                               9: self.histColor[i], 10: self.histMoments[i], 11: self.histLen[i]})
        self._dispatch("variance", ubo, None, {0: gc, 1: self.histMoments[i], 2: self.histLen[i], 3: self.variance, 5: gw, 6: gn})
        phiColor, phiNormal, phiDepth = float(fl[516 // 4]), float(fl[520 // 4]), float(fl[524 // 4])
        src = self.histColor[i]
        for it in range(s.atrousIterations):
            dst = self.ping if it % 2 == 0 else self.pong
            self._dispatch("atrous", ubo, struct.pack("<ifff", 1 << it, phiColor, phiNormal, phiDepth),
                           {0: src, 1: self.variance, 2: gw, 3: gn, 4: dst})
            src = dst
        denoised = src
        self.histWorldPos[i][0][...] = gw[0]
        self.histNormal[i][0][...] = gn[0]
        j, q = self.j, 1 - self.j
        final = denoised
        if s.enableTAA:
            self._dispatch("taa", ubo, struct.pack("<ffff", s.jitter[0], s.jitter[1], s.feedbackMin, s.feedbackMax),
                           {0: denoised, 1: self.taaHist[q], 2: gm, 3: gw, 4: self.taaOut, 5: self.taaHist[j]})
            final = self.taaOut
        if s.enableTonemapping:
            self._dispatch("tonemap", ubo, struct.pack("<ffif", s.exposure, s.saturationBoost, s.tonemapOperator, s.whitePoint),
                           {0: final, 1: self.tonemapOut})
            final = self.tonemapOut
            if s.enableSharpening:
                self._dispatch("sharpen", ubo, struct.pack("<ffff", s.sharpenStrength, 0.0, 0.0, 0.0), {0: self.tonemapOut, 1: self.sharpenOut})
                final = self.sharpenOut
        out = {"historyColor": self.histColor[i][0], "moments": self.histMoments[i][0], "historyLength": self.histLen[i][0],
               "variance": self.variance[0], "denoised": denoised[0], "taaHistory": self.taaHist[j][0], "taaOutput": self.taaOut[0],
               "tonemapped": self.tonemapOut[0], "final": final[0]}
        self.i, self.j = p, q
        return out


def float_to_half_bits(lib, a):
    a = np.ascontiguousarray(a, np.float32)
    out = np.empty(a.shape, np.uint16)
    lib.ref_float_to_half.argtypes = [_vp, _vp, C.c_size_t]
    lib.ref_float_to_half(_ptr(a), _ptr(out), a.size)
    return out
