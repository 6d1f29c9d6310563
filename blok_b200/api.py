"""ctypes binding of include/blokb200.h plus a thin host-side mirror of the reference's seams.

Class / method names follow the reference so that callers (and the parity tests) read like code written
against blok itself:

    Renderer.add_world(world)        <- Renderer::addWorld(WorldSvoGpu&)          blok/include/renderer.hpp:40-54
    Renderer.update_world(world)     <- Renderer::updateWorld()                    blok/include/renderer.hpp:55-71
    Renderer.render(camera | ubo)    <- Renderer::render(const Camera&, float dt)  blok/include/renderer.hpp:34
    Renderer.reset_accum()           <- CudaTracer::resetAccum()                   blok/src/cuda_tracer.cu:478-482
    Renderer.denoise(settings)       <- Denoiser::denoise                          blok/src/renderer_denoising.cpp:714-769
    Renderer.postprocess(settings)   <- PostProcess::process                       blok/src/renderer_postprocess.cpp:505-558
    Renderer.end_frame()             <- copyCurrentGeometryToHistory + swapHistoryBuffers (renderer_draw.cpp:182, 327-328)
    PostSettings                     <- Denoiser::Settings::atrousIterations + PostProcess::Settings
    WorldSvoGpu(nodes, subs, mats)   <- blok::WorldSvoGpu                          blok/include/resources.hpp:195-210
    Camera(position, yaw, pitch, fov)<- blok::Camera                               blok/include/camera.hpp:15-59

There is no CPU fallback: if libblokb200.so is missing this module raises at import of the library, and every
compute call fails loudly (BlokbError) when no CUDA device is present.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BLOKB_LIB_PATH") or os.path.join(HERE, "libblokb200.so")  # override: kernel experiments (tools/)

NODE_DT = np.dtype([("childMask", "<u4"), ("firstChild", "<u4"), ("materialId", "<u4"), ("occupancy", "<f4")])
SUB_DT = np.dtype([("nodeOffset", "<u4"), ("rootNodeIndex", "<u4"), ("nodeCount", "<u4"), ("startDepth", "<u4"),
                   ("worldMin", "<f4", 3), ("subChunkSize", "<f4"), ("worldMax", "<f4", 3), ("pad0", "<f4")])
MAT_DT = np.dtype([("albedo", "<f4", 3), ("flags", "<u4"), ("emission", "<f4", 3), ("ior", "<f4")])
HIT_DT = np.dtype([("node", "<u4"), ("material", "<u4"), ("t", "<f4"), ("facePrim", "<u4")])
assert NODE_DT.itemsize == 16 and SUB_DT.itemsize == 48 and MAT_DT.itemsize == 32 and HIT_DT.itemsize == 16
PACKED_DT = np.dtype([("nodeFace", "<u4"), ("t", "<f4")])  # blokb_hit_packed
PACKED_MAX_NODES = 0x1FFFFFFF

IMG_HIT, IMG_SHADOW, IMG_COLOR, IMG_WORLDPOS, IMG_NORMALROUGH, IMG_ALBEDOMETAL, IMG_MOTION, IMG_ACCUM = range(8)
(IMG_HISTORY_COLOR, IMG_MOMENTS, IMG_HISTORY_LENGTH, IMG_VARIANCE, IMG_DENOISED, IMG_TAA_HISTORY, IMG_TAA_OUTPUT, IMG_TONEMAPPED,
 IMG_FINAL, IMG_HIT_PACKED) = range(8, 18)
REDUCE_ALL, REDUCE_SLICE = 0, 1

WORLD_NODES, WORLD_BLOCKFIRSTLEAF, WORLD_VOXELMAT, WORLD_SUBCHUNKS = range(4)
CNODE_DT = np.dtype([("mask", "<u8"), ("base", "<u4"), ("aux", "<u4")])  # blok_b200/csrc/scene.h Node

_vp = C.c_void_p


class BlokbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"blokb error {code}: {msg}")
        self.code = code


class FrameUBO(C.Structure):
    """blok::FrameUBO (resources.hpp:103-150), 560 bytes."""
    _fields_ = [(n, C.c_float * 16) for n in ("view", "proj", "invView", "invProj", "prevView", "prevProj", "prevViewProj")] + [
        ("camPos", C.c_float * 3), ("deltaTime", C.c_float), ("prevCamPos", C.c_float * 3), ("depth", C.c_uint32),
        ("frameCount", C.c_uint32), ("sampleCount", C.c_uint32), ("screenWidth", C.c_uint32), ("screenHeight", C.c_uint32),
        ("temporalAlpha", C.c_float), ("momentAlpha", C.c_float), ("varianceClipGamma", C.c_float), ("depthThreshold", C.c_float),
        ("normalThreshold", C.c_float), ("phiColor", C.c_float), ("phiNormal", C.c_float), ("phiDepth", C.c_float),
        ("atrousIteration", C.c_int32), ("stepSize", C.c_int32), ("varianceBoost", C.c_float), ("minHistoryLength", C.c_int32),
        ("jitterOffset", C.c_float * 2), ("_padding", C.c_float * 2)]

    @staticmethod
    def from_bytes(b):
        b = bytes(np.asarray(b, dtype=np.uint8).tobytes()) if not isinstance(b, (bytes, bytearray)) else bytes(b)
        assert len(b) == 560
        return FrameUBO.from_buffer_copy(b)

    def to_numpy(self):
        return np.frombuffer(bytes(self), dtype=np.uint8).copy()


assert C.sizeof(FrameUBO) == 560


class PostSettings(C.Structure):
    """blokb_post_settings: Denoiser::Settings::atrousIterations (renderer_denoising.hpp:71) + PostProcess::Settings
    (renderer_postprocess.hpp:102-119); defaults are the reference's."""
    _fields_ = [("atrousIterations", C.c_int32), ("enableTAA", C.c_int32), ("feedbackMin", C.c_float), ("feedbackMax", C.c_float),
                ("enableTonemapping", C.c_int32), ("exposure", C.c_float), ("saturationBoost", C.c_float), ("tonemapOperator", C.c_int32),
                ("whitePoint", C.c_float), ("enableSharpening", C.c_int32), ("sharpenStrength", C.c_float)]

    def __init__(self, **kw):
        super().__init__(4, 1, 0.93, 0.98, 1, 1.0, 1.15, 1, 0.0, 1, 0.5)
        for k, v in kw.items():
            assert k in dict(self._fields_), k
            setattr(self, k, v)


PATH_WAVEFRONT, PATH_FUSED = 0, 1


class PathOptions(C.Structure):
    """blokb_path_options: how render / trace_path schedule a sample's rays (wavefront with sorted bounce rays, or fused)."""
    _fields_ = [("mode", C.c_int32), ("batchSamples", C.c_uint32), ("dirCells", C.c_uint32), ("originCellShift", C.c_int32),
                ("mortonBitsPerAxis", C.c_uint32), ("refillBelow", C.c_uint32)]


class DenseChunkC(C.Structure):
    """blokb_dense_chunk: one blok::Chunk (chunk.hpp:33-42) — index + its density / materialIds arrays (128^3 each)."""
    _fields_ = [("cx", C.c_int32), ("cy", C.c_int32), ("cz", C.c_int32), ("density", _vp), ("materialIds", _vp)]


class CameraC(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("yaw", C.c_float), ("pitch", C.c_float), ("fov", C.c_float)]


class Buffers(C.Structure):
    _fields_ = [("hit", _vp), ("shadow", _vp), ("color", _vp), ("worldPos", _vp), ("normalRough", _vp), ("albedoMetal", _vp),
                ("motion", _vp), ("accum", _vp), ("width", C.c_uint32), ("height", C.c_uint32), ("hitPacked", _vp)]


class WorldStats(C.Structure):
    _fields_ = [("nodes", C.c_uint64), ("subChunks", C.c_uint64), ("materials", C.c_uint64), ("bricks", C.c_uint64),
                ("blocks", C.c_uint64), ("voxels", C.c_uint64), ("compactBytes", C.c_uint64), ("gridOrigin", C.c_int32 * 3),
                ("gridDim", C.c_int32 * 3), ("levels", C.c_int32), ("buildMs", C.c_float), ("uploadMs", C.c_float)]

    def as_dict(self):
        d = {}
        for n, _ in self._fields_:
            v = getattr(self, n)
            d[n] = list(v) if hasattr(v, "__len__") else v
        return d


# Every symbol include/blokb200.h declares (tests check that the library exports each one).
EXPORTS = ["blokb_last_error", "blokb_create", "blokb_destroy", "blokb_upload_world", "blokb_world_stats_get", "blokb_set_frame",
           "blokb_camera_frame", "blokb_camera_frame_seq", "blokb_trace_primary", "blokb_trace_primary_shadow", "blokb_render", "blokb_trace_path",
           "blokb_resolve", "blokb_reset_accum", "blokb_trace_rays", "blokb_trace_rays_host", "blokb_get_buffers",
           "blokb_readback", "blokb_sync", "blokb_launch_count", "blokb_stats_enable", "blokb_stats_read", "blokb_stream",
           "blokb_worldgen_pack", "blokb_worldgen_counts", "blokb_worldgen_copy", "blokb_worldgen_nodes", "blokb_worldgen_subs",
           "blokb_worldgen_free", "blokb_pipeline_create", "blokb_pipeline_submit", "blokb_pipeline_wait", "blokb_pipeline_drain",
           "blokb_pipeline_destroy", "blokb_host_alloc", "blokb_host_free", "blokb_post_settings_default", "blokb_denoise",
           "blokb_postprocess", "blokb_end_frame", "blokb_reset_history", "blokb_write_image", "blokb_upload_voxels", "blokb_upload_dense", "blokb_world_export", "blokb_vox_load", "blokb_vox_parse",
           "blokb_vox_counts", "blokb_vox_xyz", "blokb_vox_material_ids", "blokb_vox_materials", "blokb_vox_free", "blokb_vox_last_error", "blokb_upload_vox", "blokb_set_strips", "blokb_ipc_export", "blokb_set_peers", "blokb_bind_image",
           "blokb_world_save", "blokb_world_load", "blokb_update_dense", "blokb_trace_rays_reordered", "blokb_trace_rays_reordered_static",
           "blokb_pipeline_create_packed", "blokb_pipeline_submit_packed", "blokb_node_prims", "blokb_unpack_hits",
           "blokb_path_options_default", "blokb_set_path_options", "blokb_get_path_options", "blokb_trace_primary_packed",
           "blokb_comm_unique_id", "blokb_comm_init", "blokb_set_comm", "blokb_comm_rank", "blokb_reduce_accum", "blokb_reduce_slice",
           "blokb_readback_slice", "blokb_readback_strips", "blokb_bind_fence", "blokb_frame_fence", "blokb_fence_status",
           "blokb_bind_accum_for_reduce"]

_lib = None
_host = None
HOST_LIB_PATH = os.path.join(HERE, "libblokbhost.so")


def _bind_host(L):
    """Prototypes of the host-only helpers (present in libblokb200.so and, without any CUDA, in libblokbhost.so)."""
    L.blokb_vox_load.argtypes = [C.c_char_p, C.c_uint32, _vp, C.POINTER(_vp)]
    L.blokb_vox_parse.argtypes = [_vp, C.c_size_t, C.c_uint32, _vp, C.POINTER(_vp)]
    L.blokb_vox_counts.argtypes = [_vp, _vp, _vp, _vp, _vp]
    L.blokb_vox_counts.restype = None
    for f, t in (("blokb_vox_xyz", _vp), ("blokb_vox_material_ids", _vp), ("blokb_vox_materials", _vp), ("blokb_vox_last_error", C.c_char_p)):
        getattr(L, f).restype = t
    for f in ("blokb_vox_xyz", "blokb_vox_material_ids", "blokb_vox_materials", "blokb_vox_free"):
        getattr(L, f).argtypes = [_vp]
    L.blokb_vox_free.restype = None
    L.blokb_camera_frame.argtypes = [C.POINTER(CameraC), C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(FrameUBO)]
    L.blokb_camera_frame_seq.argtypes = [C.POINTER(CameraC), C.POINTER(CameraC), C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int32,
                                         C.POINTER(FrameUBO)]
    L.blokb_worldgen_pack.argtypes = [_vp, _vp, C.c_size_t]
    L.blokb_worldgen_pack.restype = _vp
    L.blokb_worldgen_counts.argtypes = [_vp, _vp, _vp]
    L.blokb_worldgen_counts.restype = None
    L.blokb_worldgen_copy.argtypes = [_vp, _vp, _vp]
    L.blokb_worldgen_copy.restype = None
    L.blokb_worldgen_free.argtypes = [_vp]
    L.blokb_worldgen_free.restype = None


def load_host_library():
    """The host-only helpers (scene generator, Camera -> FrameUBO, .vox import). Served by libblokb200.so when it is already loaded
    (or when libblokbhost.so is absent); otherwise by libblokbhost.so — same sources compiled without CUDA — so that a process
    that only prepares inputs (bench.py --impl reference, the CPU tests) never maps the product's CUDA binary."""
    global _host
    if _lib is not None:
        return _lib
    if _host is not None:
        return _host
    if not os.path.exists(HOST_LIB_PATH):
        return load_library()
    L = C.CDLL(HOST_LIB_PATH)
    _bind_host(L)
    _host = L
    return L


class _Missing:
    """Stand-in for an entry point an experimental library does not export: prototypes can be assigned, calling it raises."""

    def __init__(self, name):
        object.__setattr__(self, "_name", name)

    def __setattr__(self, k, v):
        pass

    def __call__(self, *a):
        raise BlokbError(-3, f"{LIB_PATH} does not export {self._name}")


class _Tolerant:
    def __init__(self, lib):
        object.__setattr__(self, "_lib", lib)

    def __getattr__(self, name):
        try:
            return getattr(self._lib, name)
        except AttributeError:
            m = _Missing(name)
            object.__setattr__(self, name, m)
            return m


def load_library():
    """Load libblokb200.so (built in-tree by blok_b200/csrc/Makefile). Raises if it is missing: no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} not built — run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "or `make -C blok_b200/csrc`; blok_b200 has no CPU fallback")
    L = C.CDLL(LIB_PATH)
    if os.environ.get("BLOKB_LIB_PATH"):
        L = _Tolerant(L)   # a variant or older build named explicitly (kernel experiments): symbols it lacks fail when called, not at load
    L.blokb_last_error.restype = C.c_char_p
    L.blokb_create.argtypes = [C.c_int, C.POINTER(_vp)]
    L.blokb_destroy.argtypes = [_vp]
    L.blokb_destroy.restype = None
    L.blokb_upload_world.argtypes = [_vp, _vp, C.c_size_t, _vp, C.c_size_t, _vp, C.c_size_t]
    L.blokb_world_stats_get.argtypes = [_vp, C.POINTER(WorldStats)]
    L.blokb_set_frame.argtypes = [_vp, C.POINTER(FrameUBO)]
    L.blokb_upload_voxels.argtypes = [_vp, _vp, _vp, C.c_size_t, _vp, C.c_size_t]
    L.blokb_world_export.argtypes = [_vp, C.c_int, _vp, C.c_size_t, C.POINTER(C.c_size_t)]
    L.blokb_upload_dense.argtypes = [_vp, _vp, C.c_size_t, _vp, C.c_size_t]
    _bind_host(L)
    L.blokb_upload_vox.argtypes = [_vp, C.c_char_p, C.c_uint32, _vp]
    L.blokb_comm_unique_id.argtypes = [_vp]
    L.blokb_comm_init.argtypes = [_vp, _vp, C.c_int, C.c_int]
    L.blokb_set_comm.argtypes = [_vp, _vp]
    L.blokb_comm_rank.argtypes = [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.blokb_reduce_accum.argtypes = [_vp, C.c_int, _vp]
    L.blokb_reduce_slice.argtypes = [_vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.blokb_readback_slice.argtypes = [_vp, C.c_int, _vp, C.c_size_t]
    L.blokb_readback_strips.argtypes = [_vp, C.c_int, _vp, C.c_size_t]
    L.blokb_bind_fence.argtypes = [_vp, _vp, _vp, C.c_uint32]
    L.blokb_bind_accum_for_reduce.argtypes = [_vp, _vp, _vp]
    L.blokb_frame_fence.argtypes = [_vp, _vp]
    L.blokb_fence_status.argtypes = [_vp]
    for f in ("blokb_trace_primary", "blokb_trace_primary_shadow", "blokb_trace_primary_packed", "blokb_render", "blokb_resolve", "blokb_reset_accum"):
        getattr(L, f).argtypes = [_vp, _vp]
    L.blokb_trace_path.argtypes = [_vp, C.c_uint32, C.c_uint32, _vp]
    L.blokb_set_strips.argtypes = [_vp, C.c_uint32, C.c_uint32]
    L.blokb_path_options_default.argtypes = [C.POINTER(PathOptions)]
    L.blokb_path_options_default.restype = None
    L.blokb_set_path_options.argtypes = [_vp, C.POINTER(PathOptions)]
    L.blokb_get_path_options.argtypes = [_vp, C.POINTER(PathOptions)]
    L.blokb_ipc_export.argtypes = [_vp, C.c_int, _vp]
    L.blokb_world_save.argtypes = [_vp, C.c_char_p]
    L.blokb_update_dense.argtypes = [_vp, _vp, C.c_size_t]
    L.blokb_world_load.argtypes = [_vp, C.c_char_p]
    L.blokb_bind_image.argtypes = [_vp, C.c_int, _vp, _vp]
    L.blokb_set_peers.argtypes = [_vp, C.c_int, _vp, C.c_uint32, C.c_uint32]
    L.blokb_trace_rays.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp, _vp]
    L.blokb_trace_rays_reordered.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp, _vp]
    L.blokb_trace_rays_reordered_static.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp, _vp]
    L.blokb_trace_rays_host.argtypes = [_vp, _vp, C.c_size_t, C.c_int, _vp]
    L.blokb_get_buffers.argtypes = [_vp, C.POINTER(Buffers)]
    L.blokb_readback.argtypes = [_vp, C.c_int, _vp, C.c_size_t]
    L.blokb_sync.argtypes = [_vp]
    L.blokb_post_settings_default.argtypes = [C.POINTER(PostSettings)]
    L.blokb_post_settings_default.restype = None
    L.blokb_denoise.argtypes = [_vp, C.POINTER(PostSettings), _vp]
    L.blokb_postprocess.argtypes = [_vp, C.POINTER(PostSettings), _vp]
    L.blokb_end_frame.argtypes = [_vp, _vp]
    L.blokb_reset_history.argtypes = [_vp, _vp]
    L.blokb_write_image.argtypes = [_vp, C.c_int, _vp, C.c_size_t]
    L.blokb_launch_count.argtypes = [_vp]
    L.blokb_launch_count.restype = C.c_uint64
    L.blokb_stats_enable.argtypes = [_vp, C.c_int]
    L.blokb_stats_read.argtypes = [_vp, _vp]
    L.blokb_stream.argtypes = [_vp]
    L.blokb_stream.restype = _vp
    L.blokb_pipeline_create.argtypes = [_vp, C.c_int, C.c_uint32, C.c_uint32, C.POINTER(_vp)]
    L.blokb_pipeline_submit.argtypes = [_vp, C.POINTER(FrameUBO), _vp, C.POINTER(C.c_uint64)]
    L.blokb_pipeline_create_packed.argtypes = L.blokb_pipeline_create.argtypes
    L.blokb_pipeline_submit_packed.argtypes = L.blokb_pipeline_submit.argtypes
    L.blokb_node_prims.argtypes = [_vp, C.c_size_t, _vp, C.c_size_t, _vp]
    L.blokb_unpack_hits.argtypes = [_vp, C.c_size_t, _vp, C.c_size_t, _vp, _vp]
    L.blokb_pipeline_wait.argtypes = [_vp, C.c_uint64]
    L.blokb_pipeline_drain.argtypes = [_vp]
    L.blokb_pipeline_destroy.argtypes = [_vp]
    L.blokb_pipeline_destroy.restype = None
    L.blokb_host_alloc.argtypes = [C.c_size_t]
    L.blokb_host_alloc.restype = _vp
    L.blokb_host_free.argtypes = [_vp]
    L.blokb_host_free.restype = None
    _lib = L
    return L


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_vp)


def _ck(rc):
    if rc != 0:
        raise BlokbError(rc, load_library().blokb_last_error().decode())


class WorldSvoGpu:
    """blok::WorldSvoGpu (resources.hpp:195-210): the three host arrays the reference uploads."""

    def __init__(self, global_nodes, global_sub_chunks, materials):
        self.globalNodes = np.ascontiguousarray(global_nodes, dtype=NODE_DT)
        self.globalSubChunks = np.ascontiguousarray(global_sub_chunks, dtype=SUB_DT)
        self.materials = np.ascontiguousarray(materials, dtype=MAT_DT)

    @staticmethod
    def from_voxels(xyz, material_ids, materials):
        """ChunkManager(128,1).setVoxelMaterial per voxel + rebuildDirtyChunks + packChunksToGpuSvo (host, C++)."""
        L = load_host_library()
        xyz = np.ascontiguousarray(xyz, np.int32).reshape(-1, 3)
        mid = np.ascontiguousarray(material_ids, np.uint32)
        assert len(mid) == len(xyz)
        h = L.blokb_worldgen_pack(_ptr(xyz), _ptr(mid), len(mid))
        if not h:
            raise MemoryError("blokb_worldgen_pack")
        try:
            n = (C.c_uint64 * 2)()
            L.blokb_worldgen_counts(h, C.byref(n, 0), C.byref(n, 8))
            nodes = np.empty(n[0], NODE_DT)
            subs = np.empty(n[1], SUB_DT)
            L.blokb_worldgen_copy(h, _ptr(nodes), _ptr(subs))
        finally:
            L.blokb_worldgen_free(h)
        return WorldSvoGpu(nodes, subs, materials)

    def node_prims(self) -> np.ndarray:
        """primOfNode (blokb_node_prims): the sub-chunk (gl_PrimitiveID) whose subtree holds each node; cached."""
        if getattr(self, "_prims", None) is None:
            L = load_library()
            prims = np.empty(len(self.globalNodes), np.uint32)
            _ck(L.blokb_node_prims(_ptr(self.globalNodes), len(self.globalNodes), _ptr(self.globalSubChunks), len(self.globalSubChunks),
                                   _ptr(prims)))
            self._prims = prims
        return self._prims

    def unpack_hits(self, packed: np.ndarray, with_prim: bool = True) -> np.ndarray:
        """blokb_unpack_hits: PACKED_DT records -> HIT_DT records (material and gl_PrimitiveID looked up in this world's arrays)."""
        L = load_library()
        src = np.ascontiguousarray(packed, PACKED_DT)
        out = np.empty(src.shape, HIT_DT)
        prims = self.node_prims() if with_prim else None
        _ck(L.blokb_unpack_hits(_ptr(src), src.size, _ptr(self.globalNodes), len(self.globalNodes), _ptr(prims) if with_prim else None, _ptr(out)))
        return out


def load_vox(source, model_index=0, offset=(0.0, 0.0, 0.0)):
    """loadVoxFile + importVoxMaterials + importVoxToChunks (blok/src/vox_loader.cpp) as data: returns (xyz (n,3) i32, material ids
    (n,) u32, materials (256,) MAT_DT, info dict). source: a path or the file's bytes. Host only."""
    L = load_host_library()
    off = (C.c_float * 3)(*[float(v) for v in offset])
    h = _vp()
    if isinstance(source, (bytes, bytearray)):
        buf = (C.c_char * len(source)).from_buffer_copy(bytes(source))
        rc = L.blokb_vox_parse(buf, len(source), model_index, off, C.byref(h))
    else:
        rc = L.blokb_vox_load(str(source).encode(), model_index, off, C.byref(h))
    if rc != 0:
        raise BlokbError(rc, L.blokb_vox_last_error().decode())
    try:
        n, nm = C.c_uint64(), C.c_uint64()
        nmod, size = C.c_uint32(), (C.c_uint32 * 3)()
        L.blokb_vox_counts(h, C.byref(n), C.byref(nm), C.byref(nmod), size)
        xyz = np.ctypeslib.as_array(C.cast(L.blokb_vox_xyz(h), C.POINTER(C.c_int32)), (n.value, 3)).copy() if n.value else np.zeros((0, 3), np.int32)
        ids = np.ctypeslib.as_array(C.cast(L.blokb_vox_material_ids(h), C.POINTER(C.c_uint32)), (n.value,)).copy() if n.value else np.zeros(0, np.uint32)
        mats = np.frombuffer(C.string_at(L.blokb_vox_materials(h), nm.value * 32), dtype=MAT_DT).copy()
    finally:
        L.blokb_vox_free(h)
    return xyz, ids, mats, {"models": int(nmod.value), "size": tuple(int(v) for v in size)}


class Camera:
    """blok::Camera (camera.hpp:15-59)."""

    def __init__(self, position=(0.0, 10.0, -5.0), yaw=0.0, pitch=0.0, fov=60.0):
        self.position, self.yaw, self.pitch, self.fov = tuple(float(v) for v in position), float(yaw), float(pitch), float(fov)

    def _c(self):
        return CameraC((C.c_float * 3)(*self.position), self.yaw, self.pitch, self.fov)

    def frame(self, width, height, frame_count=0, sample_count=8, prev: "Camera" = None, jitter_index=-1) -> FrameUBO:
        """view()/projection() + fillFrameUBO; sample_count defaults to the reference's 8 (renderer_denoising.cpp:683).
        prev = the previous frame's camera (prev* matrices for reprojection), jitter_index >= 0 = TAA Halton jitter."""
        c = self._c()
        out = FrameUBO()
        L = load_host_library()
        if prev is None and jitter_index < 0:
            rc = L.blokb_camera_frame(C.byref(c), width, height, frame_count, sample_count, C.byref(out))
        else:
            pc = prev._c() if prev is not None else None
            rc = L.blokb_camera_frame_seq(C.byref(c), C.byref(pc) if pc is not None else None, width, height, frame_count,
                                          sample_count, jitter_index, C.byref(out))
        if rc != 0:
            raise BlokbError(rc, "blokb_camera_frame: bad argument")
        return out


class Renderer:
    """Headless stand-in for blok::Renderer's ray-tracing path on one GPU (one context per device)."""

    def __init__(self, width=1280, height=720, device=0):  # app.cpp:95 default size
        self.L = load_library()
        self.width, self.height, self.device = width, height, device
        self._ctx = _vp()
        _ck(self.L.blokb_create(device, C.byref(self._ctx)))
        self._frame = None

    def close(self):
        if self._ctx:
            self.L.blokb_destroy(self._ctx)
            self._ctx = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- scene in
    def add_world(self, world: WorldSvoGpu):
        _ck(self.L.blokb_upload_world(self._ctx, _ptr(world.globalNodes), len(world.globalNodes), _ptr(world.globalSubChunks),
                                      len(world.globalSubChunks), _ptr(world.materials), len(world.materials)))

    update_world = add_world

    def add_voxels(self, xyz, material_ids, materials):
        """Scene in, built on the GPU: the voxel list (what ChunkManager::setVoxelMaterial would be fed) goes to the device as
        is; equivalent to WorldSvoGpu.from_voxels(...) + add_world(...) with the same node numbering."""
        xyz = np.ascontiguousarray(xyz, np.int32).reshape(-1, 3)
        mid = np.ascontiguousarray(material_ids, np.uint32)
        mats = np.ascontiguousarray(materials, dtype=MAT_DT)
        assert len(mid) == len(xyz)
        _ck(self.L.blokb_upload_voxels(self._ctx, _ptr(xyz), _ptr(mid), len(mid), _ptr(mats), len(mats)))

    def save_world(self, path):
        """Scene dump: the device-resident compact world to one file (blokb_world_save)."""
        _ck(self.L.blokb_world_save(self._ctx, os.fspath(path).encode()))

    def load_world(self, path):
        """File -> device, no builder involved (blokb_world_load); traces exactly like the world that was saved."""
        _ck(self.L.blokb_world_load(self._ctx, os.fspath(path).encode()))

    @staticmethod
    def _dense_array(chunks):
        keep = []
        arr = (DenseChunkC * max(len(chunks), 1))()
        for k, ((cx, cy, cz), (den, ids)) in enumerate(chunks.items()):
            den = np.ascontiguousarray(den, np.float32).reshape(-1)
            ids = np.ascontiguousarray(ids, np.uint32).reshape(-1)
            assert den.size == 128 ** 3 and ids.size == 128 ** 3
            keep += [den, ids]
            arr[k] = DenseChunkC(cx, cy, cz, den.ctypes.data, ids.ctypes.data)
        return arr, keep

    def add_dense(self, chunks, materials):
        """chunks: {(cx, cy, cz): (density f32[128^3], materialIds u32[128^3])} — ChunkManager::chunks as the reference holds them
        (index x + 128 y + 16384 z). rebuildDirtyChunks + packChunksToGpuSvo + addWorld, on the device."""
        arr, keep = self._dense_array(chunks)
        mats = np.ascontiguousarray(materials, dtype=MAT_DT)
        _ck(self.L.blokb_upload_dense(self._ctx, C.cast(arr, _vp), len(chunks), _ptr(mats), len(mats)))

    def update_dense(self, dirty_chunks):
        """Renderer::updateWorld after an edit: only the dirty chunks (same dict form as add_dense) travel; the world is rebuilt on
        the device from the voxel list kept since add_dense. Equal to add_dense of all chunks."""
        arr, keep = self._dense_array(dirty_chunks)
        _ck(self.L.blokb_update_dense(self._ctx, C.cast(arr, _vp), len(dirty_chunks)))

    def add_vox(self, path, model_index=0, offset=(0.0, 0.0, 0.0)):
        """loadAndImportVox + rebuild + pack + addWorld (blok/src/app.cpp:100-126): the file's voxels go straight to the device builder."""
        off = (C.c_float * 3)(*[float(v) for v in offset])
        _ck(self.L.blokb_upload_vox(self._ctx, str(path).encode(), model_index, off))

    def world_export(self, which: int) -> np.ndarray:
        """Device-resident world array -> numpy (WORLD_NODES: (n,) {mask,base,aux}; WORLD_BLOCKFIRSTLEAF: u32; WORLD_VOXELMAT: raw
        bytes as u8; WORLD_SUBCHUNKS: SUB_DT)."""
        n = C.c_size_t()
        _ck(self.L.blokb_world_export(self._ctx, which, None, 0, C.byref(n)))
        dt = {WORLD_NODES: CNODE_DT, WORLD_BLOCKFIRSTLEAF: np.dtype("<u4"), WORLD_VOXELMAT: np.dtype("u1"), WORLD_SUBCHUNKS: SUB_DT}[which]
        out = np.empty(n.value // dt.itemsize, dt)
        _ck(self.L.blokb_world_export(self._ctx, which, _ptr(out), out.nbytes, None))
        return out

    def world_stats(self) -> dict:
        s = WorldStats()
        _ck(self.L.blokb_world_stats_get(self._ctx, C.byref(s)))
        return s.as_dict()

    # ---- frame in
    def set_frame(self, frame):
        if isinstance(frame, Camera):
            frame = frame.frame(self.width, self.height)
        elif not isinstance(frame, FrameUBO):
            frame = FrameUBO.from_bytes(frame)
        self._frame = frame
        self.width, self.height = frame.screenWidth, frame.screenHeight
        _ck(self.L.blokb_set_frame(self._ctx, C.byref(frame)))

    # ---- dispatch (asynchronous; stream = cudaStream_t as int, None = the context's stream)
    def render(self, frame=None, stream=None):
        """One frame of raygen.rgen: frame.sampleCount spp -> the five G-buffer images."""
        if frame is not None:
            self.set_frame(frame)
        _ck(self.L.blokb_render(self._ctx, stream))

    def set_strips(self, first=0, stride=1):
        """Tile scheduler: later dispatches render the frame's 8-row strips first, first + stride, ... only (this GPU's
        share of ONE frame, sharding.strip_set); (0, 1) = the whole frame."""
        _ck(self.L.blokb_set_strips(self._ctx, first, stride))

    def bind_image(self, image: int, local_ptr=None, multicast_ptr=None, reduce_only=False):
        """Place IMG_HIT / IMG_SHADOW / IMG_ACCUM in caller-owned (symmetric) memory; with multicast_ptr the kernels store each
        record once to the NVLink multicast address (sharding.connect_multicast). bind_image(image) restores the own buffer.
        reduce_only (IMG_ACCUM of sample-sharded frames): dispatches keep storing locally, the multicast address only serves
        reduce_accum's in-switch sum (blokb_bind_accum_for_reduce)."""
        if reduce_only:
            assert image == IMG_ACCUM
            _ck(self.L.blokb_bind_accum_for_reduce(self._ctx, local_ptr or None, multicast_ptr or None))
        else:
            _ck(self.L.blokb_bind_image(self._ctx, image, local_ptr or None, multicast_ptr or None))

    def ipc_export(self, image: int) -> bytes:
        """64-byte CUDA IPC handle of one of this context's exchangeable images (IMG_HIT, IMG_SHADOW, IMG_ACCUM)."""
        h = C.create_string_buffer(64)
        _ck(self.L.blokb_ipc_export(self._ctx, image, h))
        return h.raw

    def set_peers(self, image: int, handles, self_index: int):
        """handles: the ranks' ipc_export(image) results in rank order (this rank's own entry is ignored). Later dispatches
        store `image` into every rank's buffer (fused exchange); handles=None unmaps."""
        if not handles:
            _ck(self.L.blokb_set_peers(self._ctx, image, None, 0, 0))
            return
        blob = b"".join(handles)
        assert len(blob) == 64 * len(handles)
        _ck(self.L.blokb_set_peers(self._ctx, image, blob, len(handles), self_index))

    def trace_primary(self, frame=None, shadow=False, stream=None, packed=False):
        """Sample-0 primary rays -> IMG_HIT (+ IMG_SHADOW with shadow=True), or 8-byte records -> IMG_HIT_PACKED with packed=True."""
        if frame is not None:
            self.set_frame(frame)
        fn = self.L.blokb_trace_primary_packed if packed else (self.L.blokb_trace_primary_shadow if shadow else self.L.blokb_trace_primary)
        _ck(fn(self._ctx, stream))

    # ---- cross-GPU reduction of the accumulation buffers (sample-sharded path tracing)
    @staticmethod
    def comm_unique_id() -> bytes:
        """128-byte ncclUniqueId (rank 0 makes it and ships it to the other ranks)."""
        b = C.create_string_buffer(128)
        _ck(load_library().blokb_comm_unique_id(b))
        return b.raw

    def comm_init(self, unique_id: bytes, rank: int, nranks: int):
        """Library-owned NCCL communicator for this context (collective: every rank calls it with the same id)."""
        assert len(unique_id) == 128
        _ck(self.L.blokb_comm_init(self._ctx, unique_id, rank, nranks))

    def set_comm(self, nccl_comm_ptr: int):
        """Adopt a caller-owned ncclComm_t (0 / None detaches)."""
        _ck(self.L.blokb_set_comm(self._ctx, nccl_comm_ptr or None))

    def comm_rank(self):
        r, n = C.c_int(), C.c_int()
        _ck(self.L.blokb_comm_rank(self._ctx, C.byref(r), C.byref(n)))
        return r.value, n.value

    def reduce_accum(self, mode=REDUCE_ALL, stream=None):
        """colour = resolve(sum over GPUs of accum): REDUCE_SLICE leaves each rank with its slice of pixels, REDUCE_ALL with the frame."""
        _ck(self.L.blokb_reduce_accum(self._ctx, mode, stream))

    def reduce_slice(self):
        """(first pixel, pixel count) of the slice of the frame (row-major) this rank owns after reduce_accum."""
        a, b = C.c_uint64(), C.c_uint64()
        _ck(self.L.blokb_reduce_slice(self._ctx, C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)

    def readback_slice(self, image: int = IMG_COLOR) -> np.ndarray:
        """This rank's slice of an rgba32f image as (count, 4) float32."""
        first, count = self.reduce_slice()
        out = np.empty((count, 4), np.float32)
        _ck(self.L.blokb_readback_slice(self._ctx, image, _ptr(out), out.nbytes))
        return out

    def readback_slice_into(self, image: int, host_ptr: int, nbytes: int):
        _ck(self.L.blokb_readback_slice(self._ctx, image, host_ptr, nbytes))

    def bind_fence(self, local_ptr=None, multicast_ptr=None, nranks=0):
        """Frame fence over a multicast word pair in symmetric memory (sharding.connect_fence); nranks = 0 unbinds."""
        _ck(self.L.blokb_bind_fence(self._ctx, local_ptr or None, multicast_ptr or None, nranks))

    def frame_fence(self, stream=None):
        """Enqueue the cross-GPU fence: completes when every rank's earlier work on its fence stream has been performed."""
        _ck(self.L.blokb_frame_fence(self._ctx, stream))

    def fence_status(self):
        _ck(self.L.blokb_fence_status(self._ctx))

    def readback_strips_into(self, image: int, host_ptr: int, nbytes: int):
        """Only this context's strips (set_strips) of an image, copied to their full-frame positions in a whole-image host buffer."""
        _ck(self.L.blokb_readback_strips(self._ctx, image, host_ptr, nbytes))

    def set_path_options(self, mode=None, batch_samples=None, dir_cells=None, origin_cell_shift=None, morton_bits_per_axis=None,
                         refill_below=None):
        """blokb_set_path_options; arguments left None keep their current value. Returns the options now in force."""
        o = PathOptions()
        _ck(self.L.blokb_get_path_options(self._ctx, C.byref(o)))
        for name, v in (("mode", mode), ("batchSamples", batch_samples), ("dirCells", dir_cells), ("originCellShift", origin_cell_shift),
                        ("mortonBitsPerAxis", morton_bits_per_axis), ("refillBelow", refill_below)):
            if v is not None:
                setattr(o, name, int(v))
        _ck(self.L.blokb_set_path_options(self._ctx, C.byref(o)))
        _ck(self.L.blokb_get_path_options(self._ctx, C.byref(o)))
        return {n: getattr(o, n) for n, _ in PathOptions._fields_}

    def trace_path(self, sample_begin, sample_end, stream=None):
        _ck(self.L.blokb_trace_path(self._ctx, sample_begin, sample_end, stream))

    def resolve(self, stream=None):
        _ck(self.L.blokb_resolve(self._ctx, stream))

    def reset_accum(self, stream=None):
        _ck(self.L.blokb_reset_accum(self._ctx, stream))

    # ---- denoiser + post-process (enqueue only)
    def denoise(self, settings: "PostSettings" = None, stream=None):
        """Denoiser::denoise on the current G-buffer -> IMG_DENOISED (+ history / moments / variance)."""
        self._post = settings or getattr(self, "_post", None) or PostSettings()
        _ck(self.L.blokb_denoise(self._ctx, C.byref(self._post), stream))

    def postprocess(self, settings: "PostSettings" = None, stream=None):
        """PostProcess::process on the denoised image -> IMG_FINAL."""
        self._post = settings or getattr(self, "_post", None) or PostSettings()
        _ck(self.L.blokb_postprocess(self._ctx, C.byref(self._post), stream))

    def end_frame(self, stream=None):
        _ck(self.L.blokb_end_frame(self._ctx, stream))

    def reset_history(self, stream=None):
        _ck(self.L.blokb_reset_history(self._ctx, stream))

    def write_image(self, image: int, data: np.ndarray):
        """Host -> device copy into one of the G-buffer images (feeds denoise / postprocess from a G-buffer made elsewhere)."""
        a = np.ascontiguousarray(data)
        _ck(self.L.blokb_write_image(self._ctx, image, _ptr(a), a.nbytes))

    def trace_rays(self, rays: np.ndarray, any_hit=False) -> np.ndarray:
        """Host rays (n,8) f32 {o,d,tmin,tmax} -> (n,) HIT_DT, through the host-buffer C-ABI call."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        out = np.empty(len(rays), HIT_DT)
        _ck(self.L.blokb_trace_rays_host(self._ctx, _ptr(rays), len(rays), int(any_hit), _ptr(out)))
        return out

    def trace_rays_device(self, d_rays: int, n: int, d_out: int, any_hit=False, stream=None, reorder=False):
        """Device ray list -> device hit records. reorder=True sorts the list by direction cell + origin Morton code first
        (blokb_trace_rays_reordered): same records, faster for incoherent rays."""
        fn = self.L.blokb_trace_rays_reordered if reorder else self.L.blokb_trace_rays
        _ck(fn(self._ctx, d_rays, n, int(any_hit), d_out, stream))

    # ---- images out
    def buffers(self) -> Buffers:
        b = Buffers()
        _ck(self.L.blokb_get_buffers(self._ctx, C.byref(b)))
        return b

    def readback(self, image: int) -> np.ndarray:
        h, w = self.height, self.width
        shape_dt = {IMG_HIT: ((h, w), HIT_DT), IMG_HIT_PACKED: ((h, w), PACKED_DT), IMG_SHADOW: ((h, w), np.uint8), IMG_COLOR: ((h, w, 4), np.float32),
                    IMG_WORLDPOS: ((h, w, 4), np.float32), IMG_NORMALROUGH: ((h, w, 4), np.float16),
                    IMG_ALBEDOMETAL: ((h, w, 4), np.uint8), IMG_MOTION: ((h, w, 2), np.float16), IMG_ACCUM: ((h, w, 4), np.float32),
                    IMG_HISTORY_COLOR: ((h, w, 4), np.float32), IMG_MOMENTS: ((h, w, 2), np.float32), IMG_HISTORY_LENGTH: ((h, w), np.float16),
                    IMG_VARIANCE: ((h, w), np.float32), IMG_DENOISED: ((h, w, 4), np.float32), IMG_TAA_HISTORY: ((h, w, 4), np.float32),
                    IMG_TAA_OUTPUT: ((h, w, 4), np.float32), IMG_TONEMAPPED: ((h, w, 4), np.uint8), IMG_FINAL: ((h, w, 4), np.uint8)}
        shape, dt = shape_dt[image]
        if image == IMG_FINAL and getattr(self, "_post", None) is not None and not self._post.enableTonemapping:
            dt = np.float32
        out = np.empty(shape, dt)
        _ck(self.L.blokb_readback(self._ctx, image, _ptr(out), out.nbytes))
        return out

    def readback_into(self, image: int, host_ptr: int, nbytes: int):
        """Stream-ordered copy of one image into caller-owned host memory (pinned for full PCIe speed) + sync."""
        _ck(self.L.blokb_readback(self._ctx, image, host_ptr, nbytes))

    def device_array(self, image: int):
        """Zero-copy view of a context-owned device image as an object exposing __cuda_array_interface__ (for
        torch.as_tensor(..., device="cuda") — used to hand the accumulation buffer to torch.distributed / NCCL). The 16-B
        hit records are exposed as (h, w, 4) int32 words (NCCL and torch index ops have no uint32)."""
        b = self.buffers()
        h, w = int(b.height), int(b.width)
        table = {IMG_HIT: (b.hit, (h, w, 4), "<i4"), IMG_HIT_PACKED: (b.hitPacked, (h, w, 2), "<i4"), IMG_SHADOW: (b.shadow, (h, w), "|u1"), IMG_COLOR: (b.color, (h, w, 4), "<f4"),
                 IMG_WORLDPOS: (b.worldPos, (h, w, 4), "<f4"), IMG_NORMALROUGH: (b.normalRough, (h, w, 4), "<f2"),
                 IMG_ALBEDOMETAL: (b.albedoMetal, (h, w, 4), "|u1"), IMG_MOTION: (b.motion, (h, w, 2), "<f2"),
                 IMG_ACCUM: (b.accum, (h, w, 4), "<f4")}
        ptr, shape, typestr = table[image]

        class _View:
            __cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (int(ptr), False), "version": 2, "strides": None}
        v = _View()
        v._owner = self  # keep the context alive as long as the view is
        return v

    def sync(self):
        _ck(self.L.blokb_sync(self._ctx))

    @property
    def stream(self) -> int:
        return self.L.blokb_stream(self._ctx)

    def launch_count(self) -> int:
        return int(self.L.blokb_launch_count(self._ctx))

    def stats_enable(self, on=True):
        _ck(self.L.blokb_stats_enable(self._ctx, int(on)))

    def stats_read(self):
        v = np.zeros(2, np.uint64)
        _ck(self.L.blokb_stats_read(self._ctx, _ptr(v)))
        return int(v[0]), int(v[1])


class FramePipeline:
    """Frames in flight (the reference keeps 2, blok/include/renderer.hpp:185): submit(frame) enqueues the primary pass
    and the device->host copy of its hit records into one of `depth` pinned host images; the copy of frame i overlaps
    the tracing of frame i+1. result(ticket) waits for that frame's copy and returns its (h, w) HIT_DT view (valid until
    `depth` more frames have been submitted)."""

    def __init__(self, renderer: Renderer, width: int, height: int, depth: int = 2, packed: bool = False):
        self.L, self.r, self.width, self.height, self.depth = renderer.L, renderer, width, height, depth
        self._p = _vp()
        self.packed = packed  # 8-byte blokb_hit_packed records (PACKED_DT) instead of HIT_DT; WorldSvoGpu.unpack_hits restores the rest
        self.dtype = PACKED_DT if packed else HIT_DT
        create = self.L.blokb_pipeline_create_packed if packed else self.L.blokb_pipeline_create
        _ck(create(renderer._ctx, depth, width, height, C.byref(self._p)))
        self.nbytes = width * height * self.dtype.itemsize
        self._host = []
        for _ in range(depth):
            h = self.L.blokb_host_alloc(self.nbytes)
            if not h:
                raise MemoryError("blokb_host_alloc")
            self._host.append(h)

    def submit(self, frame) -> int:
        if isinstance(frame, Camera):
            frame = frame.frame(self.width, self.height)
        t = C.c_uint64()
        n = getattr(self, "_n", 0)
        submit = self.L.blokb_pipeline_submit_packed if self.packed else self.L.blokb_pipeline_submit
        _ck(submit(self._p, C.byref(frame), self._host[n % self.depth], C.byref(t)))
        self._n = n + 1
        return int(t.value)

    def result(self, ticket: int) -> np.ndarray:
        _ck(self.L.blokb_pipeline_wait(self._p, ticket))
        buf = (C.c_char * self.nbytes).from_address(self._host[ticket % self.depth])
        return np.frombuffer(buf, dtype=self.dtype).reshape(self.height, self.width)

    def drain(self):
        _ck(self.L.blokb_pipeline_drain(self._p))

    def close(self):
        if self._p:
            self.L.blokb_pipeline_destroy(self._p)
            self._p = _vp()
            for h in self._host:
                self.L.blokb_host_free(h)
            self._host = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
