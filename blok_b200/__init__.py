"""blok_b200 — B200-native voxel ray traversal behind the reference's scene / camera / framebuffer seams.

Only what the hot path needs: csrc/ (CUDA kernels + C-ABI, built into libblokb200.so), api.py (ctypes mirror of the
reference's Renderer / Camera / WorldSvoGpu interface), scenes.py (deterministic synthetic worlds for bench and tests),
sharding.py (tile / sample partition across GPUs).
"""
from .api import (BlokbError, Camera, FramePipeline, FrameUBO, PostSettings, Renderer, load_vox, WorldSvoGpu, load_library, HIT_DT, MAT_DT, NODE_DT, PACKED_DT, SUB_DT,  # noqa: F401
                  IMG_ACCUM, IMG_ALBEDOMETAL, IMG_COLOR, IMG_HIT, IMG_MOTION, IMG_NORMALROUGH, IMG_SHADOW, IMG_WORLDPOS,
                  IMG_DENOISED, IMG_FINAL, IMG_HISTORY_COLOR, IMG_HISTORY_LENGTH, IMG_MOMENTS, IMG_TAA_HISTORY, IMG_TAA_OUTPUT, IMG_TONEMAPPED,
                  IMG_VARIANCE, IMG_HIT_PACKED, REDUCE_ALL, REDUCE_SLICE, PATH_FUSED, PATH_WAVEFRONT, PathOptions, WORLD_BLOCKFIRSTLEAF, WORLD_NODES, WORLD_SUBCHUNKS, WORLD_VOXELMAT)
