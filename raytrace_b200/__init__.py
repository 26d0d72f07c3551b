"""raytrace_b200 — B200-native voxel ray-tracing hot path (libvxrt) and the host-side mirror of the
reference's `render` module interface.  Importing this package loads raytrace_b200/libvxrt.so; if the
CUDA library has not been built the import fails (there is no fallback)."""
from . import _native  # noqa: F401  (loads libvxrt.so, raises if missing)
from .render import (Camera, FramePipeline, Material, MATERIALS, Pipeline, Renderer, VxrtError, constants, create_instance,
                     compute_triple_euler_vector, fill_uniform_data, load_materials_csv, plan_bands, plan_bands_weighted, plan_frames, plan_passes)
from .game import Game
from .world import ChunkStorage, PackedChunkData
from .terrain_upload import TerrainUploadManager

__all__ = ["Camera", "FramePipeline", "Material", "MATERIALS", "Pipeline", "Renderer", "VxrtError", "constants", "create_instance",
           "compute_triple_euler_vector", "fill_uniform_data", "load_materials_csv", "plan_bands", "plan_bands_weighted", "plan_frames", "plan_passes", "Game", "ChunkStorage",
           "PackedChunkData", "TerrainUploadManager"]
