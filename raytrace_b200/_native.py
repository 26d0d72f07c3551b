"""ctypes binding of libvxrt.so — the C ABI declared in include/vxrt.h.

The library is the product: if it is missing this module raises at import (there is no CPU or
PyTorch fallback anywhere in the package).  Build it with ``python raytrace_b200/build.py`` or
``__graft_entry__.build()``.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# VXRT_LIB: another build of the same library (A/B runs of kernel variants, misc/r1_runs/run_k.sh); never a fallback
LIB_PATH = os.environ.get("VXRT_LIB") or os.path.join(_HERE, "libvxrt.so")

ABI_VERSION = 1

# VxrtStatus
OK, ERR_INVALID, ERR_CUDA, ERR_STATE, ERR_SIZE, ERR_COMM = 0, -1, -2, -3, -4, -5
# VxrtMode (truncations of raytrace.comp:main, :320-350)
MODE_PRIMARY, MODE_SHADOW, MODE_FULL = 0, 1, 2
# vxrt_create flags
FLAG_HIT_BUFFERS, FLAG_DENOISE, FLAG_FINALIZE, FLAG_ACCUMULATE, FLAG_PONG_SWAP, FLAG_GENERAL_MARCH, FLAG_REPROJECT = 1, 2, 4, 8, 16, 64, 128
FLAG_RAY_HANDOFF = 256
TRACE_AUTO, TRACE_SEQUENTIAL, TRACE_POOL, TRACE_WAVEFRONT = 0, 1, 2, 3   # vxrt_set_trace_form
# VxrtBufferId
(BUF_OUTPUT, BUF_LIGHTING, BUF_DEPTH, BUF_NORMAL, BUF_ALBEDO, BUF_EMISSION, BUF_FOG, BUF_HIT_VOXEL, BUF_HIT_FACE,
 BUF_HIT_WORD, BUF_HIT_T, BUF_STEPS, BUF_RAYS, BUF_ACCUM, BUF_LIGHTING_RAW, BUF_HIT_POS, BUF_HISTORY, BUF_COMPLETED) = range(18)


class VxrtConfig(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("width", C.c_int32), ("height", C.c_int32),
                ("nx", C.c_int32), ("ny", C.c_int32), ("nz", C.c_int32), ("mode", C.c_int32),
                ("flags", C.c_uint32), ("max_steps", C.c_int32), ("device", C.c_int32),
                ("row_begin", C.c_int32), ("row_end", C.c_int32), ("reserved", C.c_int32 * 4)]


class VxrtFrameParams(C.Structure):
    """Byte-compatible with RaytraceUniformData (src/render/pipeline/structs.rs:3-31), 192 bytes."""
    _fields_ = [("sun_angle", C.c_float), ("seed", C.c_uint32), ("_padding0", C.c_uint64),
                ("origin", C.c_float * 3), ("_padding1", C.c_uint32),
                ("forward", C.c_float * 3), ("_padding2", C.c_uint32),
                ("up", C.c_float * 3), ("_padding3", C.c_uint32),
                ("right", C.c_float * 3), ("_padding4", C.c_uint32),
                ("old_origin", C.c_float * 3), ("_padding5", C.c_uint32),
                ("old_transform_c0", C.c_float * 3), ("_padding6", C.c_uint32),
                ("old_transform_c1", C.c_float * 3), ("_padding7", C.c_uint32),
                ("old_transform_c2", C.c_float * 3), ("_padding8", C.c_uint32),
                ("region_offset", C.c_int32 * 3), ("_padding9", C.c_uint32),
                ("rotation", C.c_int32 * 3), ("_padding10", C.c_uint32),
                ("space_offset", C.c_int32 * 3), ("_padding11", C.c_uint32)]


assert C.sizeof(VxrtFrameParams) == 192


class VxrtEdit(C.Structure):
    _fields_ = [("x", C.c_int32), ("y", C.c_int32), ("z", C.c_int32), ("word", C.c_uint32)]


class VxrtFrameStats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("steps", C.c_uint64), ("hits", C.c_uint64), ("noise_fetches", C.c_uint64),
                ("limit_exhausted", C.c_uint64), ("trace_ms", C.c_float), ("denoise_ms", C.c_float),
                ("finalize_ms", C.c_float), ("total_ms", C.c_float)]


# every symbol include/vxrt.h declares: name -> (restype, argtypes)
_H = C.c_void_p
SYMBOLS = {
    "vxrt_create": (C.c_int, [C.POINTER(VxrtConfig), C.POINTER(_H)]),
    "vxrt_create_shared": (C.c_int, [C.POINTER(VxrtConfig), _H, C.POINTER(_H)]),
    "vxrt_destroy": (None, [_H]),
    "vxrt_last_error": (C.c_char_p, [_H]),
    "vxrt_abi_version": (C.c_int, []),
    "vxrt_upload_chunk": (C.c_int, [_H, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "vxrt_upload_world": (C.c_int, [_H, C.c_void_p, C.c_void_p]),
    "vxrt_upload_slice": (C.c_int, [_H, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "vxrt_generate_world": (C.c_int, [_H, C.c_int, C.c_uint32]),
    "vxrt_rebuild_minefield": (C.c_int, [_H]),
    "vxrt_read_world": (C.c_int, [_H, C.c_void_p, C.c_void_p]),
    "vxrt_validate_chunk": (C.c_int, [_H, C.c_int, C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_int64)]),
    "vxrt_upload_blue_noise": (C.c_int, [_H, C.c_void_p, C.c_size_t]),
    "vxrt_apply_edits": (C.c_int, [_H, C.c_void_p, C.c_size_t]),
    "vxrt_apply_edits_device": (C.c_int, [_H, C.c_void_p, C.c_size_t]),
    "vxrt_apply_edits_static": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]),
    "vxrt_set_frame": (C.c_int, [_H, C.c_void_p]),
    "vxrt_render_frame": (C.c_int, [_H]),
    "vxrt_sync": (C.c_int, [_H]),
    "vxrt_reset_accumulation": (C.c_int, [_H]),
    "vxrt_reset_history": (C.c_int, [_H]),
    "vxrt_read_output": (C.c_int, [_H, C.c_void_p, C.c_size_t]),
    "vxrt_read_output_async": (C.c_int, [_H, C.c_void_p, C.c_size_t, C.POINTER(C.c_uint64)]),
    "vxrt_wait_output": (C.c_int, [_H, C.c_uint64]),
    "vxrt_host_register": (C.c_int, [_H, C.c_void_p, C.c_size_t]),
    "vxrt_host_unregister": (C.c_int, [_H, C.c_void_p]),
    "vxrt_read_buffer": (C.c_int, [_H, C.c_int, C.c_void_p, C.c_size_t]),
    "vxrt_buffer_bytes": (C.c_size_t, [_H, C.c_int]),
    "vxrt_device_buffer": (C.c_void_p, [_H, C.c_int]),
    "vxrt_cuda_stream": (C.c_void_p, [_H]),
    "vxrt_frame_stats": (C.c_int, [_H, C.POINTER(VxrtFrameStats)]),
    "vxrt_launches_per_frame": (C.c_int, [_H]),
    "vxrt_set_trace_form": (C.c_int, [_H, C.c_int]),
    "vxrt_plan_bands": (C.c_int, [C.c_int, C.c_int, C.c_void_p]),
    "vxrt_plan_bands_weighted": (C.c_int, [C.c_int, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "vxrt_row_costs": (C.c_int, [_H, C.c_void_p, C.c_void_p, C.c_size_t]),
    "vxrt_plan_passes": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_void_p]),
    "vxrt_comm_unique_id": (C.c_int, [C.c_void_p, C.c_size_t]),
    "vxrt_comm_init": (C.c_int, [_H, C.c_void_p, C.c_size_t, C.c_int, C.c_int]),
    "vxrt_comm_gather_output": (C.c_int, [_H, C.c_int]),
    "vxrt_comm_broadcast_edits": (C.c_int, [_H, C.c_void_p, C.c_size_t, C.c_int]),
    "vxrt_chunk_file_name": (C.c_int, [C.c_int64, C.c_int64, C.c_int64, C.c_char_p, C.c_size_t]),
    "vxrt_chunk_encode_bound": (C.c_size_t, []),
    "vxrt_chunk_encode": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "vxrt_chunk_decode": (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]),
    "vxrt_peer_export": (C.c_int, [_H, C.c_void_p, C.c_size_t]),
    "vxrt_peer_barrier": (C.c_int, [_H]),
    "vxrt_peer_attach": (C.c_int, [_H, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_int]),
    "vxrt_selftest_division": (C.c_int, [C.c_int, C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64)]),
}
STATIC_MAX_EDITS, STATIC_MAX_CHUNKS = 16384, 32768
COMM_ID_BYTES = 128
PEER_BLOB_BYTES = 512


def load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: libvxrt (the CUDA implementation) has not been built and there is no "
            "fallback path.  Run `python raytrace_b200/build.py`.")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError here = header and library disagree
        fn.restype = res
        fn.argtypes = args
    if lib.vxrt_abi_version() != ABI_VERSION:
        raise ImportError("libvxrt ABI version mismatch")
    return lib


lib = load()
