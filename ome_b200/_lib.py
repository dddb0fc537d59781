"""ctypes declarations for libxetgpu.so — the C ABI in include/xetgpu.h.

Loading fails loudly: there is no CPU fallback in this package.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("XETGPU_LIBRARY") or os.path.join(_HERE, "libxetgpu.so")


class XetError(C.Structure):  # pkg/xet/xet.h:86-90
    _fields_ = [("code", C.c_int32), ("message", C.c_char_p), ("details", C.c_char_p)]


class XetProgressUpdate(C.Structure):  # pkg/xet/xet.h:21-30
    _fields_ = [("phase", C.c_int), ("total_bytes", C.c_uint64), ("completed_bytes", C.c_uint64),
                ("total_files", C.c_uint32), ("completed_files", C.c_uint32), ("current_file", C.c_char_p),
                ("current_file_completed_bytes", C.c_uint64), ("current_file_total_bytes", C.c_uint64)]


ProgressCallback = C.CFUNCTYPE(None, C.POINTER(XetProgressUpdate), C.c_void_p)
CancellationCallback = C.CFUNCTYPE(C.c_bool, C.c_void_p)


class XetCancellationToken(C.Structure):
    _fields_ = [("callback", CancellationCallback), ("user_data", C.c_void_p)]


class XetGpuConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("stream", C.c_void_p), ("staging_bytes", C.c_uint64),
                ("max_batch_bytes", C.c_uint64), ("flags", C.c_uint32)]


class XetGpuFileResult(C.Structure):
    _fields_ = [("file_hash", C.c_uint8 * 32), ("merkle_root", C.c_uint8 * 32), ("size", C.c_uint64),
                ("n_chunks", C.c_uint64), ("first_chunk", C.c_uint64)]


class XetGpuXorb(C.Structure):
    _fields_ = [("hash", C.c_uint8 * 32), ("file", C.c_uint64), ("first_chunk", C.c_uint64), ("n_chunks", C.c_uint64),
                ("bytes", C.c_uint64)]


class XetGpuResult(C.Structure):
    _fields_ = [("n_files", C.c_size_t), ("files", C.POINTER(XetGpuFileResult)), ("n_chunks", C.c_uint64),
                ("chunk_ends", C.POINTER(C.c_uint64)), ("chunk_hashes", C.POINTER(C.c_uint8)),
                ("n_xorbs", C.c_uint64), ("xorbs", C.POINTER(XetGpuXorb))]


class XetGpuRangeSpec(C.Structure):
    _fields_ = [("file_len", C.c_uint64), ("range_begin", C.c_uint64), ("range_end", C.c_uint64), ("entry_cut", C.c_uint64)]


class XetGpuTensorInfo(C.Structure):
    _fields_ = [("name", C.c_char_p), ("dtype", C.c_char_p), ("begin", C.c_uint64), ("end", C.c_uint64),
                ("n_elements", C.c_uint64), ("rank", C.c_uint32)]


class XetGpuSafetensorsIndex(C.Structure):
    _fields_ = [("file_size", C.c_uint64), ("header_len", C.c_uint64), ("data_start", C.c_uint64),
                ("n_tensors", C.c_uint64), ("tensors", C.POINTER(XetGpuTensorInfo)), ("parameter_count", C.c_int64)]


class XetGpuTimings(C.Structure):
    _fields_ = [("h2d_ms", C.c_float), ("gear_ms", C.c_float), ("resolve_ms", C.c_float),
                ("blake3_ms", C.c_float), ("merkle_ms", C.c_float), ("d2h_ms", C.c_float),
                ("total_ms", C.c_float), ("gear_kernel_ms", C.c_float),
                ("leaf_kernel_ms", C.c_float), ("bytes", C.c_uint64), ("n_chunks", C.c_uint64),
                ("n_candidates", C.c_uint64), ("kernel_launches", C.c_uint32)]


ERRP = C.POINTER(XetError)
CTX = C.c_void_p

# name -> (restype, argtypes); mirrors include/xetgpu.h one to one
SIGNATURES = {
    "xetgpu_version_0_2_0": (None, []),
    "xetgpu_context_new": (ERRP, [C.POINTER(XetGpuConfig), C.POINTER(CTX)]),
    "xetgpu_context_free": (None, [CTX]),
    "xetgpu_context_set_progress_callback": (ERRP, [CTX, ProgressCallback, C.c_void_p, C.c_uint32]),
    "xetgpu_hash_buffers": (ERRP, [CTX, C.c_void_p, C.POINTER(C.c_uint64), C.c_size_t, C.c_int32, C.c_uint32,
                                   C.POINTER(C.POINTER(XetGpuResult))]),
    "xetgpu_hash_buffers_ranges": (ERRP, [CTX, C.c_void_p, C.POINTER(C.c_uint64), C.c_size_t, C.c_int32, C.c_uint32,
                                          C.POINTER(XetGpuRangeSpec), C.POINTER(C.c_uint64),
                                          C.POINTER(C.POINTER(XetGpuResult))]),
    "xetgpu_hash_files": (ERRP, [CTX, C.POINTER(C.c_char_p), C.c_size_t, C.c_uint32,
                                 C.POINTER(XetCancellationToken), C.POINTER(C.POINTER(XetGpuResult))]),
    "xetgpu_verify_file": (ERRP, [CTX, C.c_char_p, C.c_char_p]),
    "xetgpu_verify_files": (ERRP, [CTX, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p), C.c_size_t,
                                   C.POINTER(XetCancellationToken), C.POINTER(C.c_uint8)]),
    "xetgpu_verify_buffer": (ERRP, [CTX, C.c_void_p, C.c_uint64, C.c_int32, C.c_char_p]),
    "xetgpu_hash_to_hex": (None, [C.POINTER(C.c_uint8), C.c_char_p]),
    "xetgpu_hash_from_hex": (ERRP, [C.c_char_p, C.POINTER(C.c_uint8)]),
    "xetgpu_gear_candidates": (ERRP, [CTX, C.c_void_p, C.c_uint64, C.c_int32, C.POINTER(C.POINTER(C.c_uint64)),
                                      C.POINTER(C.c_uint64)]),
    "xetgpu_hash_range": (ERRP, [CTX, C.c_void_p, C.c_int32, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64,
                                 C.POINTER(C.c_uint64), C.POINTER(C.POINTER(XetGpuResult))]),
    "xetgpu_range_exit_cut": (ERRP, [CTX, C.c_void_p, C.c_int32, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64,
                                     C.POINTER(C.c_uint64)]),
    "xetgpu_merkle_from_chunks": (ERRP, [CTX, C.POINTER(C.c_uint8), C.POINTER(C.c_uint64), C.c_uint64,
                                         C.POINTER(C.c_uint8), C.POINTER(C.c_uint8)]),
    "xetgpu_merkle_from_chunks_static": (ERRP, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                                C.c_void_p]),
    "xetgpu_multi_new": (ERRP, [C.POINTER(C.c_int32), C.c_size_t, C.POINTER(XetGpuConfig), C.POINTER(CTX)]),
    "xetgpu_multi_free": (None, [CTX]),
    "xetgpu_multi_device_count": (C.c_size_t, [CTX]),
    "xetgpu_multi_set_progress_callback": (ERRP, [CTX, ProgressCallback, C.c_void_p, C.c_uint32]),
    "xetgpu_multi_hash_files": (ERRP, [CTX, C.POINTER(C.c_char_p), C.c_size_t, C.c_uint32,
                                       C.POINTER(XetCancellationToken), C.POINTER(C.POINTER(XetGpuResult))]),
    "xetgpu_multi_verify_files": (ERRP, [CTX, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p), C.c_size_t,
                                         C.POINTER(XetCancellationToken), C.POINTER(C.c_uint8)]),
    "xetgpu_multi_plan": (ERRP, [CTX, C.POINTER(C.c_char_p), C.c_size_t, C.c_uint32, C.POINTER(C.c_uint64), C.c_size_t,
                                 C.POINTER(C.c_size_t)]),
    "xetgpu_safetensors_open": (ERRP, [C.c_char_p, C.POINTER(C.POINTER(XetGpuSafetensorsIndex))]),
    "xetgpu_safetensors_parse": (ERRP, [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.POINTER(XetGpuSafetensorsIndex))]),
    "xetgpu_safetensors_free": (None, [C.POINTER(XetGpuSafetensorsIndex)]),
    "xetgpu_safetensors_split": (ERRP, [C.POINTER(XetGpuSafetensorsIndex), C.c_uint64, C.c_size_t, C.POINTER(C.c_uint64)]),
    "xetgpu_fill_splitmix": (ERRP, [CTX, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64]),
    "xetgpu_last_timings": (ERRP, [CTX, C.POINTER(XetGpuTimings)]),
    "xetgpu_free_error": (None, [ERRP]),
    "xetgpu_free_result": (None, [C.POINTER(XetGpuResult)]),
    "xetgpu_free_u64": (None, [C.POINTER(C.c_uint64)]),
}

_lib = None


def load() -> C.CDLL:
    """dlopen libxetgpu.so and bind every symbol of the header.  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `make -C ome_b200/csrc` (or __graft_entry__.build()). "
            "ome_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
