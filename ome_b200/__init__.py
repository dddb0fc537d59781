"""ome_b200 — B200-native Xet content addressing (chunk boundaries, chunk hashes, file hash).

Host-side mirror of the Go surface OME's pkg/xet would gain: thin wrappers
over the C ABI in include/xetgpu.h.  The compute lives in hand-written sm_100a CUDA kernels
(ome_b200/csrc); this package never computes a hash on the CPU.
"""
from .hasher import (ChunkList, FileHash, FileList, RangeResult, XetGpuError, XetHasher, XetMultiHasher, hash_from_hex, hash_to_hex,  # noqa: F401
                     merkle_from_chunks_static)

__all__ = ["XetHasher", "XetMultiHasher", "XetGpuError", "FileHash", "FileList", "ChunkList", "RangeResult", "hash_to_hex", "hash_from_hex"]
