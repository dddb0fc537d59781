"""ctypes/numpy front-end of the CPU oracle (oracle/xet_oracle.c).

TEST INFRASTRUCTURE ONLY — see the header of xet_oracle.c.  Imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs; never by
anything under ome_b200/.

The functions restate huggingface/xet-core (pinned by /root/reference
pkg/xet/Cargo.lock:237) as specified in SURVEY.md Appendix A; parity is pinned against
hf_xet 1.5.0 through tests/golden/xet_golden.json (scripts/gen_golden.py).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libxet_oracle.so")
_lib = None

_u8p = ctypes.POINTER(ctypes.c_uint8)
_u64p = ctypes.POINTER(ctypes.c_uint64)


def build(force: bool = False) -> str:
    """Compile libxet_oracle.so with the committed Makefile (gcc only)."""
    src = os.path.join(_HERE, "xet_oracle.c")
    stale = (not os.path.exists(_LIB_PATH)) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src)
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-B" if force else "-s"], check=True,
                       stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = ctypes.CDLL(_LIB_PATH)
        L.xo_blake3_keyed.argtypes = [_u8p, _u8p, ctypes.c_size_t, _u8p]
        L.xo_chunk_hash.argtypes = [_u8p, ctypes.c_size_t, _u8p]
        L.xo_chunk_boundaries.argtypes = [_u8p, ctypes.c_uint64, _u64p, ctypes.c_size_t]
        L.xo_chunk_boundaries.restype = ctypes.c_size_t
        L.xo_window_candidates.argtypes = [_u8p, ctypes.c_uint64, _u64p, ctypes.c_size_t]
        L.xo_window_candidates.restype = ctypes.c_size_t
        L.xo_merkle_root.argtypes = [_u8p, _u64p, ctypes.c_size_t, _u8p]
        L.xo_file_hash.argtypes = [_u8p, _u64p, ctypes.c_size_t, _u8p]
        L.xo_hash_file.argtypes = [_u8p, ctypes.c_uint64, _u64p, _u8p, ctypes.c_size_t, _u8p, _u8p]
        L.xo_hash_file.restype = ctypes.c_size_t
        L.xo_hex.argtypes = [_u8p, ctypes.c_char_p]
        L.xo_from_hex.argtypes = [ctypes.c_char_p, _u8p]
        L.xo_from_hex.restype = ctypes.c_int
        L.xo_xorbs.argtypes = [_u8p, _u64p, ctypes.c_size_t, _u64p, _u8p, ctypes.c_size_t]
        L.xo_xorbs.restype = ctypes.c_size_t
        L.xo_splitmix_fill.argtypes = [ctypes.c_uint64, ctypes.c_uint64, _u8p, ctypes.c_uint64]
        _lib = L
    return _lib


def _as_u8(data) -> np.ndarray:
    if isinstance(data, (bytes, bytearray, memoryview)):
        a = np.frombuffer(data, dtype=np.uint8)
    else:
        a = np.ascontiguousarray(data, dtype=np.uint8)
    return a


def _p8(a: np.ndarray):
    return a.ctypes.data_as(_u8p)


def _p64(a: np.ndarray):
    return a.ctypes.data_as(_u64p)


def blake3_keyed(key: bytes, data) -> bytes:
    a = _as_u8(data)
    k = np.frombuffer(key, dtype=np.uint8)
    assert k.size == 32
    out = np.zeros(32, dtype=np.uint8)
    lib().xo_blake3_keyed(_p8(k), _p8(a), a.size, _p8(out))
    return out.tobytes()


def chunk_hash(data) -> bytes:
    a = _as_u8(data)
    out = np.zeros(32, dtype=np.uint8)
    lib().xo_chunk_hash(_p8(a), a.size, _p8(out))
    return out.tobytes()


def chunk_boundaries(data) -> np.ndarray:
    """Exclusive end offsets of the chunks of one file (uint64)."""
    a = _as_u8(data)
    cap = a.size // 8192 + 2
    ends = np.zeros(cap, dtype=np.uint64)
    n = lib().xo_chunk_boundaries(_p8(a), a.size, _p64(ends), cap)
    assert n <= cap
    return ends[:n].copy()


def window_candidates(data) -> np.ndarray:
    """Byte indices i whose 64-byte gear window hash has the 16 top bits clear."""
    a = _as_u8(data)
    cap = 1024
    while True:
        idx = np.zeros(cap, dtype=np.uint64)
        n = lib().xo_window_candidates(_p8(a), a.size, _p64(idx), cap)
        if n <= cap:
            return idx[:n].copy()
        cap = n


def merkle_root(hashes: np.ndarray, sizes: np.ndarray) -> bytes:
    h = np.ascontiguousarray(hashes, dtype=np.uint8).reshape(-1, 32)
    s = np.ascontiguousarray(sizes, dtype=np.uint64)
    assert h.shape[0] == s.size
    out = np.zeros(32, dtype=np.uint8)
    lib().xo_merkle_root(_p8(h), _p64(s), s.size, _p8(out))
    return out.tobytes()


def file_hash_from_chunks(hashes: np.ndarray, sizes: np.ndarray) -> bytes:
    h = np.ascontiguousarray(hashes, dtype=np.uint8).reshape(-1, 32)
    s = np.ascontiguousarray(sizes, dtype=np.uint64)
    out = np.zeros(32, dtype=np.uint8)
    lib().xo_file_hash(_p8(h), _p64(s), s.size, _p8(out))
    return out.tobytes()


def xorbs(hashes: np.ndarray, sizes: np.ndarray):
    """Greedy xorb formation over one file's chunk list (<= 64 MiB and <= 8192 chunks per xorb; grouping parity
    unpinned, see xo_xorbs).  Returns (starts uint64 [n_xorbs + 1], xorb hashes uint8 [n_xorbs, 32])."""
    h = np.ascontiguousarray(hashes, dtype=np.uint8).reshape(-1, 32)
    s = np.ascontiguousarray(sizes, dtype=np.uint64)
    n = s.size
    cap = n + 1
    starts = np.zeros(cap + 1, dtype=np.uint64)
    roots = np.zeros((cap, 32), dtype=np.uint8)
    c = lib().xo_xorbs(_p8(h), _p64(s), n, _p64(starts), _p8(roots), cap)
    return starts[:c + 1].copy(), roots[:c].copy()


@dataclass
class FileResult:
    chunk_ends: np.ndarray      # uint64 [n_chunks], exclusive end offsets
    chunk_hashes: np.ndarray    # uint8 [n_chunks, 32], raw digest order
    merkle_root: bytes          # aggregated_node_hash (== xorb_hash of all chunks)
    file_hash: bytes            # raw digest order; hex via to_hex()


def hash_file(data) -> FileResult:
    a = _as_u8(data)
    cap = a.size // 8192 + 2
    ends = np.zeros(cap, dtype=np.uint64)
    hashes = np.zeros((cap, 32), dtype=np.uint8)
    root = np.zeros(32, dtype=np.uint8)
    fh = np.zeros(32, dtype=np.uint8)
    n = lib().xo_hash_file(_p8(a), a.size, _p64(ends), _p8(hashes), cap, _p8(root), _p8(fh))
    if n == 0:
        root[:] = 0
    return FileResult(ends[:n].copy(), hashes[:n].copy(), root.tobytes(), fh.tobytes())


def to_hex(h: bytes) -> str:
    a = np.frombuffer(h, dtype=np.uint8)
    out = ctypes.create_string_buffer(65)
    lib().xo_hex(_p8(a), out)
    return out.value.decode()


def from_hex(s: str) -> bytes:
    out = np.zeros(32, dtype=np.uint8)
    if lib().xo_from_hex(s.encode(), _p8(out)) != 0:
        raise ValueError("not a 64-digit hex MerkleHash")
    return out.tobytes()


def splitmix_bytes(seed: int, n: int, byte_off: int = 0) -> np.ndarray:
    out = np.empty(n, dtype=np.uint8)
    if n:
        lib().xo_splitmix_fill(seed & (2**64 - 1), byte_off, _p8(out), n)
    return out
