"""XetHasher: the host-side handle over libxetgpu's C ABI.

Naming follows the reference's Go binding (pkg/xet/xet.go): a handle created once and closed
explicitly (NewClient/Close, xet.go:241-308), errors carried as a typed exception with the
reference's code/message/details triple (XetError, xet.go:227-238, 507-522), progress delivered
through a handler callback (SetProgressHandler, xet.go:93-151).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Callable, Optional, Sequence, Tuple

import numpy as np

from . import _lib as L

MEM_AUTO, MEM_HOST, MEM_DEVICE = 0, 1, 2
WANT_CHUNKS = 1
WANT_XORBS = 2
CTX_STAGE_TIMING = 1
CTX_NO_OVERLAP = 2

MIN_CHUNK, TARGET_CHUNK, MAX_CHUNK = 8192, 65536, 131072
RANGE_LEAD = MAX_CHUNK + 64

ERR_INVALID_CONFIG, ERR_NOT_FOUND, ERR_CHECKSUM_MISMATCH, ERR_CANCELLED, ERR_IO, ERR_UNKNOWN = 1, 4, 6, 7, 8, 99


class XetGpuError(RuntimeError):
    """Mirror of the Go XetError (pkg/xet/xet.go:227-238): code, message, details."""

    def __init__(self, code: int, message: str, details: str = ""):
        self.code, self.message, self.details = code, message, details
        super().__init__(f"xet error {code}: {message}" + (f" ({details})" if details else ""))


def _check(errp) -> None:
    if not errp:
        return
    e = errp.contents
    code = e.code
    msg = (e.message or b"").decode(errors="replace")
    det = (e.details or b"").decode(errors="replace")
    L.load().xetgpu_free_error(errp)
    raise XetGpuError(code, msg, det)


def hash_to_hex(h: bytes) -> str:
    """MerkleHash::hex — four LE u64 words as %016x (what the Hub prints as X-Xet-Hash)."""
    assert len(h) == 32
    out = C.create_string_buffer(65)
    L.load().xetgpu_hash_to_hex((C.c_uint8 * 32).from_buffer_copy(h), out)
    return out.value.decode()


def hash_from_hex(s: str) -> bytes:
    out = (C.c_uint8 * 32)()
    _check(L.load().xetgpu_hash_from_hex(s.encode(), out))
    return bytes(out)


@dataclass
class FileHash:
    file_hash: bytes      # raw digest order
    merkle_root: bytes
    size: int
    n_chunks: int
    first_chunk: int

    @property
    def hex(self) -> str:
        return hash_to_hex(self.file_hash)


_FILE_DTYPE = np.dtype([("file_hash", "u1", 32), ("merkle_root", "u1", 32), ("size", "<u8"), ("n_chunks", "<u8"),
                        ("first_chunk", "<u8")])
assert _FILE_DTYPE.itemsize == C.sizeof(L.XetGpuFileResult)


class FileList:
    """Per-file results of one call: a read-only sequence of FileHash built on demand from one copied
    record array (a batch of 200,000 small files should not cost 200,000 Python objects up front)."""

    def __init__(self, records: np.ndarray):
        self.records = records

    def __len__(self) -> int:
        return int(self.records.shape[0])

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[j] for j in range(*i.indices(len(self)))]
        r = self.records[i]
        return FileHash(r["file_hash"].tobytes(), r["merkle_root"].tobytes(), int(r["size"]), int(r["n_chunks"]),
                        int(r["first_chunk"]))

    def __iter__(self):
        return (self[i] for i in range(len(self)))

    def __eq__(self, other):
        return list(self) == list(other)

    def __repr__(self):
        return f"FileList({len(self)} files)"


_XORB_DTYPE = np.dtype([("hash", "u1", 32), ("file", "<u8"), ("first_chunk", "<u8"), ("n_chunks", "<u8"), ("bytes", "<u8")])
assert _XORB_DTYPE.itemsize == C.sizeof(L.XetGpuXorb)


@dataclass
class ChunkList:
    files: FileList                  # sequence of FileHash
    chunk_ends: Optional[np.ndarray]    # uint64 [n_chunks], file-relative exclusive ends
    chunk_hashes: Optional[np.ndarray]  # uint8 [n_chunks, 32]
    xorbs: Optional[np.ndarray] = None  # structured [n_xorbs]: hash, file, first_chunk (within the file), n_chunks, bytes

    def chunks_of(self, f: int):
        fr = self.files[f]
        sl = slice(fr.first_chunk, fr.first_chunk + fr.n_chunks)
        return self.chunk_ends[sl], self.chunk_hashes[sl]


@dataclass
class RangeResult:
    exit_cut: int
    chunk_ends: np.ndarray      # file-relative
    chunk_hashes: np.ndarray


def _ptr_of(data):
    """(address, nbytes, keepalive) of bytes / numpy / torch tensor / raw int address."""
    if isinstance(data, int):
        return data, None, None
    if hasattr(data, "data_ptr"):       # torch tensor (host or device)
        return data.data_ptr(), data.numel() * data.element_size(), data
    if isinstance(data, np.ndarray):
        a = np.ascontiguousarray(data)
        return a.ctypes.data, a.nbytes, a
    a = np.frombuffer(data, dtype=np.uint8)
    return a.ctypes.data, a.nbytes, a


STATIC_MERKLE_MAX_CHUNKS = 65536


def merkle_from_chunks_static(hashes_dev_ptr: int, sizes_dev_ptr: int, n_chunks: int, out_root_dev_ptr: int,
                              out_file_hash_dev_ptr: int, stream: int = 0) -> None:
    """Merkle root + file hash of a device-resident chunk list without any device allocation.

    For callers that may not allocate on the GPU (memory owned by another allocator, latency
    paths, no XetGpuContext at hand).  All five pointers are device addresses the caller owns;
    hash and output pointers 4-byte aligned, the size array 8-byte aligned;
    at most STATIC_MERKLE_MAX_CHUNKS chunks; the kernel is only enqueued on `stream` — synchronise
    before reading the outputs, and do not let two calls overlap on the GPU (they share one
    fixed scratch array inside the library).
    """
    _check(L.load().xetgpu_merkle_from_chunks_static(hashes_dev_ptr, sizes_dev_ptr, n_chunks, out_root_dev_ptr,
                                                     out_file_hash_dev_ptr, stream or None))


class XetHasher:
    """One context on one GPU.  Thread-compatible like the reference client: calls serialise."""

    def __init__(self, device: int = -1, stream: int = 0, staging_bytes: int = 0, max_batch_bytes: int = 0,
                 stage_timing: bool = False, overlap: bool = True):
        self._lib = L.load()
        cfg = L.XetGpuConfig(device, stream or None, staging_bytes, max_batch_bytes,
                             (CTX_STAGE_TIMING if stage_timing else 0) | (0 if overlap else CTX_NO_OVERLAP))
        ctx = L.CTX()
        _check(self._lib.xetgpu_context_new(C.byref(cfg), C.byref(ctx)))
        self._ctx = ctx
        self._cb = None
        self._stream = int(stream or 0)

    # -- lifecycle (Close in pkg/xet/xet.go:288-308)
    def close(self) -> None:
        if getattr(self, "_ctx", None):
            self._lib.xetgpu_context_free(self._ctx)
            self._ctx = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- progress (SetProgressHandler in pkg/xet/xet.go:93-151)
    def set_progress_handler(self, handler: Optional[Callable[[dict], None]], throttle_ms: int = 200) -> None:
        if handler is None:
            self._cb = None
            _check(self._lib.xetgpu_context_set_progress_callback(self._ctx, L.ProgressCallback(0), None, 0))
            return

        def bridge(update, _user):
            u = update.contents
            handler(dict(phase=u.phase, total_bytes=u.total_bytes, completed_bytes=u.completed_bytes,
                         total_files=u.total_files, completed_files=u.completed_files,
                         current_file=(u.current_file or b"").decode(errors="replace")))

        self._cb = L.ProgressCallback(bridge)
        _check(self._lib.xetgpu_context_set_progress_callback(self._ctx, self._cb, None, throttle_ms))

    def _order_after_torch(self, data) -> None:
        """Device inputs must be complete with respect to the context's stream (include/xetgpu.h, XETGPU_MEM_DEVICE).
        A CUDA tensor may still have work pending on torch's current stream for its device; unless the context was
        created on that very stream, wait for it here — the context's own stream is non-blocking and takes no
        implicit ordering from torch's streams."""
        if getattr(data, "is_cuda", False):
            import torch
            s = torch.cuda.current_stream(data.device)
            if not self._stream or s.cuda_stream != self._stream:
                s.synchronize()

    # -- results
    def _take(self, resp, want_chunks: bool) -> ChunkList:
        r = resp.contents
        if r.n_files:
            raw = C.string_at(r.files, r.n_files * _FILE_DTYPE.itemsize)
            files = FileList(np.frombuffer(raw, dtype=_FILE_DTYPE))
        else:
            files = FileList(np.zeros(0, dtype=_FILE_DTYPE))
        ends = hashes = None
        if want_chunks:
            n = r.n_chunks
            if n:
                ends = np.ctypeslib.as_array(r.chunk_ends, shape=(n,)).copy()
                hashes = np.ctypeslib.as_array(r.chunk_hashes, shape=(n * 32,)).reshape(n, 32).copy()
            else:
                ends = np.zeros(0, dtype=np.uint64)
                hashes = np.zeros((0, 32), dtype=np.uint8)
        xorbs = None
        if r.n_xorbs:
            xorbs = np.frombuffer(C.string_at(r.xorbs, r.n_xorbs * _XORB_DTYPE.itemsize), dtype=_XORB_DTYPE)
        self._lib.xetgpu_free_result(resp)
        return ChunkList(files, ends, hashes, xorbs)

    # -- hashing
    def hash_buffers(self, data, offsets: Sequence[int], memory: int = MEM_AUTO, want_chunks: bool = False,
                     want_xorbs: bool = False) -> ChunkList:
        """Files packed back to back in `data` (bytes, numpy, torch tensor or raw address)."""
        addr, _, keep = _ptr_of(data)
        self._order_after_torch(data)
        off = np.ascontiguousarray(offsets, dtype=np.uint64)
        n_files = max(off.size - 1, 0)
        resp = C.POINTER(L.XetGpuResult)()
        _check(self._lib.xetgpu_hash_buffers(self._ctx, addr, off.ctypes.data_as(C.POINTER(C.c_uint64)), n_files,
                                             memory, (WANT_CHUNKS if want_chunks else 0) | (WANT_XORBS if want_xorbs else 0),
                                             C.byref(resp)))
        del keep
        return self._take(resp, want_chunks)

    def hash_buffers_ranges(self, data, offsets: Sequence[int], specs: Sequence[Tuple[int, int, int, int]],
                            memory: int = MEM_AUTO, want_chunks: bool = False):
        """Whole files and byte-range items in one batched call (xetgpu_hash_buffers_ranges).  specs[i] =
        (file_len, range_begin, range_end, entry_cut); a whole file is (len, 0, len, 0).  Returns (ChunkList, exit
        cuts per item); the ChunkList carries the chunk lists (file-relative ends) of the byte-range items, and of the
        whole files too with want_chunks."""
        addr, _, keep = _ptr_of(data)
        self._order_after_torch(data)
        off = np.ascontiguousarray(offsets, dtype=np.uint64)
        n = max(off.size - 1, 0)
        assert len(specs) == n
        arr = (L.XetGpuRangeSpec * max(n, 1))(*[L.XetGpuRangeSpec(*[int(x) for x in sp]) for sp in specs])
        exits = (C.c_uint64 * max(n, 1))()
        resp = C.POINTER(L.XetGpuResult)()
        _check(self._lib.xetgpu_hash_buffers_ranges(self._ctx, addr, off.ctypes.data_as(C.POINTER(C.c_uint64)), n, memory,
                                                    WANT_CHUNKS if want_chunks else 0, arr, exits, C.byref(resp)))
        del keep
        return self._take(resp, True), [int(exits[i]) for i in range(n)]

    def hash_bytes(self, data, memory: int = MEM_AUTO, want_chunks: bool = False, want_xorbs: bool = False) -> ChunkList:
        _, n, _ = _ptr_of(data)
        return self.hash_buffers(data, [0, n], memory, want_chunks, want_xorbs)

    def hash_files(self, paths: Sequence[str], want_chunks: bool = False,
                   should_cancel: Optional[Callable[[], bool]] = None, want_xorbs: bool = False) -> ChunkList:
        arr = (C.c_char_p * len(paths))(*[p.encode() for p in paths])
        tok = None
        if should_cancel is not None:
            cb = L.CancellationCallback(lambda _u: bool(should_cancel()))
            tok = L.XetCancellationToken(cb, None)
        resp = C.POINTER(L.XetGpuResult)()
        _check(self._lib.xetgpu_hash_files(self._ctx, arr, len(paths),
                                           (WANT_CHUNKS if want_chunks else 0) | (WANT_XORBS if want_xorbs else 0),
                                           C.byref(tok) if tok is not None else None, C.byref(resp)))
        return self._take(resp, want_chunks)

    def hash_file(self, path: str) -> str:
        """Xet file hash of one file as the 64-digit hex the Hub advertises."""
        return self.hash_files([path]).files[0].hex

    # -- verification (shape precedent: hub.VerifyChecksum, pkg/hfutil/hub/utils.go:139-158)
    def verify_file(self, path: str, expected_hex: str) -> None:
        _check(self._lib.xetgpu_verify_file(self._ctx, path.encode(), expected_hex.encode()))

    def verify_files(self, paths: Sequence[str], expected_hex: Sequence[Optional[str]]) -> list:
        """Verify a snapshot in one batch.  Returns the per-file ok flags; raises XetGpuError(code 6)
        listing the mismatches when any file differs.  None / "" skips a file."""
        n = len(paths)
        assert len(expected_hex) == n
        cp = (C.c_char_p * n)(*[p.encode() for p in paths])
        ch = (C.c_char_p * n)(*[(h.encode() if h else None) for h in expected_hex])
        ok = (C.c_uint8 * max(n, 1))()
        self._last_ok = ok
        _check(self._lib.xetgpu_verify_files(self._ctx, cp, ch, n, None, ok))
        return [bool(ok[i]) for i in range(n)]

    def verify_buffer(self, data, expected_hex: str, memory: int = MEM_AUTO) -> None:
        addr, n, keep = _ptr_of(data)
        self._order_after_torch(data)
        _check(self._lib.xetgpu_verify_buffer(self._ctx, addr, n, memory, expected_hex.encode()))
        del keep

    # -- stage-level
    def gear_candidates(self, data, memory: int = MEM_AUTO) -> np.ndarray:
        addr, n, keep = _ptr_of(data)
        self._order_after_torch(data)
        out = C.POINTER(C.c_uint64)()
        cnt = C.c_uint64()
        _check(self._lib.xetgpu_gear_candidates(self._ctx, addr, n, memory, C.byref(out), C.byref(cnt)))
        res = np.ctypeslib.as_array(out, shape=(cnt.value,)).copy() if cnt.value else np.zeros(0, dtype=np.uint64)
        self._lib.xetgpu_free_u64(out)
        del keep
        return res

    def hash_range(self, data, file_len: int, range_begin: int, range_end: int, entry_cut: int,
                   memory: int = MEM_AUTO) -> RangeResult:
        addr, _, keep = _ptr_of(data)
        self._order_after_torch(data)
        resp = C.POINTER(L.XetGpuResult)()
        exit_cut = C.c_uint64()
        _check(self._lib.xetgpu_hash_range(self._ctx, addr, memory, file_len, range_begin, range_end, entry_cut,
                                           C.byref(exit_cut), C.byref(resp)))
        del keep
        cl = self._take(resp, True)
        return RangeResult(exit_cut.value, cl.chunk_ends, cl.chunk_hashes)

    def range_exit_cut(self, data, file_len: int, range_begin: int, range_end: int, entry_cut: int,
                       memory: int = MEM_AUTO) -> int:
        addr, _, keep = _ptr_of(data)
        self._order_after_torch(data)
        exit_cut = C.c_uint64()
        _check(self._lib.xetgpu_range_exit_cut(self._ctx, addr, memory, file_len, range_begin, range_end, entry_cut,
                                               C.byref(exit_cut)))
        del keep
        return exit_cut.value

    def merkle_from_chunks(self, chunk_hashes, chunk_sizes):
        """(merkle root, file hash) of a chunk list: numpy arrays, or torch tensors on the host or the device
        (uint8 [n, 32] and int64/uint64 [n])."""
        root = (C.c_uint8 * 32)()
        fh = (C.c_uint8 * 32)()
        if hasattr(chunk_hashes, "data_ptr"):
            h, s = chunk_hashes.contiguous(), chunk_sizes.contiguous()
            self._order_after_torch(h)
            assert h.element_size() == 1 and s.element_size() == 8 and h.numel() == 32 * s.numel()
            hp = C.cast(h.data_ptr(), C.POINTER(C.c_uint8))
            sp = C.cast(s.data_ptr(), C.POINTER(C.c_uint64))
            n = s.numel()
        else:
            h = np.ascontiguousarray(chunk_hashes, dtype=np.uint8).reshape(-1)
            s = np.ascontiguousarray(chunk_sizes, dtype=np.uint64)
            hp, sp, n = h.ctypes.data_as(C.POINTER(C.c_uint8)), s.ctypes.data_as(C.POINTER(C.c_uint64)), s.size
        _check(self._lib.xetgpu_merkle_from_chunks(self._ctx, hp, sp, n, root, fh))
        return bytes(root), bytes(fh)

    def fill_splitmix(self, device_ptr: int, seed: int, byte_off: int, length: int) -> None:
        _check(self._lib.xetgpu_fill_splitmix(self._ctx, device_ptr, seed & (2**64 - 1), byte_off, length))

    def last_timings(self) -> dict:
        t = L.XetGpuTimings()
        _check(self._lib.xetgpu_last_timings(self._ctx, C.byref(t)))
        return {k: getattr(t, k) for k, _ in L.XetGpuTimings._fields_}


class XetMultiHasher:
    """Every listed GPU of the node from ONE process (xetgpu_multi_* in include/xetgpu.h): what a single model-agent
    process would hold.  Files are laid end to end and cut into one byte range per device; see the header."""

    def __init__(self, devices: Optional[Sequence[int]] = None, staging_bytes: int = 0, max_batch_bytes: int = 0):
        self._lib = L.load()
        cfg = L.XetGpuConfig(-1, None, staging_bytes, max_batch_bytes, 0)
        ctx = L.CTX()
        if devices:
            arr = (C.c_int32 * len(devices))(*devices)
            _check(self._lib.xetgpu_multi_new(arr, len(devices), C.byref(cfg), C.byref(ctx)))
        else:
            _check(self._lib.xetgpu_multi_new(None, 0, C.byref(cfg), C.byref(ctx)))
        self._ctx = ctx
        self._cb = None

    @property
    def n_devices(self) -> int:
        return int(self._lib.xetgpu_multi_device_count(self._ctx))

    def close(self) -> None:
        if getattr(self, "_ctx", None):
            self._lib.xetgpu_multi_free(self._ctx)
            self._ctx = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_progress_handler(self, handler: Optional[Callable[[dict], None]], throttle_ms: int = 200) -> None:
        if handler is None:
            self._cb = None
            _check(self._lib.xetgpu_multi_set_progress_callback(self._ctx, L.ProgressCallback(0), None, 0))
            return

        def bridge(update, _user):
            u = update.contents
            handler(dict(phase=u.phase, total_bytes=u.total_bytes, completed_bytes=u.completed_bytes,
                         total_files=u.total_files, completed_files=u.completed_files,
                         current_file=(u.current_file or b"").decode(errors="replace")))

        self._cb = L.ProgressCallback(bridge)
        _check(self._lib.xetgpu_multi_set_progress_callback(self._ctx, self._cb, None, throttle_ms))

    def hash_files(self, paths: Sequence[str], want_chunks: bool = False,
                   should_cancel: Optional[Callable[[], bool]] = None) -> ChunkList:
        arr = (C.c_char_p * len(paths))(*[p.encode() for p in paths])
        tok = None
        if should_cancel is not None:
            cb = L.CancellationCallback(lambda _u: bool(should_cancel()))
            tok = L.XetCancellationToken(cb, None)
        resp = C.POINTER(L.XetGpuResult)()
        _check(self._lib.xetgpu_multi_hash_files(self._ctx, arr, len(paths), WANT_CHUNKS if want_chunks else 0,
                                                 C.byref(tok) if tok is not None else None, C.byref(resp)))
        return XetHasher._take(self, resp, want_chunks)

    def plan(self, paths: Sequence[str], tensor_snap: bool = True) -> list:
        """[(device index, file index, begin, end)] — the byte ranges hash_files gives each device."""
        arr = (C.c_char_p * len(paths))(*[p.encode() for p in paths])
        cap = 2 * len(paths) + 4 * self.n_devices + 4
        out = (C.c_uint64 * (4 * cap))()
        n = C.c_size_t()
        _check(self._lib.xetgpu_multi_plan(self._ctx, arr, len(paths), 0 if tensor_snap else 4, out, cap, C.byref(n)))
        return [tuple(int(out[4 * i + k]) for k in range(4)) for i in range(min(n.value, cap))]

    def verify_files(self, paths: Sequence[str], expected_hex: Sequence[Optional[str]]) -> list:
        n = len(paths)
        assert len(expected_hex) == n
        cp = (C.c_char_p * n)(*[p.encode() for p in paths])
        ch = (C.c_char_p * n)(*[(h.encode() if h else None) for h in expected_hex])
        ok = (C.c_uint8 * max(n, 1))()
        _check(self._lib.xetgpu_multi_verify_files(self._ctx, cp, ch, n, None, ok))
        return [bool(ok[i]) for i in range(n)]
