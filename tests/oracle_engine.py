"""Oracle-backed stand-in for XetHasher, for CPU tests of the multi-rank host logic only.

It answers hash_range / hash_buffers / merkle_from_chunks with the CPU oracle (SURVEY.md A.2
candidate walk in Python + oracle chunk hashes), so that ome_b200.sharded can be exercised over
gloo without a GPU.  Test infrastructure: never imported by the product.
"""
from dataclasses import dataclass

import numpy as np

import xet_oracle

MIN, MAX, LEAD = 8192, 131072, 131072 + 64


@dataclass
class _Range:
    exit_cut: int
    chunk_ends: np.ndarray
    chunk_hashes: np.ndarray


@dataclass
class _File:
    file_hash: bytes
    n_chunks: int


@dataclass
class _List:
    files: list


def _np(data):
    if hasattr(data, "numpy"):
        return data.cpu().numpy()
    return np.frombuffer(data, dtype=np.uint8) if isinstance(data, (bytes, bytearray, memoryview)) else np.asarray(data)


class OracleEngine:
    def __init__(self):
        self.range_calls = []

    def hash_buffers(self, data, offsets, memory=0, want_chunks=False):
        a = _np(data)
        files = []
        for f in range(len(offsets) - 1):
            r = xet_oracle.hash_file(a[int(offsets[f]):int(offsets[f + 1])])
            files.append(_File(r.file_hash, int(r.chunk_ends.size)))
        return _List(files)

    def hash_buffers_ranges(self, data, offsets, specs, memory=0):
        """Whole files and byte-range items in one call; returns (list-like with .files / .chunks_of, exit cuts)."""
        a = _np(data)
        files, chunks, exits = [], [], []
        for i, (flen, begin, end, entry) in enumerate(specs):
            d = a[int(offsets[i]):int(offsets[i + 1])]
            if begin == 0 and end == flen:
                r = xet_oracle.hash_file(d)
                files.append(_File(r.file_hash, int(r.chunk_ends.size)))
                chunks.append((r.chunk_ends, r.chunk_hashes))
                exits.append(int(r.chunk_ends[-1]) if r.chunk_ends.size else 0)
            else:
                r = self.hash_range(d, flen, begin, end, entry)
                files.append(_File(bytes(32), int(r.chunk_ends.size)))
                chunks.append((r.chunk_ends, r.chunk_hashes))
                exits.append(int(r.exit_cut))
        out = _List(files)
        out.chunks_of = lambda i: chunks[i]
        return out, exits

    def merkle_from_chunks(self, hashes, sizes):
        if hasattr(hashes, "numpy"):
            hashes, sizes = hashes.cpu().numpy().reshape(-1, 32), sizes.cpu().numpy().astype(np.uint64)
        return xet_oracle.merkle_root(hashes, sizes) if len(sizes) else bytes(32), \
            xet_oracle.file_hash_from_chunks(hashes, sizes)

    def hash_range(self, data, file_len, begin, end, entry, memory=0):
        a = _np(data)
        lead = min(begin, LEAD)
        pos0 = begin - lead
        assert a.size >= end - pos0 and pos0 <= entry <= begin
        self.range_calls.append((begin, end, entry))
        a = a[: end - pos0]
        cand = [int(c) + pos0 for c in xet_oracle.window_candidates(a) if c >= 63 or pos0 == 0]
        eof = end == file_len
        ends, start, i = [], entry, 0
        while True:
            if eof and start >= end:
                break
            hi = start + MAX
            if eof:
                hi = min(hi, end)
            while i < len(cand) and cand[i] + 1 < start + MIN:
                i += 1
            cut = cand[i] + 1 if i < len(cand) and cand[i] + 1 <= hi else hi
            if not eof and cut > end:
                break
            ends.append(cut)
            start = cut
        hashes = np.zeros((len(ends), 32), dtype=np.uint8)
        s = entry
        for k, e in enumerate(ends):
            hashes[k] = np.frombuffer(xet_oracle.chunk_hash(a[s - pos0:e - pos0]), dtype=np.uint8)
            s = e
        return _Range(ends[-1] if ends else entry, np.asarray(ends, dtype=np.uint64), hashes)
