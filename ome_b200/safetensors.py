"""Safetensors headers, as OME's model-config parser reads them (pkg/hfutil/modelconfig/safetensors.go:67-141).

Thin binding of xetgpu_safetensors_* (include/xetgpu.h).  `ParseSafetensors(path)` keeps the reference function's
name and result — the parameter count of one shard — and its error texts; `open_index` adds the tensor table with
absolute byte extents and `split` the tensor-aligned byte ranges the multi-device file call uses.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List

from . import _lib as L
from .hasher import _check


@dataclass
class Tensor:
    name: str
    dtype: str
    begin: int        # byte extent in the file
    end: int
    n_elements: int
    rank: int


@dataclass
class SafetensorsIndex:
    file_size: int
    header_len: int
    data_start: int
    parameter_count: int
    tensors: List[Tensor]

    def split(self, n_ranges: int, file_size: int = 0) -> List[int]:
        """n_ranges + 1 ascending bounds with interior bounds on tensor boundaries (xetgpu_safetensors_split)."""
        return _split(self, n_ranges, file_size or self.file_size)


def _take(ixp) -> SafetensorsIndex:
    ix = ixp.contents
    tensors = [Tensor(ix.tensors[i].name.decode(), ix.tensors[i].dtype.decode(), ix.tensors[i].begin, ix.tensors[i].end,
                      ix.tensors[i].n_elements, ix.tensors[i].rank) for i in range(ix.n_tensors)]
    out = SafetensorsIndex(ix.file_size, ix.header_len, ix.data_start, ix.parameter_count, tensors)
    L.load().xetgpu_safetensors_free(ixp)
    return out


def open_index(path: str) -> SafetensorsIndex:
    ixp = C.POINTER(L.XetGpuSafetensorsIndex)()
    _check(L.load().xetgpu_safetensors_open(path.encode(), C.byref(ixp)))
    return _take(ixp)


def parse_index(head: bytes, file_size: int = 0) -> SafetensorsIndex:
    ixp = C.POINTER(L.XetGpuSafetensorsIndex)()
    buf = C.create_string_buffer(bytes(head), len(head))
    _check(L.load().xetgpu_safetensors_parse(buf, len(head), file_size, C.byref(ixp)))
    return _take(ixp)


def ParseSafetensors(path: str) -> int:
    """modelconfig.ParseSafetensors: the parameter count of one shard."""
    return open_index(path).parameter_count


def _split(index: SafetensorsIndex, n_ranges: int, file_size: int) -> List[int]:
    # rebuild a C index from the Python one (names are not needed for the split)
    arr = (L.XetGpuTensorInfo * max(1, len(index.tensors)))()
    for i, t in enumerate(index.tensors):
        arr[i].begin, arr[i].end = t.begin, t.end
    ix = L.XetGpuSafetensorsIndex(index.file_size, index.header_len, index.data_start, len(index.tensors), arr,
                                  index.parameter_count)
    out = (C.c_uint64 * (n_ranges + 1))()
    _check(L.load().xetgpu_safetensors_split(C.byref(ix), file_size, n_ranges, out))
    return [int(x) for x in out]
