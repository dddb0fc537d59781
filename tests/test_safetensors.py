"""xetgpu_safetensors_* — the shard header as pkg/hfutil/modelconfig/safetensors.go:67-141 reads it, plus the
tensor-aligned byte-range split.  Host code of the library: runs without a GPU."""
import json
import os
import struct

import pytest

import ome_b200
from ome_b200 import safetensors as st

FIX = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_fixtures")


def shard(tensors, metadata=None, pad=8):
    """tensors: [(name, dtype, shape, nbytes)] laid out back to back -> bytes of a safetensors file."""
    hdr, off = {}, 0
    if metadata is not None:
        hdr["__metadata__"] = metadata
    for name, dtype, shape, nbytes in tensors:
        hdr[name] = {"dtype": dtype, "shape": shape, "data_offsets": [off, off + nbytes]}
        off += nbytes
    j = json.dumps(hdr, separators=(",", ":")).encode()
    j += b" " * (-len(j) % pad)
    return struct.pack("<Q", len(j)) + j + bytes(off)


def test_reference_fixture_header():
    """The reference's own fixture (tiny-random-PhiModel shard): independent count with Python's json module."""
    p = os.path.join(FIX, "model-1-of-2.safetensors")
    raw = open(p, "rb").read()
    n = struct.unpack("<Q", raw[:8])[0]
    hdr = json.loads(raw[8:8 + n])
    want = 0
    for k, v in hdr.items():
        if k == "__metadata__" or not v["shape"]:
            continue
        c = 1
        for d in v["shape"]:
            c *= d
        want += c
    ix = st.open_index(p)
    assert st.ParseSafetensors(p) == ix.parameter_count == want == 46282
    assert ix.header_len == n and ix.data_start == 8 + n and ix.file_size == len(raw)
    assert len(ix.tensors) == len(hdr) - 1
    assert [t.begin for t in ix.tensors] == sorted(t.begin for t in ix.tensors)
    for t in ix.tensors:
        assert [t.begin - ix.data_start, t.end - ix.data_start] == hdr[t.name]["data_offsets"] and t.dtype == hdr[t.name]["dtype"]
    assert ix.tensors[-1].end == len(raw)
    assert st.parse_index(raw[:8 + n], len(raw)).parameter_count == want


def test_error_behaviour_follows_the_reference(tmp_path):
    def err(data, name="x.safetensors"):
        p = tmp_path / name
        p.write_bytes(data)
        with pytest.raises(ome_b200.XetGpuError) as ei:
            st.ParseSafetensors(str(p))
        return ei.value

    assert "failed to read header length" in err(b"\x01\x02").message
    e = err(struct.pack("<Q", 11 * 1024 * 1024) + b"{}")
    assert e.code == 1 and "exceeds maximum allowed size of 10485760 bytes" in e.message
    assert "failed to read JSON header" in err(struct.pack("<Q", 100) + b"{}").message
    assert "failed to parse JSON header" in err(struct.pack("<Q", 5) + b"{\"a\":").message
    bad = shard([("w", "F32", [4, 0], 0)])
    assert "tensor 'w'" in err(bad).message and "invalid dimension 0 at index 1" in err(bad).message
    big = shard([("w", "F32", [1 << 40, 1 << 40], 0)])
    assert "dimension overflow for tensor 'w'" in err(big).message
    two = shard([("a", "F32", [1 << 62], 0), ("b", "F32", [1 << 62], 0)])
    assert "parameter count overflow" in err(two).message
    with pytest.raises(ome_b200.XetGpuError) as ei:
        st.ParseSafetensors(str(tmp_path / "missing.safetensors"))
    assert ei.value.code == 4 and "failed to open safetensors file" in ei.value.message
    with pytest.raises(ome_b200.XetGpuError) as ei:
        st.ParseSafetensors("")
    assert "safetensors file path cannot be empty" in ei.value.message
    # scalars (empty shape) and metadata are skipped, as in the reference
    ok = tmp_path / "ok.safetensors"
    ok.write_bytes(shard([("s", "F32", [], 4), ("w", "BF16", [3, 5], 30)], metadata={"format": "pt", "note": 'a "quoted" {brace}'}))
    assert st.ParseSafetensors(str(ok)) == 15
    # data_offsets beyond the file are refused (the partitioner relies on them)
    raw = bytearray(shard([("w", "F32", [4], 16)]))
    assert "data_offsets outside the file" in err(bytes(raw[:-8])).message


def test_split_places_bounds_on_tensor_boundaries(tmp_path):
    mib = 1 << 20
    tensors = [(f"layer{i}.w", "BF16", [mib // 2 * (1 + i % 3)], mib * (1 + i % 3)) for i in range(24)]   # 1..3 MiB each
    p = tmp_path / "m.safetensors"
    p.write_bytes(shard(tensors))
    ix = st.open_index(str(p))
    marks = {t.begin for t in ix.tensors} | {t.end for t in ix.tensors}
    for n in (2, 3, 8):
        b = ix.split(n)
        assert len(b) == n + 1 and b[0] == 0 and b[-1] == ix.file_size and b == sorted(b)
        assert all(x in marks for x in b[1:-1]), (n, b)
        assert all(abs(b[r] - ix.file_size * r // n) <= ix.file_size // n // 4 for r in range(1, n))
        assert all(y - x >= 131136 + 4096 for x, y in zip(b, b[1:]) if y > x)
    # one giant tensor: no boundary within reach -> the aligned equal split
    q = tmp_path / "g.safetensors"
    q.write_bytes(shard([("emb", "F32", [16 * mib // 4], 16 * mib)]))
    g = st.open_index(str(q))
    b = g.split(4)
    assert all(x % 4096 == 0 for x in b[1:-1]) and b[1] == g.file_size // 4 // 4096 * 4096
    # tiny file: nothing to share
    assert ix.split(3, 100_000)[1:] == [100_000, 100_000, 100_000] or ix.split(3, 100_000)[1] in (0, 100_000)
