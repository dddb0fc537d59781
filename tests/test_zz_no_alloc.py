"""xetgpu_merkle_from_chunks_static: the entry point for callers that cannot allocate on the GPU.

Collected last on purpose (file name), so a run that stops at the first failure has already
run every other test.  The check: its root and file hash equal those of the allocating entry
point xetgpu_merkle_from_chunks AND those of the CPU oracle on the same lists, bit for bit (the
tolerance of every parity test here), on two
calls in a row that reuse the library's fixed scratch array — the second list is shorter than
the first, so anything the first call left behind would show.  Runs the same alone in a fresh
process or after the rest of the suite: it creates its own inputs and only needs a context for
the comparison call.
"""
import numpy as np
import pytest

import xet_oracle

pytestmark = pytest.mark.gpu


def run_static(torch, ome_b200, hashes, sizes):
    """Call through the C ABI with caller-owned device buffers; returns (root, file_hash)."""
    d_h = torch.from_numpy(hashes.reshape(-1).copy()).cuda()
    d_s = torch.from_numpy(sizes.astype(np.int64)).cuda()   # same 8 bytes per entry as uint64
    d_root = torch.full((32,), 0xAA, dtype=torch.uint8, device="cuda")
    d_fh = torch.full((32,), 0xAA, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    ome_b200.merkle_from_chunks_static(d_h.data_ptr(), d_s.data_ptr(), sizes.size, d_root.data_ptr(),
                                       d_fh.data_ptr(), stream)
    torch.cuda.synchronize()
    return bytes(d_root.cpu().numpy()), bytes(d_fh.cpu().numpy())


def test_static_scratch_merkle_matches_allocating_entry_point(hasher):
    import torch
    import ome_b200
    rng = np.random.default_rng(2024)
    # two calls in a row on the shared scratch: long list first, then shorter ones
    for n in (50_001, 1_237, 9, 2, 1, 65_536):
        hashes = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        sizes = rng.integers(1, 131_073, n).astype(np.uint64)
        want_root, want_fh = hasher.merkle_from_chunks(hashes, sizes)
        got_root, got_fh = run_static(torch, ome_b200, hashes, sizes)
        assert got_root == want_root, n
        assert got_fh == want_fh, n
        assert got_root == xet_oracle.merkle_root(hashes, sizes), n
        assert got_fh == xet_oracle.file_hash_from_chunks(hashes, sizes), n
    # empty list: zero hash, like merklehash::file_hash
    root, fh = run_static(torch, ome_b200, np.zeros((0, 32), np.uint8), np.zeros(0, np.uint64))
    assert root == bytes(32) and fh == bytes(32)


def test_static_scratch_merkle_rejects_lists_beyond_its_limit():
    import torch
    import ome_b200
    n = ome_b200.hasher.STATIC_MERKLE_MAX_CHUNKS + 1
    d_h = torch.zeros(n * 32, dtype=torch.uint8, device="cuda")
    d_s = torch.ones(n, dtype=torch.int64, device="cuda")
    d_o = torch.zeros(64, dtype=torch.uint8, device="cuda")
    with pytest.raises(ome_b200.XetGpuError) as ei:
        ome_b200.merkle_from_chunks_static(d_h.data_ptr(), d_s.data_ptr(), n, d_o.data_ptr(), d_o.data_ptr() + 32)
    assert ei.value.code == 1 and "65536" in ei.value.details


def test_static_scratch_merkle_rejects_misaligned_pointers():
    import torch
    import ome_b200
    d_h = torch.zeros(4 * 32 + 8, dtype=torch.uint8, device="cuda")
    d_s = torch.ones(5, dtype=torch.int64, device="cuda")
    d_o = torch.zeros(72, dtype=torch.uint8, device="cuda")
    for args in ((d_h.data_ptr() + 1, d_s.data_ptr(), d_o.data_ptr(), d_o.data_ptr() + 32),
                 (d_h.data_ptr(), d_s.data_ptr() + 4, d_o.data_ptr(), d_o.data_ptr() + 32),
                 (d_h.data_ptr(), d_s.data_ptr(), d_o.data_ptr() + 2, d_o.data_ptr() + 32)):
        with pytest.raises(ome_b200.XetGpuError) as ei:
            ome_b200.merkle_from_chunks_static(args[0], args[1], 4, args[2], args[3])
        assert ei.value.code == 1 and "aligned" in ei.value.message
