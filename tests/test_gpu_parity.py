"""Parity of the CUDA path (through the C ABI) with the oracle and the hf_xet golden vectors.

Bar: bit-exact — every chunk boundary, chunk hash, merkle root and file hash.
"""
import hashlib
import os

import numpy as np
import pytest

import xet_inputs
import xet_oracle
from conftest import load_golden

pytestmark = pytest.mark.gpu

GOLDEN = load_golden()
SMALL = [c for c in GOLDEN["cases"] if not c.get("large")]
LARGE = [c for c in GOLDEN["cases"] if c.get("large")]


def check_against_oracle(cl, f, data):
    """File f of ChunkList cl vs the oracle run on `data`."""
    import ome_b200
    ref = xet_oracle.hash_file(data)
    fr = cl.files[f]
    ends, hashes = cl.chunks_of(f)
    assert fr.size == data.size
    assert fr.n_chunks == ref.chunk_ends.size
    assert np.array_equal(ends, ref.chunk_ends)
    assert np.array_equal(hashes, ref.chunk_hashes)
    assert fr.merkle_root == ref.merkle_root
    assert fr.file_hash == ref.file_hash
    return ome_b200.hash_to_hex(fr.file_hash)


@pytest.mark.parametrize("case", SMALL, ids=[c["name"] for c in SMALL])
def test_golden_case_host_input(hasher, case):
    data = xet_inputs.build(case["spec"])
    cl = hasher.hash_bytes(data, want_chunks=True)
    fr = cl.files[0]
    assert fr.hex == case["file_hash"]                      # hf_xet ground truth
    assert fr.n_chunks == case["n_chunks"]
    ends, hashes = cl.chunks_of(0)
    assert [int(x) for x in ends[:16]] == case["cuts_head"]
    assert hashlib.sha256(ends.astype("<u8").tobytes()).hexdigest() == case["cuts_sha256"]
    assert hashlib.sha256(hashes.tobytes()).hexdigest() == case["chunk_hashes_sha256"]
    import ome_b200
    assert ome_b200.hash_to_hex(fr.merkle_root) == case["merkle_root"]
    if data.size <= (4 << 20):
        check_against_oracle(cl, 0, data)


@pytest.mark.parametrize("name", ["splitmix_seed1_1MiB", "dense_W_x16384", "deadzone_after_forced",
                                  "zeros_8x131072", "safetensors_shaped_2MiB", "splitmix_seed2_10MiB"])
def test_gear_candidates_match_oracle(hasher, name):
    case = next(c for c in SMALL if c["name"] == name)
    data = xet_inputs.build(case["spec"])
    got = hasher.gear_candidates(data)
    want = xet_oracle.window_candidates(data)
    # positions < 63 depend on bytes in front of the buffer (there are none): never used by K2
    assert np.array_equal(got[got >= 63], want[want >= 63])


def test_dense_candidates_overflow_path(hasher):
    """64 MiB of the 64-byte W pattern: 1M candidates, every tile overflows the staged list."""
    case = next(c for c in SMALL if c["name"] == "dense_W_x1048576")
    data = xet_inputs.build(case["spec"])
    got = hasher.gear_candidates(data)
    want = xet_oracle.window_candidates(data)
    assert np.array_equal(got[got >= 63], want[want >= 63])
    assert hasher.hash_bytes(data).files[0].hex == case["file_hash"]


def test_all_golden_cases_in_one_batch(hasher):
    """Many small files packed in one call (BASELINE config 5 shape): per-file state must reset."""
    cases = [c for c in SMALL if c["n"] <= (2 << 20)]
    datas = [xet_inputs.build(c["spec"]) for c in cases]
    offsets = np.concatenate([[0], np.cumsum([d.size for d in datas])]).astype(np.uint64)
    packed = np.concatenate([d for d in datas if d.size] or [np.zeros(0, np.uint8)])
    cl = hasher.hash_buffers(packed, offsets, want_chunks=True)
    assert len(cl.files) == len(cases)
    for f, (c, d) in enumerate(zip(cases, datas)):
        assert cl.files[f].hex == c["file_hash"], c["name"]
        check_against_oracle(cl, f, d)


def test_ragged_random_files_batch(hasher):
    rng = np.random.default_rng(11)
    sizes = [0, 1, 5, 63, 64, 1024, 4096, 4097, 8191, 8192, 8193, 0, 20000, 70001, 131072, 131073, 300000, 1 << 20]
    sizes += [int(x) for x in np.exp(rng.uniform(np.log(1), np.log(3e6), 30))]
    datas = [xet_inputs.splitmix_bytes(500 + i, n) for i, n in enumerate(sizes)]
    offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
    packed = np.concatenate(datas)
    cl = hasher.hash_buffers(packed, offsets, want_chunks=True)
    for f, d in enumerate(datas):
        check_against_oracle(cl, f, d)


SEG = 2 << 20   # kSegBytes of the boundary-resolution kernels


@pytest.mark.parametrize("spec", [
    # chains that never re-merge: candidate-free run after an unaligned prefix, across many segments
    [["splitmix", 61, (1 << 20) + 5], ["zeros", 9 << 20], ["splitmix", 62, 3 << 20]],
    # dense candidates (every 64 bytes) in the middle of random data
    [["splitmix", 63, (3 << 20) + 77], ["repeat", xet_inputs.W_HEX, 40000], ["splitmix", 64, 2 << 20]],
    # file sizes at the segment size and around it
    [["splitmix", 65, SEG]], [["splitmix", 66, SEG - 1]], [["splitmix", 67, SEG + 1]], [["splitmix", 68, 3 * SEG]],
    [["zeros", 3 * SEG + 4097]], [["mod251", 2 * SEG + 100]],
    # forced cuts that straddle a segment end exactly
    [["zeros", SEG - 131072], ["splitmix", 69, 131072 + 9000]],
], ids=lambda s: "+".join(f"{seg[0]}{seg[-1]}" for seg in s))
def test_boundary_speculation_paths(hasher, spec):
    data = xet_inputs.build(spec)
    cl = hasher.hash_bytes(data, want_chunks=True)
    check_against_oracle(cl, 0, data)


def test_nonzero_first_offset_and_unaligned_views(hasher):
    data = xet_inputs.splitmix_bytes(77, (3 << 20) + 123)
    for lead in (1, 7, 16, 33):
        cl = hasher.hash_buffers(data, [lead, data.size], want_chunks=True)
        check_against_oracle(cl, 0, data[lead:])


def test_device_resident_input(hasher):
    import torch
    data = xet_inputs.build([["splitmix", 5, (9 << 20) + 17]])
    t = torch.from_numpy(data).cuda()
    cl = hasher.hash_buffers(t, [0, data.size], want_chunks=True)
    check_against_oracle(cl, 0, data)
    # unaligned device view (base + 3 bytes) takes the device-to-device realignment path
    cl = hasher.hash_buffers(t[3:], [0, data.size - 3], want_chunks=True)
    check_against_oracle(cl, 0, data[3:])


def test_fill_splitmix_matches_host_generator(hasher):
    import torch
    n = (1 << 20) + 13
    buf = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
    for seed, dst_off, stream_off in [(1, 0, 0), (9, 3, 0), (2**64 - 3, 8, 5), (12, 1, 7)]:
        hasher.fill_splitmix(buf.data_ptr() + dst_off, seed, stream_off, n)
        torch.cuda.synchronize()
        got = buf[dst_off:dst_off + n].cpu().numpy()
        assert np.array_equal(got, xet_inputs.splitmix_bytes(seed, n, stream_off))


def test_merkle_from_chunks(hasher):
    rng = np.random.default_rng(5)
    for n in [0, 1, 2, 3, 4, 8, 9, 10, 11, 19, 100, 1000, 8192, 50001]:
        hashes = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        sizes = rng.integers(1, 131073, n).astype(np.uint64)
        root, fh = hasher.merkle_from_chunks(hashes, sizes)
        if n:
            assert root == xet_oracle.merkle_root(hashes, sizes)
        assert fh == xet_oracle.file_hash_from_chunks(hashes, sizes)
    # sizes beyond 32 bits (upper merkle levels of > 4 GiB files print > 10 decimal digits)
    hashes = rng.integers(0, 256, (40, 32), dtype=np.uint8)
    sizes = rng.integers(1 << 33, 1 << 40, 40).astype(np.uint64)
    root, fh = hasher.merkle_from_chunks(hashes, sizes)
    assert root == xet_oracle.merkle_root(hashes, sizes)
    # uniform fan-out: every hash ends a group / no hash ends a group
    for byte24 in (0, 1):
        hashes = np.full((300, 32), byte24, dtype=np.uint8)
        sizes = np.full(300, 8192, dtype=np.uint64)
        assert hasher.merkle_from_chunks(hashes, sizes)[0] == xet_oracle.merkle_root(hashes, sizes)


@pytest.mark.parametrize("nranges", [2, 3, 5])
def test_byte_range_partition_reassembles(hasher, nranges):
    """BASELINE config 4 shape on one GPU: ranges processed in order with the carried cut."""
    data = xet_inputs.build([["splitmix", 31, 3 << 20], ["zeros", 400000], ["splitmix", 32, (2 << 20) + 999]])
    n = data.size
    ref = xet_oracle.hash_file(data)
    bounds = [n * i // nranges for i in range(nranges + 1)]
    bounds = [b + (17 if 0 < b < n else 0) for b in bounds]   # unaligned split points
    entry, ends, hashes = 0, [], []
    for r in range(nranges):
        b0, b1 = bounds[r], bounds[r + 1]
        lead = min(b0, 131072 + 64)
        piece = data[b0 - lead:b1]
        spec_exit = hasher.range_exit_cut(piece, n, b0, b1, entry)
        res = hasher.hash_range(piece, n, b0, b1, entry)
        assert res.exit_cut == spec_exit
        assert all(b0 < int(e) <= b1 for e in res.chunk_ends)
        ends.append(res.chunk_ends)
        hashes.append(res.chunk_hashes)
        entry = res.exit_cut
    ends = np.concatenate(ends)
    hashes = np.concatenate(hashes)
    assert np.array_equal(ends, ref.chunk_ends)
    assert np.array_equal(hashes, ref.chunk_hashes)
    sizes = np.diff(np.concatenate([[0], ends])).astype(np.uint64)
    root, fh = hasher.merkle_from_chunks(hashes, sizes)
    assert root == ref.merkle_root and fh == ref.file_hash


def test_hash_files_and_verify(hasher, tmp_path):
    import ome_b200
    specs = {"a.safetensors": [["splitmix", 90, (5 << 20) + 3]], "empty.bin": [], "small.json": [["ascii", "{}"]],
             "zeros.bin": [["zeros", 300000]]}
    paths, datas = [], []
    for name, spec in specs.items():
        d = xet_inputs.build(spec)
        p = tmp_path / name
        d.tofile(p)
        paths.append(str(p))
        datas.append(d)
    seen = []
    hasher.set_progress_handler(lambda u: seen.append(u), throttle_ms=0)
    cl = hasher.hash_files(paths, want_chunks=True)
    hasher.set_progress_handler(None)
    assert seen and seen[-1]["completed_bytes"] == sum(d.size for d in datas)
    for f, d in enumerate(datas):
        hexhash = check_against_oracle(cl, f, d)
        hasher.verify_file(paths[f], hexhash)
    assert hasher.hash_file(paths[0]) == cl.files[0].hex
    bad = "0" * 64
    with pytest.raises(ome_b200.XetGpuError) as ei:
        hasher.verify_file(paths[0], bad)
    assert ei.value.code == 6 and cl.files[0].hex in ei.value.details   # ChecksumMismatch
    with pytest.raises(ome_b200.XetGpuError) as ei:
        hasher.verify_file(str(tmp_path / "missing"), bad)
    assert ei.value.code == 4                                           # NotFound
    with pytest.raises(ome_b200.XetGpuError) as ei:
        hasher.verify_buffer(datas[0], "xyz")
    assert ei.value.code == 1                                           # InvalidConfig
    hasher.verify_buffer(datas[0], cl.files[0].hex)
    # snapshot verification in one batch: all good / one skipped / one wrong
    good = [f.hex for f in cl.files]
    assert hasher.verify_files(paths, good) == [True] * len(paths)
    assert hasher.verify_files(paths, [good[0], None, "", good[3]]) == [True] * len(paths)
    with pytest.raises(ome_b200.XetGpuError) as ei:
        hasher.verify_files(paths, [good[0], good[1], bad, good[3]])
    assert ei.value.code == 6 and "1 of 4" in ei.value.message and paths[2] in ei.value.details
    assert [bool(x) for x in hasher._last_ok][:4] == [True, True, False, True]
    # small batches force several device round trips
    small = ome_b200.XetHasher(staging_bytes=1 << 20, max_batch_bytes=1 << 20)
    cl2 = small.hash_files(paths, want_chunks=True)
    small.close()
    for f, d in enumerate(datas):
        check_against_oracle(cl2, f, d)
    assert hasher.hash_files(paths, should_cancel=lambda: False).files[0].hex == cl.files[0].hex
    with pytest.raises(ome_b200.XetGpuError) as ei:
        hasher.hash_files(paths, should_cancel=lambda: True)
    assert ei.value.code == 7                                           # Cancelled


def test_invalid_arguments(hasher):
    import ome_b200
    with pytest.raises(ome_b200.XetGpuError) as ei:
        hasher.hash_buffers(np.zeros(10, np.uint8), [5, 2])
    assert ei.value.code == 1
    with pytest.raises(ome_b200.XetGpuError) as ei:
        hasher.hash_range(np.zeros(100, np.uint8), 100, 50, 40, 0)
    assert ei.value.code == 1


@pytest.mark.parametrize("case", LARGE, ids=[c["name"] for c in LARGE])
def test_large_golden_cases_device_generated(hasher, case):
    """hf_xet golden vectors too big to ship as bytes, rebuilt on the device: the 1 GiB BASELINE
    config-1 size, and N > 2^32 (u64 offsets everywhere, > 10-digit sizes in upper merkle levels)."""
    import torch
    n = case["n"]
    buf = torch.zeros(n + 256, dtype=torch.uint8, device="cuda")
    pos = 0
    for seg in case["spec"]:
        if seg[0] == "splitmix":
            hasher.fill_splitmix(buf.data_ptr() + pos, seg[1], 0, seg[2])
            pos += seg[2]
        elif seg[0] == "zeros":
            pos += seg[1]
        else:
            raise AssertionError("large cases are built from splitmix and zeros only")
    assert pos == n
    cl = hasher.hash_buffers(buf[:n], [0, n], want_chunks=True)
    assert cl.files[0].hex == case["file_hash"]                # hf_xet ground truth
    assert cl.files[0].n_chunks == case["n_chunks"]
    ends, hashes = cl.chunks_of(0)
    assert [int(x) for x in ends[:16]] == case["cuts_head"]
    assert hashlib.sha256(ends.astype("<u8").tobytes()).hexdigest() == case["cuts_sha256"]
    assert hashlib.sha256(hashes.tobytes()).hexdigest() == case["chunk_hashes_sha256"]


def test_full_size_shard_properties(hasher):
    """1 GiB device-resident shard (BASELINE config 1 size), checked through size-independent
    properties: (a) the two halves hashed as byte ranges with a carried cut reassemble to the
    whole-file chunk list; (b) merkle of the returned chunk list reproduces the file hash;
    (c) chunk sizes respect min/max; (d) a 64 MiB prefix agrees with the oracle."""
    import torch
    n = 1 << 30
    buf = torch.empty(n, dtype=torch.uint8, device="cuda")
    hasher.fill_splitmix(buf.data_ptr(), 1, 0, n)
    cl = hasher.hash_buffers(buf, [0, n], want_chunks=True)
    ends, hashes = cl.chunks_of(0)
    sizes = np.diff(np.concatenate([[0], ends])).astype(np.uint64)
    assert int(ends[-1]) == n and sizes.max() <= 131072 and sizes[:-1].min() >= 8192
    root, fh = hasher.merkle_from_chunks(hashes, sizes)
    assert root == cl.files[0].merkle_root and fh == cl.files[0].file_hash
    # (a) two ranges
    mid = (n // 2) + 12345
    lead = 131072 + 64
    r0 = hasher.hash_range(buf, n, 0, mid, 0)
    r1 = hasher.hash_range(buf[mid - lead:], n, mid, n, r0.exit_cut)
    assert np.array_equal(np.concatenate([r0.chunk_ends, r1.chunk_ends]), ends)
    assert np.array_equal(np.concatenate([r0.chunk_hashes, r1.chunk_hashes]), hashes)
    # (d) prefix parity: first 64 MiB is golden case splitmix_seed1_64MiB
    case = next(c for c in SMALL if c["name"] == "splitmix_seed1_64MiB")
    k = case["n_chunks"] - 1      # all but the prefix's last (EOF-flushed) chunk are shared
    assert [int(x) for x in ends[:16]] == case["cuts_head"]
    pref = xet_oracle.hash_file(xet_inputs.splitmix_bytes(1, 64 << 20))
    assert np.array_equal(ends[:k], pref.chunk_ends[:k])
    assert np.array_equal(hashes[:k], pref.chunk_hashes[:k])


def test_xetsum_cli_cpp_consumer(tmp_path):
    """tools/xetsum: a plain C++ program over the C ABI (the role of OME's cgo wrappers)."""
    import subprocess
    from conftest import ROOT
    exe = os.path.join(ROOT, "tools", "xetsum")
    if not os.path.exists(exe):
        pytest.skip("tools/xetsum not built")
    cases = {c["name"]: c for c in SMALL}
    names = ["hello_world", "empty", "zeros_131073", "splitmix_seed1_1MiB"]
    paths = []
    for n in names:
        p = tmp_path / (n + ".bin")
        xet_inputs.build(cases[n]["spec"]).tofile(p)
        paths.append(str(p))
    out = subprocess.run([exe] + paths, capture_output=True, text=True, check=True).stdout.splitlines()
    assert [l.split("  ")[0] for l in out] == [cases[n]["file_hash"] for n in names]
    listing = tmp_path / "sums.txt"
    listing.write_text("\n".join(out) + "\n")
    ok = subprocess.run([exe, "-c", str(listing)], capture_output=True, text=True)
    assert ok.returncode == 0 and ok.stdout.count(": OK") == len(names)
    listing.write_text("0" * 64 + "  " + paths[0] + "\n")
    bad = subprocess.run([exe, "-c", str(listing)], capture_output=True, text=True)
    assert bad.returncode == 1 and "FAILED" in bad.stdout
    ch = subprocess.run([exe, "--chunks", paths[3]], capture_output=True, text=True, check=True).stdout.splitlines()
    assert len(ch) == 1 + cases["splitmix_seed1_1MiB"]["n_chunks"]
    assert ch[1].split() == ["25125", cases["splitmix_seed1_1MiB"]["chunk_hashes"][0]]


@pytest.fixture()
def small_group_hasher(monkeypatch):
    """A context whose file groups are tiny, so that ordinary test batches take the two-lane path
    (groups alternating between two streams, unaligned group starts, staggered gear passes)."""
    import ome_b200
    monkeypatch.setenv("XETGPU_MIN_GROUP_BYTES", str(1 << 20))
    h = ome_b200.XetHasher(stage_timing=True)
    yield h
    h.close()


@pytest.mark.parametrize("where", ["host", "device"])
def test_two_lane_file_groups(small_group_hasher, where):
    import torch
    rng = np.random.default_rng(23)
    # >= 16 files: device-resident batches take the two-lane path too
    sizes = [int(x) for x in np.exp(rng.uniform(np.log(1000), np.log(6e6), 40))] + [0, 1, 8191, 3 << 20, 5 << 20]
    datas = [xet_inputs.splitmix_bytes(1500 + i, n) for i, n in enumerate(sizes)]
    # one group of dense candidates in the middle: its candidate list overflows and is re-run alone
    datas[20] = xet_inputs.build([["repeat", xet_inputs.W_HEX, 30000]])
    sizes[20] = datas[20].size
    offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
    packed = np.concatenate(datas)
    src = torch.from_numpy(packed).cuda() if where == "device" else packed
    cl = small_group_hasher.hash_buffers(src, offsets, want_chunks=True)
    assert len(cl.files) == len(sizes)
    for f, d in enumerate(datas):
        check_against_oracle(cl, f, d)
    t = small_group_hasher.last_timings()
    assert t["n_chunks"] == sum(f.n_chunks for f in cl.files) and t["bytes"] == packed.size
    # and without chunk lists (first_chunk bookkeeping differs)
    cl2 = small_group_hasher.hash_buffers(src, offsets)
    assert [f.file_hash for f in cl2.files] == [f.file_hash for f in cl.files]


def test_randomised_inputs_against_oracle(hasher):
    """Seeded fuzz: files stitched from random, zero, periodic and dense-candidate pieces of random
    lengths, hashed in one batch; every boundary and hash must equal the oracle's."""
    rng = np.random.default_rng(20261106)
    kinds = ["splitmix", "zeros", "mod251", "W"]
    datas = []
    for i in range(48):
        spec = []
        for _ in range(int(rng.integers(1, 5))):
            k = kinds[int(rng.integers(0, 4))]
            n = int(np.exp(rng.uniform(np.log(1), np.log(2.5e6))))
            if k == "splitmix":
                spec.append(["splitmix", int(rng.integers(0, 1 << 30)), n])
            elif k == "W":
                spec.append(["repeat", xet_inputs.W_HEX, max(1, n // 64)])
            else:
                spec.append([k, n])
        datas.append(xet_inputs.build(spec))
    offsets = np.concatenate([[0], np.cumsum([d.size for d in datas])]).astype(np.uint64)
    cl = hasher.hash_buffers(np.concatenate(datas), offsets, want_chunks=True)
    for f, d in enumerate(datas):
        check_against_oracle(cl, f, d)


def test_context_shared_by_threads(hasher):
    """Calls on one context serialise (pkg/xet/README.md:76,88 says the same of the reference client)."""
    import threading
    datas = [xet_inputs.splitmix_bytes(3000 + i, (1 << 20) + 1000 * i) for i in range(8)]
    want = [xet_oracle.hash_file(d).file_hash for d in datas]
    got = [None] * len(datas)

    def work(i):
        got[i] = hasher.hash_bytes(datas[i]).files[0].file_hash

    ts = [threading.Thread(target=work, args=(i,)) for i in range(len(datas))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert got == want


def test_degenerate_batches(hasher):
    """No files at all; only empty files; an empty file between non-empty ones."""
    cl = hasher.hash_buffers(b"", [0], want_chunks=True)
    assert cl.files == [] and cl.chunk_ends.size == 0
    cl = hasher.hash_buffers(b"", [0, 0, 0, 0], want_chunks=True)
    assert [f.file_hash for f in cl.files] == [bytes(32)] * 3 and all(f.n_chunks == 0 for f in cl.files)
    d = xet_inputs.splitmix_bytes(9, 300000)
    cl = hasher.hash_buffers(np.concatenate([d, d]), [0, 300000, 300000, 600000], want_chunks=True)
    check_against_oracle(cl, 0, d)
    assert cl.files[1].n_chunks == 0 and cl.files[1].file_hash == bytes(32)
    check_against_oracle(cl, 2, d)
    assert hasher.gear_candidates(b"").size == 0


def test_cpp_host_mirror(tmp_path):
    """include/xetgpu.hpp (GpuHasher: HashFiles / HashBytes / VerifyFile / VerifyFiles, XetError) — the C++
    stand-in for the Go wrappers, checked by tests/cpp/host_test.cpp against hf_xet golden values."""
    import subprocess
    from conftest import ROOT
    exe = os.path.join(ROOT, "tools", "xetgpu_host_test")
    if not os.path.exists(exe):
        pytest.skip("tools/xetgpu_host_test not built")
    r = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("ok")


def test_reference_fixture_files_on_gpu(hasher):
    """The reference's own safetensors fixtures (pkg/hfutil/modelconfig/testdata/tiny-random-PhiModel), shipped as
    bytes under tests/golden/reference_fixtures (model-1-of-2 and model-2-of-2 are byte-identical), hashed by the GPU
    path through xetgpu_hash_files and compared with the hf_xet hashes recorded in the golden file."""
    fix = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_fixtures")
    exp = GOLDEN["reference_fixture_files"]
    base = "pkg/hfutil/modelconfig/testdata/tiny-random-PhiModel/"
    paths = [os.path.join(fix, "model-1-of-2.safetensors"), os.path.join(fix, "config.json")]
    cl = hasher.hash_files(paths, want_chunks=True)
    assert cl.files[0].hex == exp[base + "model-1-of-2.safetensors"]["file_hash"] == exp[base + "model-2-of-2.safetensors"]["file_hash"]
    assert cl.files[1].hex == exp[base + "config.json"]["file_hash"]
    assert cl.files[0].size == exp[base + "model-1-of-2.safetensors"]["n"]
    for f, p in enumerate(paths):
        check_against_oracle(cl, f, np.fromfile(p, dtype=np.uint8))


def _oracle_xorbs(data):
    ref = xet_oracle.hash_file(data)
    sizes = np.diff(np.concatenate([[0], ref.chunk_ends]).astype(np.uint64)).astype(np.uint64)
    starts, roots = xet_oracle.xorbs(ref.chunk_hashes, sizes)
    return ref, sizes, starts, roots


def _check_xorbs(cl, datas):
    """XETGPU_WANT_XORBS output vs the oracle twin (oracle/xet_oracle.c xo_xorbs).  Grouping policy parity is
    unpinned (hf_xet does not expose xorbs); the xorb hash function is the pinned aggregated_node_hash."""
    x = cl.xorbs
    pos = 0
    for f, d in enumerate(datas):
        ref, sizes, starts, roots = _oracle_xorbs(d)
        n = roots.shape[0]
        mine = x[pos:pos + n] if n else x[:0]
        assert n == 0 or (mine["file"] == f).all(), f
        assert [int(v) for v in mine["first_chunk"]] == [int(v) for v in starts[:-1]], f
        assert [int(v) for v in mine["n_chunks"]] == [int(starts[i + 1] - starts[i]) for i in range(n)], f
        assert [int(v) for v in mine["bytes"]] == [int(sizes[int(starts[i]):int(starts[i + 1])].sum()) for i in range(n)], f
        assert np.array_equal(mine["hash"], roots), f
        assert all(int(b) <= (64 << 20) and int(c) <= 8192 for b, c in zip(mine["bytes"], mine["n_chunks"]))
        pos += n
    assert pos == (0 if x is None else x.shape[0])


def test_xorb_formation_matches_oracle(hasher, tmp_path):
    specs = [[["splitmix", 61, (150 << 20) + 3]],            # three xorbs, the byte limit binds
             [],                                              # empty file: no xorb
             [["splitmix", 62, 70001]],                       # one xorb
             [["repeat", xet_inputs.W_HEX, (80 << 20) // 64]],  # candidates everywhere: 8192-byte chunks, both limits bind
             [["ascii", "x"]]]                                # single one-byte chunk: xorb hash == chunk hash
    datas = [xet_inputs.build(s) for s in specs]
    off = np.concatenate([[0], np.cumsum([d.size for d in datas])]).astype(np.uint64)
    cl = hasher.hash_buffers(np.concatenate(datas), off, want_chunks=True, want_xorbs=True)
    for f, d in enumerate(datas):
        check_against_oracle(cl, f, d)
    _check_xorbs(cl, datas)
    assert cl.xorbs.shape[0] >= 6
    # the same through the file entry point, without the chunk lists
    paths = []
    for i, d in enumerate(datas):
        p = tmp_path / f"x{i}.bin"
        d.tofile(p)
        paths.append(str(p))
    small = __import__("ome_b200").XetHasher(max_batch_bytes=200 << 20)   # two batches: file indices are remapped
    try:
        cl2 = small.hash_files(paths, want_xorbs=True)
    finally:
        small.close()
    _check_xorbs(cl2, datas)
    assert hasher.hash_files(paths).xorbs is None


def test_whole_files_and_byte_range_items_in_one_batch(hasher):
    """xetgpu_hash_buffers_ranges: what a rank of a byte-range partitioned job holds, in ONE batched call."""
    import ome_b200
    LEAD = ome_b200.hasher.RANGE_LEAD
    A = xet_inputs.splitmix_bytes(71, (3 << 20) + 17)
    B = xet_inputs.splitmix_bytes(72, (9 << 20) + 5)
    Cc = xet_inputs.build([["splitmix", 73, 1 << 20], ["zeros", 4 << 20]])
    refB, refC = xet_oracle.hash_file(B), xet_oracle.hash_file(Cc)

    def true_entry(ref, begin):   # last cut <= begin
        e = [int(x) for x in ref.chunk_ends if x <= begin]
        return e[-1] if e else 0

    b1, b2 = 3 << 20, 6 << 20
    items = [  # (data, spec, ref, file bytes)
        (A, (A.size, 0, A.size, 0), None),
        (B[b1 - LEAD:b2], (B.size, b1, b2, true_entry(refB, b1)), refB),            # middle range, true entry
        (np.zeros(0, np.uint8), (0, 0, 0, 0), None),                               # empty whole file
        (Cc[:2 << 20], (Cc.size, 0, 2 << 20, 0), refC),                            # first range: not the end of the file
        (B[b2 - LEAD:], (B.size, b2, B.size, true_entry(refB, b2)), refB),          # last range: flushes the tail
        (B[b1 - LEAD:b2], (B.size, b1, b2, b1), None),                              # speculative entry (a cut assumed at begin)
    ]
    # items packed back to back (byte-range items start at arbitrary alignment inside the batch)
    packed = np.concatenate([d for d, _, _ in items])
    poff = np.concatenate([[0], np.cumsum([d.size for d, _, _ in items])]).astype(np.uint64)
    cl, exits = hasher.hash_buffers_ranges(packed, poff, [s for _, s, _ in items], want_chunks=True)
    check_against_oracle(cl, 0, A)
    # without want_chunks only the byte-range items carry chunk lists; hashes and exit cuts are the same
    lean, exits_lean = hasher.hash_buffers_ranges(packed, poff, [s for _, s, _ in items])
    assert exits_lean == exits and lean.files[0].file_hash == cl.files[0].file_hash
    assert lean.chunk_ends.size == sum(cl.files[k].n_chunks for k in (1, 3, 4, 5))
    for k in (1, 3, 4, 5):
        assert np.array_equal(lean.chunks_of(k)[0], cl.chunks_of(k)[0]) and np.array_equal(lean.chunks_of(k)[1], cl.chunks_of(k)[1])
    assert cl.files[2].n_chunks == 0 and cl.files[2].file_hash == bytes(32)
    for k, (d, (flen, begin, end, entry), ref) in enumerate(items):
        if ref is None:
            continue
        want = [(int(e), h) for e, h in zip(ref.chunk_ends, ref.chunk_hashes) if begin < e <= end]
        ends, hashes = cl.chunks_of(k)
        assert [int(e) for e in ends] == [e for e, _ in want], k
        assert np.array_equal(hashes, np.array([h for _, h in want]).reshape(-1, 32)), k
        assert exits[k] == (want[-1][0] if want else entry), k
    # the speculative item: same protocol result as the single-item entry point
    single = hasher.hash_range(B[b1 - LEAD:b2], B.size, b1, b2, b1)
    ends, hashes = cl.chunks_of(5)
    assert np.array_equal(ends, single.chunk_ends) and np.array_equal(hashes, single.chunk_hashes) and exits[5] == single.exit_cut
    # device-resident input gives the same answer
    import torch
    dev = torch.from_numpy(packed).cuda()
    cl2, exits2 = hasher.hash_buffers_ranges(dev, poff, [s for _, s, _ in items], memory=ome_b200.hasher.MEM_DEVICE,
                                             want_chunks=True)
    assert exits2 == exits and np.array_equal(cl2.chunk_hashes, cl.chunk_hashes) and np.array_equal(cl2.chunk_ends, cl.chunk_ends)
    with pytest.raises(ome_b200.XetGpuError) as ei:
        hasher.hash_buffers_ranges(packed, poff, [(1, 0, 1, 0)] * len(items))
    assert ei.value.code == 1
