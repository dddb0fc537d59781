"""xetgpu_multi_* (several GPUs from one process) and the robustness rules of the file entry points, through the C ABI.

On a single-GPU box the multi-device context is given the same device several times (one worker each), which
exercises exactly the code that runs across real devices: the end-to-end layout, shared files with the carried cut,
fix-up window, splice and merkle combine.  With two or more GPUs the same test also runs over all of them."""
import os
import resource

import numpy as np
import pytest

import xet_inputs
import xet_oracle

pytestmark = pytest.mark.gpu


def _write(tmp_path, specs):
    paths, datas = [], []
    for i, spec in enumerate(specs):
        d = xet_inputs.build(spec)
        p = tmp_path / f"f{i}.bin"
        d.tofile(p)
        paths.append(str(p))
        datas.append(d)
    return paths, datas


def _check(cl, datas):
    for f, d in enumerate(datas):
        ref = xet_oracle.hash_file(d)
        fr = cl.files[f]
        assert fr.size == d.size and fr.n_chunks == ref.chunk_ends.size, f
        assert fr.file_hash == ref.file_hash, f
        if ref.chunk_ends.size:
            assert fr.merkle_root == ref.merkle_root, f
        if cl.chunk_ends is not None:
            ends, hashes = cl.chunks_of(f)
            assert np.array_equal(ends, ref.chunk_ends) and np.array_equal(hashes, ref.chunk_hashes), f


SPECS = [[["splitmix", 41, (9 << 20) + 1]], [], [["splitmix", 42, 11 << 20]], [["splitmix", 43, 70001]],
         [["splitmix", 44, 1 << 20], ["zeros", 6 << 20]],          # shared-file chains that cannot re-merge
         [["splitmix", 45, (10 << 20) + 77]]]


@pytest.mark.parametrize("devices", [[0], [0, 0], [0, 0, 0], None], ids=["one", "two-workers", "three-workers", "all-gpus"])
def test_multi_hash_files_matches_oracle(tmp_path, monkeypatch, devices):
    import ome_b200
    monkeypatch.setenv("XETGPU_MULTI_MIN_BYTES", str(1 << 20))
    paths, datas = _write(tmp_path, SPECS)
    with ome_b200.XetMultiHasher(devices) as m:
        seen = []
        m.set_progress_handler(seen.append, throttle_ms=0)
        cl = m.hash_files(paths, want_chunks=True)
        _check(cl, datas)
        assert seen and seen[-1]["completed_bytes"] == sum(d.size for d in datas) and seen[-1]["phase"] == 2
        _check(m.hash_files(paths), datas)
        good = [f.hex for f in cl.files]
        assert m.verify_files(paths, good) == [True] * len(paths)
        with pytest.raises(ome_b200.XetGpuError) as ei:
            m.verify_files(paths, good[:2] + ["0" * 64] + good[3:])
        assert ei.value.code == 6 and paths[2] in ei.value.details
        with pytest.raises(ome_b200.XetGpuError) as ei:
            m.hash_files(paths, should_cancel=lambda: True)
        assert ei.value.code == 7
        with pytest.raises(ome_b200.XetGpuError) as ei:
            m.hash_files(paths + [str(tmp_path / "missing")])
        assert ei.value.code == 4


def test_multi_one_large_file_split_over_workers(tmp_path, monkeypatch):
    """One file shared by every worker (the single-huge-checkpoint case)."""
    import ome_b200
    monkeypatch.setenv("XETGPU_MULTI_MIN_BYTES", str(1 << 20))
    paths, datas = _write(tmp_path, [[["splitmix", 46, (48 << 20) + 12345]]])
    with ome_b200.XetMultiHasher([0, 0, 0, 0]) as m:
        _check(m.hash_files(paths, want_chunks=True), datas)


def test_more_files_than_open_file_limit(hasher, tmp_path):
    """Files are opened piece by piece: a batch may name more files than RLIMIT_NOFILE allows to be open at once."""
    specs = [[["splitmix", 500 + i, 3000 + 37 * i]] for i in range(160)]
    paths, datas = _write(tmp_path, specs)
    soft, hard = resource.getrlimit(resource.RLIMIT_NOFILE)
    in_use = len(os.listdir("/proc/self/fd"))
    resource.setrlimit(resource.RLIMIT_NOFILE, (in_use + 40, hard))
    try:
        cl = hasher.hash_files(paths)
    finally:
        resource.setrlimit(resource.RLIMIT_NOFILE, (soft, hard))
    _check(cl, datas)


def test_file_resized_during_hash_is_an_io_error(hasher, tmp_path):
    import ome_b200
    p = tmp_path / "big.bin"
    n = 1 << 30
    with open(p, "wb") as f:
        f.truncate(n)                      # sparse: reads as zeros
    state = {"done": False}

    def shrink(_u):
        if not state["done"]:
            state["done"] = True
            os.truncate(p, n // 2)

    hasher.set_progress_handler(shrink, throttle_ms=0)
    try:
        with pytest.raises(ome_b200.XetGpuError) as ei:
            hasher.hash_files([str(p)])
    finally:
        hasher.set_progress_handler(None)
    assert ei.value.code == 8 and state["done"]          # IoError: size changed / file shrank


def test_reentrant_call_from_progress_handler_is_refused(hasher, tmp_path):
    import ome_b200
    paths, datas = _write(tmp_path, [[["splitmix", 47, 40 << 20]]])
    got = {}

    def handler(_u):
        if "err" in got:
            return
        try:
            hasher.hash_bytes(b"abc")                     # hashing entry point of the context that is running us
            got["err"] = None
        except ome_b200.XetGpuError as e:
            got["err"] = e
        got["timings"] = hasher.last_timings()            # ... while the bookkeeping calls are allowed

    hasher.set_progress_handler(handler, throttle_ms=0)
    try:
        cl = hasher.hash_files(paths)
    finally:
        hasher.set_progress_handler(None)
    _check(cl, datas)
    assert got["err"] is not None and got["err"].code == 1 and "re-entrant" in got["err"].message
    assert "total_ms" in got["timings"]


def test_shared_safetensors_shard_is_cut_on_tensor_boundaries(tmp_path, monkeypatch):
    """f4: a *.safetensors shard shared between workers has its in-file boundaries on tensor boundaries (plan), and
    its hash does not depend on where they fall."""
    import ome_b200
    from ome_b200 import safetensors as st
    from test_safetensors import shard
    monkeypatch.setenv("XETGPU_MULTI_MIN_BYTES", str(1 << 20))
    mib = 1 << 20
    tensors = [(f"model.layers.{i}.w", "BF16", [mib // 2 * (1 + i % 3)], mib * (1 + i % 3)) for i in range(20)]
    raw = bytearray(shard(tensors))
    body = xet_inputs.splitmix_bytes(48, len(raw))
    hdr_len = 8 + int.from_bytes(raw[:8], "little")
    raw[hdr_len:] = body[hdr_len:].tobytes()             # random payload behind a real header
    p = tmp_path / "model-00001-of-00001.safetensors"
    p.write_bytes(bytes(raw))
    other = tmp_path / "tokenizer.json"
    other.write_bytes(b"{}")
    ix = st.open_index(str(p))
    marks = {t.begin for t in ix.tensors} | {t.end for t in ix.tensors}
    data = np.frombuffer(bytes(raw), dtype=np.uint8)
    with ome_b200.XetMultiHasher([0, 0, 0]) as m:
        plan = m.plan([str(p), str(other)])
        cuts = sorted({b for _, f, b, e in plan if f == 0 and b > 0})
        assert len(cuts) == 2 and all(c in marks for c in cuts), (cuts, sorted(marks)[:5])
        raw_cuts = sorted({b for _, f, b, e in m.plan([str(p), str(other)], tensor_snap=False) if f == 0 and b > 0})
        assert raw_cuts != cuts and all(c % 4096 == 0 for c in raw_cuts)
        cl = m.hash_files([str(p), str(other)], want_chunks=True)
    _check(cl, [data, np.frombuffer(b"{}", dtype=np.uint8)])


def test_file_larger_than_the_device_budget_streams_in_pieces(tmp_path):
    """max_batch_bytes bounds device memory: a bigger file is hashed piece by piece (byte ranges, carried cut)."""
    import ome_b200
    specs = [[["splitmix", 49, (40 << 20) + 4321]], [["splitmix", 50, 3 << 20]], [["zeros", 20 << 20]], []]
    paths, datas = _write(tmp_path, specs)
    h = ome_b200.XetHasher(max_batch_bytes=8 << 20)
    try:
        seen = []
        h.set_progress_handler(seen.append, throttle_ms=0)
        cl = h.hash_files(paths, want_chunks=True)
        h.set_progress_handler(None)
        _check(cl, datas)
        done = [u["completed_bytes"] for u in seen]
        assert done == sorted(done) and done[-1] == sum(d.size for d in datas)
        _check(h.hash_files(paths), datas)
        h.verify_file(paths[0], cl.files[0].hex)
    finally:
        h.close()


def test_multi_reentrant_call_from_handler_is_refused(tmp_path, monkeypatch):
    import ome_b200
    monkeypatch.setenv("XETGPU_MULTI_MIN_BYTES", str(1 << 20))
    paths, datas = _write(tmp_path, [[["splitmix", 51, 24 << 20]], [["splitmix", 52, 9 << 20]]])
    got = {}
    with ome_b200.XetMultiHasher([0, 0]) as m:
        def handler(_u):
            if "err" not in got:
                try:
                    m.hash_files(paths[:1])
                    got["err"] = None
                except ome_b200.XetGpuError as e:
                    got["err"] = e

        m.set_progress_handler(handler, throttle_ms=0)
        cl = m.hash_files(paths)
        m.set_progress_handler(None)
        _check(cl, datas)
    assert got["err"] is not None and got["err"].code == 1 and "re-entrant" in got["err"].message


def test_in_process_plan_equals_the_multi_process_plan(tmp_path, monkeypatch):
    """xetgpu_multi_plan (C++, one process) lays files out exactly as sharded.plan_concatenated (Python, one process
    per GPU) does: the two forms of the config-4 partition cannot drift apart."""
    import ome_b200
    from ome_b200 import sharded
    monkeypatch.setenv("XETGPU_MULTI_MIN_BYTES", "1")
    rng = np.random.default_rng(11)
    for trial, sizes in enumerate([[3 << 20, 0, 5 << 20, 70001, 2 << 20], [20 << 20], [6 << 20] * 4,
                                   [int(x) for x in rng.integers(0, 3_000_000, 23)], [0, 0], [1 << 20, 131136 + 4096, 7]]):
        d = tmp_path / f"t{trial}"
        d.mkdir()
        paths = []
        for i, n in enumerate(sizes):
            p = d / f"f{i}.bin"
            with open(p, "wb") as f:
                f.truncate(n)               # sparse: only the sizes matter for the plan
            paths.append(str(p))
        for world in (1, 2, 3, 5):
            with ome_b200.XetMultiHasher([0] * world) as m:
                got = m.plan(paths, tensor_snap=False)
            used = min(world, max(1, sum(sizes)))    # the library uses fewer devices than listed for tiny jobs
            want = [(r, p.file, p.begin, p.end) for r, pieces in enumerate(sharded.plan_concatenated(sizes, used))
                    for p in pieces]
            assert got == want, (trial, world)
