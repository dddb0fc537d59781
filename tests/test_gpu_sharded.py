"""Two-rank NCCL run of the multi-GPU host logic with the real CUDA engine (needs >= 2 GPUs;
skipped on a single-GPU box — the CPU gloo tests in test_sharded_cpu.py cover the same logic)."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, ret):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (here, os.path.join(os.path.dirname(here), "oracle"), os.path.dirname(here)):
        if p not in sys.path:
            sys.path.insert(0, p)
    import ome_b200
    from ome_b200 import sharded
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        h = ome_b200.XetHasher(device=rank)
        # one 1.5 GiB splitmix file, byte-range partitioned; every rank generates only its own range
        n = (3 << 29) + 4321
        ranges = sharded.split_byte_ranges(n, world)
        begin, end = ranges[rank]
        cap = 1 << 18                      # headroom in front of the range, 16-byte aligned
        piece = torch.zeros(cap + (end - begin) + 256, dtype=torch.uint8, device=dev)
        h.fill_splitmix(piece.data_ptr() + cap, 7, begin, end - begin)
        lead = sharded.exchange_lead(piece[: cap + (end - begin)], cap, ranges)
        out = sharded.hash_file_by_ranges(h, piece[cap - lead: cap + (end - begin)], lead, n, begin, end,
                                          comm_device=dev)
        res = dict(rehashed=out.rehashed_bytes, n_own=int(out.chunk_ends.size))
        if rank == 0:
            whole = torch.empty(n + 256, dtype=torch.uint8, device=dev)
            h.fill_splitmix(whole.data_ptr(), 7, 0, n)
            cl = h.hash_buffers(whole, [0, n])
            res.update(fh=out.file_hash, root=out.merkle_root, n=out.n_chunks_total,
                       ref_fh=cl.files[0].file_hash, ref_root=cl.files[0].merkle_root, ref_n=cl.files[0].n_chunks)
            del whole
        # file partition: 5 files, LPT over the ranks
        sizes = [40 << 20, 3 << 20, 0, 25 << 20, 70001]
        part = sharded.FilePartition.lpt(sizes, world)
        local = []
        for i in part.assignment[rank]:
            t = torch.empty(sizes[i] + 256, dtype=torch.uint8, device=dev)
            h.fill_splitmix(t.data_ptr(), 900 + i, 0, sizes[i])
            local.append((i, t[:sizes[i]]))
        hashes, counts = sharded.hash_files_partitioned(h, local, len(sizes), comm_device=dev)
        res.update(file_hashes=hashes.tobytes(), counts=counts.tolist())
        ret[rank] = res
        h.close()
    finally:
        dist.destroy_process_group()


def test_two_rank_nccl_byte_ranges_and_file_partition():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
    r0, r1 = ret[0], ret[1]
    assert r0["fh"] == r0["ref_fh"] and r0["root"] == r0["ref_root"] and r0["n"] == r0["ref_n"]
    assert r0["n_own"] + r1["n_own"] == r0["n"]
    assert 0 < r1["rehashed"] < (64 << 20)                  # the carry fix-up re-hashes a window, not the range
    assert r0["file_hashes"] == r1["file_hashes"] and r0["counts"] == r1["counts"]
    import xet_inputs
    import xet_oracle
    hashes = np.frombuffer(r0["file_hashes"], dtype=np.uint8).reshape(-1, 32)
    for i, n in enumerate([40 << 20, 3 << 20, 0, 25 << 20, 70001]):
        if n <= (3 << 20):
            assert hashes[i].tobytes() == xet_oracle.hash_file(xet_inputs.splitmix_bytes(900 + i, n)).file_hash


def _ranges_on_one_gpu(rank, world, n, seed):
    """hash_file_by_ranges with the CUDA engine, every rank on GPU 0, exchanges over gloo (host tensors)."""
    import ome_b200
    import xet_inputs
    from ome_b200 import sharded
    dev = torch.device("cuda", 0)
    eng = ome_b200.XetHasher(device=0)
    ranges = sharded.split_byte_ranges(n, world)
    begin, end = ranges[rank]
    cap = 1 << 18
    host = torch.zeros(cap + (end - begin), dtype=torch.uint8)
    host[cap:] = torch.from_numpy(xet_inputs.splitmix_bytes(seed, n)[begin:end].copy())
    lead = sharded.exchange_lead(host, cap, ranges)                     # halo over gloo, on the host copy
    piece = torch.zeros(host.numel() + 256, dtype=torch.uint8, device=dev)
    piece[: host.numel()] = host.to(dev)
    out = sharded.hash_file_by_ranges(eng, piece[cap - lead: cap + (end - begin)], lead, n, begin, end,
                                      comm_device=torch.device("cpu"))
    eng.close()
    return dict(fh=out.file_hash, root=out.merkle_root, n=out.n_chunks_total, ends=out.chunk_ends.tolist(),
                hashes=out.chunk_hashes.tobytes(), rehashed=out.rehashed_bytes, range=(begin, end))


def test_three_range_protocol_on_one_gpu_matches_oracle():
    """The full byte-range protocol (speculative pass, carried cut, fix-up window, splice at the first common cut,
    rank-0 merkle combine) with three ranks sharing the one GPU of the test box, against the CPU oracle."""
    import xet_inputs
    import xet_oracle
    from test_sharded_cpu import run_ranks
    n, seed = (40 << 20) + 777, 86
    out = run_ranks(3, _ranges_on_one_gpu, n, seed)
    ref = xet_oracle.hash_file(xet_inputs.splitmix_bytes(seed, n))
    assert out[0]["fh"] == ref.file_hash and out[0]["root"] == ref.merkle_root and out[0]["n"] == ref.chunk_ends.size
    ends = [e for o in out for e in o["ends"]]
    assert ends == [int(x) for x in ref.chunk_ends]
    assert b"".join(o["hashes"] for o in out) == ref.chunk_hashes.tobytes()
    for o in out[1:]:                                                  # fix-up re-hashed a window, not the range
        assert 0 < o["rehashed"] <= (4 << 20) + 131136 < o["range"][1] - o["range"][0]


def test_concatenated_plan_on_one_gpu_matches_oracle():
    """BASELINE config-4 shape in miniature (files end to end, three byte ranges) with the CUDA engine."""
    import xet_inputs
    import xet_oracle
    from test_sharded_cpu import _concatenated, run_ranks
    sizes = [9 << 20, 0, 11 << 20, 70001, (7 << 20) + 5, 10 << 20]
    out = run_ranks(3, _concatenated, sizes, "cuda")
    hashes = np.frombuffer(out[0]["hashes"], dtype=np.uint8).reshape(-1, 32)
    assert all(o["hashes"] == out[0]["hashes"] for o in out) and out[0]["shared"]
    for f, n in enumerate(sizes):
        assert hashes[f].tobytes() == xet_oracle.hash_file(xet_inputs.splitmix_bytes(300 + f, n)).file_hash, f
