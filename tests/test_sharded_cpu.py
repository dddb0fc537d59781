"""World-size-2 (and 3) gloo tests of the multi-GPU host logic, on CPU with the oracle engine."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import xet_inputs
import xet_oracle
from ome_b200 import sharded


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, fn, args, ret):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (here, os.path.join(os.path.dirname(here), "oracle"), os.path.dirname(here)):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ret[rank] = fn(rank, world, *args)
    finally:
        dist.destroy_process_group()


def run_ranks(world, fn, *args):
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, free_port(), fn, args, ret), nprocs=world, join=True)
    return [ret[r] for r in range(world)]


def test_lpt_partition_balances():
    sizes = [4976, 4999, 4915, 1168, 300, 7000, 10, 0]
    p = sharded.FilePartition.lpt(sizes, 3)
    assert sorted(i for a in p.assignment for i in a) == list(range(len(sizes)))
    assert max(p.loads) - min(p.loads) <= max(sizes)
    assert p.loads == [sum(sizes[i] for i in a) for a in p.assignment]


def test_split_byte_ranges_properties():
    for n, w in [(0, 2), (5, 2), (300000, 2), (10 << 20, 3), ((64 << 20) + 7, 8), (131136, 2)]:
        r = sharded.split_byte_ranges(n, w)
        assert len(r) == w and r[0][0] == 0 and r[-1][1] == n
        assert all(r[i][1] == r[i + 1][0] for i in range(w - 1))
        for i in range(w - 1):
            if r[i + 1][1] > r[i + 1][0]:          # non-empty successor => this range can feed its halo
                assert r[i][1] - r[i][0] >= sharded.RANGE_LEAD


def _file_partition(rank, world, sizes):
    from oracle_engine import OracleEngine
    part = sharded.FilePartition.lpt(sizes, world)
    local = [(i, xet_inputs.splitmix_bytes(700 + i, sizes[i])) for i in part.assignment[rank]]
    h, c = sharded.hash_files_partitioned(OracleEngine(), local, len(sizes))
    return h.tobytes(), c.tolist()


def test_file_partition_gloo_world2():
    sizes = [300000, 0, 70000, 1 << 20, 5, 200000, 9000]
    out = run_ranks(2, _file_partition, sizes)
    assert out[0] == out[1]                                 # every rank ends with every result
    hashes = np.frombuffer(out[0][0], dtype=np.uint8).reshape(-1, 32)
    for i, n in enumerate(sizes):
        ref = xet_oracle.hash_file(xet_inputs.splitmix_bytes(700 + i, n))
        assert hashes[i].tobytes() == ref.file_hash
        assert out[0][1][i] == ref.chunk_ends.size


def _byte_ranges(rank, world, spec, use_exchange):
    from oracle_engine import OracleEngine
    data = xet_inputs.build(spec)
    n = data.size
    ranges = sharded.split_byte_ranges(n, world)
    begin, end = ranges[rank]
    cap = sharded.RANGE_LEAD
    piece = torch.zeros(cap + (end - begin), dtype=torch.uint8)
    piece[cap:] = torch.from_numpy(data[begin:end].copy())  # a rank loads only its own range ...
    if use_exchange:
        lead = sharded.exchange_lead(piece, cap, ranges)      # ... and receives the halo from its neighbour
    else:
        lead = min(begin, cap) if end > begin else 0
        piece[cap - lead:cap] = torch.from_numpy(data[begin - lead:begin].copy())
    assert bytes(piece[cap - lead:cap].numpy()) == bytes(data[begin - lead:begin])
    eng = OracleEngine()
    out = sharded.hash_file_by_ranges(eng, piece[cap - lead:], lead, n, begin, end)
    return dict(root=out.merkle_root, fh=out.file_hash, n=out.n_chunks_total, ends=out.chunk_ends.tolist(),
                rehashed=out.rehashed_bytes, calls=eng.range_calls, range=(begin, end))


@pytest.mark.parametrize("world,spec", [
    (2, [["splitmix", 81, (6 << 20) + 12345]]),
    (3, [["splitmix", 82, 9 << 20]]),
    # chains that cannot re-merge: no candidates after an unaligned prefix -> whole-range fallback
    (2, [["splitmix", 83, (1 << 20) + 7], ["zeros", 7 << 20]]),
    # tiny file: every rank but the first owns an empty range
    (2, [["splitmix", 84, 50000]]),
    (3, [["zeros", 0]]),
], ids=["w2-random", "w3-random", "w2-zeros-nomerge", "w2-tiny", "w3-empty"])
def test_byte_range_partition_gloo(world, spec):
    data = xet_inputs.build(spec)
    ref = xet_oracle.hash_file(data)
    out = run_ranks(world, _byte_ranges, spec, True)
    assert out[0]["fh"] == ref.file_hash
    assert out[0]["n"] == ref.chunk_ends.size
    if ref.chunk_ends.size:
        assert out[0]["root"] == ref.merkle_root
    assert all(o["fh"] is None for o in out[1:])            # the combine runs on rank 0 only
    ends = [e for o in out for e in o["ends"]]
    assert ends == [int(x) for x in ref.chunk_ends]
    for o in out:                                           # each rank owns exactly the chunks ending in its range
        assert all(o["range"][0] < e <= o["range"][1] for e in o["ends"])


def test_fixup_window_is_small_on_random_data():
    """After the true carry arrives a rank re-hashes one small window, not its range."""
    spec = [["splitmix", 85, 24 << 20]]
    out = run_ranks(2, _byte_ranges, spec, False)
    begin, end = out[1]["range"]
    assert 0 < out[1]["rehashed"] <= sharded.FIX_WINDOW + sharded.RANGE_LEAD < end - begin
    assert len(out[1]["calls"]) == 2                        # speculative pass + one fix-up window


def test_plan_concatenated_tiles_every_file():
    rng = np.random.default_rng(4)
    for world in (1, 2, 3, 8):
        for sizes in ([4_230_000_000] * 163, [5, 0, 300_000, 0, 7_000_000, 1_000_000, 0],
                      [int(x) for x in rng.integers(0, 3_000_000, 40)], [0, 0], [10_000_000]):
            plan = sharded.plan_concatenated(sizes, world)
            assert len(plan) == world
            owned = {f: [] for f in range(len(sizes))}
            for r, pieces in enumerate(plan):
                assert [p.file for p in pieces] == sorted(p.file for p in pieces)
                for p in pieces:
                    owned[p.file].append((r, p.begin, p.end))
                    assert p.file_len == sizes[p.file]
            for f, n in enumerate(sizes):
                ps = owned[f]
                assert ps, (world, f)                                  # every file, empty ones too, has an owner
                assert ps[0][1] == 0 and ps[-1][2] == n
                assert all(ps[i][2] == ps[i + 1][1] and ps[i + 1][0] == ps[i][0] + 1 for i in range(len(ps) - 1))
                if len(ps) > 1:                                        # shared files: every piece can feed a halo
                    assert all(e - b >= sharded.RANGE_LEAD for _, b, e in ps)
            if sum(sizes) > 10 * world * (sharded.RANGE_LEAD + 4096):  # balanced to within one file-end snap
                loads = [sum(p.end - p.begin for p in pieces) for pieces in plan]
                assert max(loads) - min(loads) <= 2 * (sharded.RANGE_LEAD + 4096) + max(sizes) * 0 + 4096 * world \
                    or max(sizes) < sum(sizes) // world


def _concatenated(rank, world, sizes, engine_kind="oracle"):
    """BASELINE config-4 shape in miniature: files end to end, cut into `world` byte ranges."""
    if engine_kind == "oracle":
        from oracle_engine import OracleEngine
        eng, dev = OracleEngine(), torch.device("cpu")
    else:  # the CUDA engine, every rank on GPU 0 (single-GPU box); the exchanges stay on gloo / host tensors
        import ome_b200
        eng, dev = ome_b200.XetHasher(device=0), torch.device("cuda", 0)
    plan = sharded.plan_concatenated(sizes, world)
    owners, shared = sharded.shared_files(plan)
    groups = {f: dist.new_group(ranks=owners[f]) for f in shared}
    lay = sharded.layout_pieces(plan[rank])
    buf = torch.zeros(lay.nbytes, dtype=torch.uint8)
    for k, p in enumerate(lay.pieces):
        if p.end > p.begin:
            o = lay.data_offset(k)
            buf[o:o + p.end - p.begin] = torch.from_numpy(xet_inputs.splitmix_bytes(300 + p.file, p.file_len)[p.begin:p.end].copy())
    buf = buf.to(dev)
    phases = {}
    hashes, rehashed = sharded.hash_concatenated(eng, buf, lay, plan, len(sizes), groups, comm_device=torch.device("cpu"),
                                                 phases=phases)
    if engine_kind != "oracle":
        eng.close()
    return dict(hashes=hashes.tobytes(), rehashed=rehashed, shared=shared, phases=sorted(phases))


@pytest.mark.parametrize("world,sizes", [
    (2, [3 << 20, 0, 5 << 20, 70001, 2 << 20]),
    (3, [6 << 20, 6 << 20, 6 << 20, 6 << 20]),
    (3, [20 << 20]),                       # one file shared by all three ranks
], ids=["w2-mixed", "w3-equal", "w3-onefile"])
def test_concatenated_plan_gloo(world, sizes):
    out = run_ranks(world, _concatenated, sizes)
    assert all(o["hashes"] == out[0]["hashes"] for o in out)      # every rank ends with every file hash
    hashes = np.frombuffer(out[0]["hashes"], dtype=np.uint8).reshape(-1, 32)
    for f, n in enumerate(sizes):
        assert hashes[f].tobytes() == xet_oracle.hash_file(xet_inputs.splitmix_bytes(300 + f, n)).file_hash, f
    assert out[0]["shared"], "the shapes above are chosen so that at least one file straddles a range boundary"
