#!/usr/bin/env python
"""BASELINE config 4: DeepSeek-V3 fp8 checkpoint shape (163 shards of ~4.23 GB, ~690 GB), the shards laid end to
end and cut into N equal byte ranges, one per GPU.  Shards inside a range are hashed whole; a shard that straddles
a range boundary is shared by the two neighbouring ranks with the byte-range protocol of ome_b200.sharded
(halo exchange and 8-byte carry over NCCL P2P, rank-local fix-up window, merkle combine on the lower rank).

usage: torchrun --nproc-per-node N scripts/bench_config4.py [n_files=163] [GB_per_file=4.23] [steps=2]
(smaller arguments give a scaled trial; the numbers in profiles/ state which shape was run)
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist

import ome_b200
from ome_b200 import sharded

n_files = int(sys.argv[1]) if len(sys.argv) > 1 else 163
gb = float(sys.argv[2]) if len(sys.argv) > 2 else 4.23
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
SEED0 = 1000
sizes = [int(gb * 1e9) + 4096 * (i % 7) + (i % 3) for i in range(n_files)]   # ~4.23 GB each, not all equal
plan = sharded.plan_concatenated(sizes, world)
mine = plan[rank]
owners = {}
for r, pieces in enumerate(plan):
    for p in pieces:
        owners.setdefault(p.file, []).append(r)
shared = sorted(f for f, rs in owners.items() if len(rs) > 1)
# one process group per shared file (every rank must create every group, in the same order)
groups = {f: dist.new_group(ranks=owners[f]) for f in shared}

h = ome_b200.XetHasher(device=lr)
CAP = 1 << 18   # headroom in front of a partial piece (>= RANGE_LEAD, keeps 16-byte alignment)
layout, off = [], 0
for p in mine:
    head = CAP if p.begin > 0 else 0
    layout.append((off, head))
    off += head + (p.end - p.begin)
    off = (off + 255) & ~255
buf = torch.empty(off + 256, dtype=torch.uint8, device=dev)
for p, (o, head) in zip(mine, layout):
    if p.end > p.begin:
        h.fill_splitmix(buf.data_ptr() + o + head, SEED0 + p.file, p.begin, p.end - p.begin)
torch.cuda.synchronize()
whole = [(p, o) for p, (o, head) in zip(mine, layout) if p.whole]
partial = [(p, o, head) for p, (o, head) in zip(mine, layout) if not p.whole]
# whole files are hashed in one batched call: offsets into buf (files are not contiguous: gaps of alignment padding
# are skipped by giving each file its own [begin, end) pair through a packed view per file run)
file_hashes = torch.zeros(n_files, 32, dtype=torch.int32, device=dev)
times, rehash = [], 0


def run_once():
    global rehash
    file_hashes.zero_()
    # 1. whole shards of this rank: consecutive in buf up to 256-byte padding -> one call per shard run
    if whole:
        # pack offsets: every whole file f occupies [o, o + len); the padding between files is hashed as part of
        # nothing because offsets are given per file through a (start, end) list with zero-length spacer files
        offs, idx = [], []
        for p, o in whole:
            offs += [o, o + p.file_len]
            idx.append(p.file)
        offsets = np.asarray(sorted(set(offs)), dtype=np.uint64)
        cl = h.hash_buffers(buf, offsets, memory=ome_b200.hasher.MEM_DEVICE)
        start_to_slot = {int(offsets[i]): i for i in range(len(offsets) - 1)}
        for p, o in whole:
            fr = cl.files[start_to_slot[o]]
            assert fr.size == p.file_len
            file_hashes[p.file] = torch.frombuffer(bytearray(fr.file_hash), dtype=torch.uint8).to(dev).to(torch.int32)
    # 2. shared shards: even-numbered boundaries first, then odd ones (no chain of waits across ranks)
    for parity in (0, 1):
        for k, f in enumerate(shared):
            if k % 2 != parity or rank not in owners[f]:
                continue
            p, o, head = next(x for x in partial if x[0].file == f)
            g = groups[f]
            ranges = [(q.begin, q.end) for r in owners[f] for q in plan[r] if q.file == f]
            view = buf[o: o + head + (p.end - p.begin)]
            lead = sharded.exchange_lead(view, head, ranges, group=g) if len(ranges) > 1 else 0
            out = sharded.hash_file_by_ranges(h, view[head - lead:], lead, p.file_len, p.begin, p.end, group=g,
                                              comm_device=dev)
            rehash = max(rehash, out.rehashed_bytes)
            if out.file_hash is not None:
                file_hashes[f] = torch.frombuffer(bytearray(out.file_hash), dtype=torch.uint8).to(dev).to(torch.int32)
    # 3. every rank ends with every file hash (disjoint rows: a sum all-reduce is a gather)
    dist.all_reduce(file_hashes)


for it in range(steps + 1):
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    run_once()
    torch.cuda.synchronize()
    dist.barrier()
    if it:
        times.append(time.perf_counter() - t0)

# check the shared shards (the interesting ones) against a single-GPU whole-file hash on their lower owner
ok = True
for f in shared:
    if owners[f][0] == rank:
        t = torch.empty(sizes[f] + 256, dtype=torch.uint8, device=dev)
        h.fill_splitmix(t.data_ptr(), SEED0 + f, 0, sizes[f])
        ref = h.hash_buffers(t, [0, sizes[f]]).files[0].file_hash
        ok &= bytes(file_hashes[f].to(torch.uint8).cpu().numpy()) == ref
        del t
okt = torch.tensor([1 if ok else 0], device=dev)
dist.all_reduce(okt, op=dist.ReduceOp.MIN)
rh = torch.tensor([rehash], dtype=torch.int64, device=dev)
dist.all_reduce(rh, op=dist.ReduceOp.MAX)
rehash = int(rh.item())
if rank == 0:
    total = sum(sizes)
    t = min(times)
    digest = bytes(file_hashes.to(torch.uint8).cpu().numpy().reshape(-1))
    import hashlib
    print(json.dumps({"config": "c4: shards laid end to end, byte-range partitioned", "n_gpus": world, "files": n_files,
                      "bytes": total, "GBps": round(total / t / 1e9, 1), "ms": round(t * 1e3, 2),
                      "shared_files": shared, "max_rehashed_bytes": rehash,
                      "shared_files_equal_single_gpu_hash": bool(okt.item()),
                      "sha256_of_all_file_hashes": hashlib.sha256(digest).hexdigest(),
                      "timing": f"host wall clock, barrier to barrier, min of {steps}"}))
h.close()
dist.destroy_process_group()
