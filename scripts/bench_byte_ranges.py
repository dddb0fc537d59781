#!/usr/bin/env python
"""BASELINE config 4 shape: ONE large file byte-range partitioned across the ranks (torchrun, NCCL).

Every rank fills its own range of a single splitmix64 stream on its GPU (device-resident), then the ranks
run ome_b200.sharded: halo exchange over NCCL P2P, speculative pass of every range in parallel, the 8-byte
carry chain with small fix-up windows, chunk lists to rank 0, merkle combine on rank 0.

usage: torchrun --nproc-per-node N scripts/bench_byte_ranges.py [GB_per_rank=16] [steps=3]
"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch, torch.distributed as dist
import ome_b200
from ome_b200 import sharded

gb = float(sys.argv[1]) if len(sys.argv) > 1 else 16.0
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
h = ome_b200.XetHasher(device=lr)
n = int(gb * 1e9) * world + 12345
ranges = sharded.split_byte_ranges(n, world)
begin, end = ranges[rank]
cap = 1 << 18
piece = torch.empty(cap + (end - begin) + 256, dtype=torch.uint8, device=dev)
h.fill_splitmix(piece.data_ptr() + cap, 4242, begin, end - begin)
times = []
for it in range(steps + 1):
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    lead = sharded.exchange_lead(piece[: cap + (end - begin)], cap, ranges)
    phases = {}
    t1 = time.perf_counter()
    out = sharded.hash_file_by_ranges(h, piece[cap - lead: cap + (end - begin)], lead, n, begin, end, comm_device=dev, phases=phases)
    phases['halo_exchange_ms'] = (t1 - t0) * 1e3
    torch.cuda.synchronize(); dist.barrier()
    if it:
        times.append(time.perf_counter() - t0)
rehashed = torch.tensor([out.rehashed_bytes], dtype=torch.int64, device=dev)
dist.all_reduce(rehashed, op=dist.ReduceOp.MAX)
if rank == 0:
    t = min(times)
    print(json.dumps({"config": "c4-shape: one file, byte-range partitioned", "n_gpus": world, "file_bytes": n,
                      "GBps": round(n / t / 1e9, 1), "ms": round(t * 1e3, 2), "chunks": out.n_chunks_total,
                      "file_hash": ome_b200.hash_to_hex(out.file_hash), "max_rehashed_bytes_per_rank": int(rehashed.item()),
                      "rank0_phases_ms": {k: round(v, 2) for k, v in phases.items()},
                      "timing": "host wall clock around the whole collective call, min of %d" % steps}))
if rank == 1:
    print('rank1 phases', {k: round(v, 2) for k, v in phases.items()}, flush=True)
h.close()
dist.destroy_process_group()
