#!/usr/bin/env python
"""BASELINE config 3: Llama-3-70B bf16 checkpoint shape (30 shards of ~4.7 GB, ~141 GB), file-partitioned over the
ranks (longest-processing-time-first), device-resident; every rank ends with all 30 file hashes.

usage: torchrun --nproc-per-node N scripts/bench_config3.py [steps=3]
"""
import hashlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch, torch.distributed as dist
import ome_b200
from ome_b200 import sharded

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
sizes = [4_700_000_000 + 1_000_003 * (i % 5) for i in range(29)] + [2_100_000_000]   # last shard is short, as in real checkpoints
part = sharded.FilePartition.lpt(sizes, world)
mine = part.assignment[rank]
h = ome_b200.XetHasher(device=lr)
offs = np.concatenate([[0], np.cumsum([sizes[i] for i in mine])]).astype(np.uint64)
buf = torch.empty(int(offs[-1]) + 256, dtype=torch.uint8, device=dev)
for k, i in enumerate(mine):
    h.fill_splitmix(buf.data_ptr() + int(offs[k]), 100 + i, 0, sizes[i])
hashes = torch.zeros(len(sizes), 32, dtype=torch.int32, device=dev)
times = []
for it in range(steps + 1):
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    hashes.zero_()
    if mine:
        cl = h.hash_buffers(buf, offs, memory=ome_b200.hasher.MEM_DEVICE)   # this rank's shards in one batched call
        for k, i in enumerate(mine):
            hashes[i] = torch.frombuffer(bytearray(cl.files[k].file_hash), dtype=torch.uint8).to(dev).to(torch.int32)
    dist.all_reduce(hashes)                                                 # disjoint rows: sum == gather
    torch.cuda.synchronize(); dist.barrier()
    if it:
        times.append(time.perf_counter() - t0)
if rank == 0:
    t, total = min(times), sum(sizes)
    print(json.dumps({"config": "c3: 30 shards file-partitioned (LPT)", "n_gpus": world, "bytes": total,
                      "GBps": round(total / t / 1e9, 1), "ms": round(t * 1e3, 2), "max_rank_bytes": max(part.loads),
                      "balance": round(total / world / max(part.loads), 3),
                      "sha256_of_all_file_hashes": hashlib.sha256(bytes(hashes.to(torch.uint8).cpu().numpy().reshape(-1))).hexdigest(),
                      "timing": f"host wall clock, barrier to barrier, min of {steps}"}))
h.close()
dist.destroy_process_group()
