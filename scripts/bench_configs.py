#!/usr/bin/env python
"""Device-resident throughput of the BASELINE.json configs that fit one GPU (not the contract bench).

usage: python scripts/bench_configs.py [config ...]   configs: c1 c2 c3slice c5
  c1       1 GiB single shard (seed 1)
  c2       Llama-3-8B shape, 4 shards, 16.06 GB (what bench.py times)
  c3slice  Llama-3-70B shape: the 8 of 30 shards (4.7 GB each) one GPU owns at N=4 (file partition)
  c5       many-small-file repo: 10,000 files, sizes log-uniform 4 KiB..64 MiB (splitmix seed 5), one empty
           file and sub-8 KiB files included; packed in one device buffer
"""
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch

import ome_b200
import xet_inputs


def sizes_for(cfg):
    if cfg == "c1":
        return [1 << 30], 1
    if cfg == "c2":
        return [4_976_698_672, 4_999_802_720, 4_915_916_176, 1_168_138_808], 10
    if cfg == "c3slice":
        return [4_700_000_000 + 1000 * i for i in range(8)], 100
    if cfg == "c5":
        u = xet_inputs.splitmix_bytes(5, 8 * 10000).view("<u8").astype(np.float64) / 2.0**64
        s = np.exp(np.log(4096) + u * (np.log(64 << 20) - np.log(4096))).astype(np.int64)
        s[17] = 0
        s[18] = 100
        return [int(x) for x in s], 50000
    raise SystemExit(f"unknown config {cfg}")


def main():
    cfgs = sys.argv[1:] or ["c1", "c2", "c5"]
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    h = ome_b200.XetHasher(device=0, stream=stream.cuda_stream, stage_timing=True)
    for cfg in cfgs:
        sizes, seed0 = sizes_for(cfg)
        off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
        total = int(off[-1])
        buf = torch.empty(total + 256, dtype=torch.uint8, device="cuda")
        if len(sizes) <= 64:
            for i, n in enumerate(sizes):
                h.fill_splitmix(buf.data_ptr() + int(off[i]), seed0 + i, 0, n)
        else:  # contents of small files: one stream, cut at the file offsets (cheap to generate, still random)
            h.fill_splitmix(buf.data_ptr(), seed0, 0, total)
        for _ in range(3):
            cl = h.hash_buffers(buf, off, memory=ome_b200.hasher.MEM_DEVICE)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        steps = 10
        torch.cuda.synchronize()
        e0.record(stream)
        for _ in range(steps):
            cl = h.hash_buffers(buf, off, memory=ome_b200.hasher.MEM_DEVICE)
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        t = h.last_timings()
        print(json.dumps({"config": cfg, "files": len(sizes), "bytes": total, "chunks": int(sum(f.n_chunks for f in cl.files)),
                          "ms": round(ms, 3), "GBps": round(total / ms / 1e6, 1),
                          "stage_ms": {k: round(t[k], 3) for k in ("gear_ms", "resolve_ms", "blake3_ms", "merkle_ms", "d2h_ms")},
                          "launches": t["kernel_launches"]}))
        del buf
        torch.cuda.empty_cache()
    h.close()


if __name__ == "__main__":
    main()
