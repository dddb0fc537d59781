#!/usr/bin/env python
"""Small fixed workload for ncu: one device-resident splitmix64 shard hashed a few times.

usage: python scripts/profile_workload.py [GiB=2] [iters=3]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import ome_b200

gib = float(sys.argv[1]) if len(sys.argv) > 1 else 2.0
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 3
n = int(gib * (1 << 30))
h = ome_b200.XetHasher(device=0, stage_timing=True)
buf = torch.empty(n + 256, dtype=torch.uint8, device="cuda")
h.fill_splitmix(buf.data_ptr(), 1, 0, n)
off = np.asarray([0, n], dtype=np.uint64)
for i in range(iters):
    cl = h.hash_buffers(buf, off, memory=ome_b200.hasher.MEM_DEVICE)
    t = h.last_timings()
    print(i, cl.files[0].hex, {k: round(v, 3) if isinstance(v, float) else v for k, v in t.items()})
h.close()
