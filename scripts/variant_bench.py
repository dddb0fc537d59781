#!/usr/bin/env python
"""Device-resident config-2 step time for one build of the library (XETGPU_LIBRARY) in the one-lane and forced
two-lane policies.  usage: XETGPU_LIBRARY=... python scripts/variant_bench.py NAME [gb_scale]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
name = sys.argv[1]
lanes = os.environ.get("XETGPU_TWO_LANES", "-")
import ome_b200
sizes = [4_976_698_672, 4_999_802_720, 4_915_916_176, 1_168_138_808]
off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
total = int(off[-1])
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
h = ome_b200.XetHasher(device=0, stream=stream.cuda_stream, stage_timing=True)
buf = torch.empty(total + 256, dtype=torch.uint8, device="cuda")
for i, n in enumerate(sizes):
    h.fill_splitmix(buf.data_ptr() + int(off[i]), 10 + i, 0, n)
for _ in range(3):
    cl = h.hash_buffers(buf, off, memory=2)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
steps = 10
torch.cuda.synchronize(); e0.record(stream)
for _ in range(steps):
    cl = h.hash_buffers(buf, off, memory=2)
e1.record(stream); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
t = h.last_timings()
print(json.dumps({"variant": name, "two_lanes": lanes, "ms": round(ms, 3), "GBps": round(total / ms / 1e6, 1),
                  "gear_kernel_ms": round(t["gear_kernel_ms"], 3), "leaf_kernel_ms": round(t["leaf_kernel_ms"], 3),
                  "launches": t["kernel_launches"], "hash0": cl.files[0].hex[:16], "hash3": cl.files[3].hex[:16]}))
h.close()
