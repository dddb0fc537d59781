#!/usr/bin/env python
"""xetgpu_multi_hash_files (ONE process, every visible GPU) on shards in tmpfs, against the single-device call and hf_xet.

usage: python scripts/bench_multi_files.py [n_files=8] [GB_per_file=2.0]
"""
import json, os, shutil, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import ome_b200

n_files = int(sys.argv[1]) if len(sys.argv) > 1 else 8
gb = float(sys.argv[2]) if len(sys.argv) > 2 else 2.0
sizes = [int(gb * 1e9) + 4096 * i + i for i in range(n_files)]
d = tempfile.mkdtemp(prefix="xetmulti-", dir="/dev/shm")
paths = []
try:
    with ome_b200.XetHasher(device=0) as h:
        for i, n in enumerate(sizes):
            t = torch.empty(n, dtype=torch.uint8, device="cuda:0")
            h.fill_splitmix(t.data_ptr(), 60 + i, 0, n)
            p = os.path.join(d, f"shard-{i}.bin")
            t.cpu().numpy().tofile(p)
            paths.append(p)
            del t
        h.hash_files(paths[:1])
        t0 = time.perf_counter(); one = h.hash_files(paths); dt1 = time.perf_counter() - t0
    total = sum(sizes)
    with ome_b200.XetMultiHasher() as m:
        m.hash_files(paths[:2])
        best = 1e9
        for _ in range(3):
            t0 = time.perf_counter(); cl = m.hash_files(paths); best = min(best, time.perf_counter() - t0)
        plan = m.plan(paths)
        ndev = m.n_devices
    import hf_xet
    t0 = time.perf_counter(); ref = hf_xet.hash_files(paths); dtr = time.perf_counter() - t0
    print(json.dumps({"files": n_files, "bytes": total, "devices": ndev, "host_cores": os.cpu_count(),
                      "multi_GBps": round(total / best / 1e9, 2), "single_device_GBps": round(total / dt1 / 1e9, 2),
                      "hf_xet_GBps": round(total / dtr / 1e9, 2),
                      "equal_single": [f.hex for f in cl.files] == [f.hex for f in one.files],
                      "equal_hf_xet": [f.hex for f in cl.files] == [r.hash for r in ref],
                      "shared_files": sorted({f for _, f, b, e in plan if b > 0})}))
finally:
    shutil.rmtree(d, ignore_errors=True)
