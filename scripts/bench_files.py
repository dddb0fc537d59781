#!/usr/bin/env python
"""Throughput of xetgpu_hash_files (paths in, hashes out) vs hf_xet.hash_files on the same tmpfs files."""
import os, sys, time, tempfile, shutil
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import ome_b200

sizes = [int(float(x) * 1e9) for x in (sys.argv[1:] or ["2", "2", "2", "1"])]
d = tempfile.mkdtemp(prefix="xetfiles-", dir="/dev/shm")
h = ome_b200.XetHasher(device=0)
paths = []
for i, n in enumerate(sizes):
    t = torch.empty(n, dtype=torch.uint8, device="cuda")
    h.fill_splitmix(t.data_ptr(), 40 + i, 0, n)
    p = os.path.join(d, f"f{i}.bin"); t.cpu().numpy().tofile(p); paths.append(p); del t
total = sum(sizes)
try:
    h.hash_files(paths[:1])
    for rep in range(3):
        t0 = time.perf_counter(); cl = h.hash_files(paths); dt = time.perf_counter() - t0
        print(f"xetgpu_hash_files: {total/dt/1e9:.2f} GB/s ({dt*1e3:.0f} ms)")
    import hf_xet
    t0 = time.perf_counter(); ref = hf_xet.hash_files(paths); dt = time.perf_counter() - t0
    print(f"hf_xet.hash_files: {total/dt/1e9:.2f} GB/s ({dt*1e3:.0f} ms), equal: {[f.hex for f in cl.files] == [r.hash for r in ref]}")
finally:
    shutil.rmtree(d, ignore_errors=True)
