#!/usr/bin/env python
"""Probe: do two hash pipelines on two streams overlap usefully on one GPU? (two contexts, two host threads)"""
import os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import ome_b200

n = 4 << 30
bufs, hs = [], []
for i in range(2):
    s = torch.cuda.Stream()
    h = ome_b200.XetHasher(device=0, stream=s.cuda_stream)
    b = torch.empty(n + 256, dtype=torch.uint8, device="cuda")
    h.fill_splitmix(b.data_ptr(), 20 + i, 0, n)
    bufs.append(b); hs.append(h)
off = np.asarray([0, n], dtype=np.uint64)
def run(i, iters):
    for _ in range(iters):
        hs[i].hash_buffers(bufs[i], off, memory=2)
for i in range(2): run(i, 2)
torch.cuda.synchronize()
t0 = time.perf_counter(); run(0, 6); run(1, 6); torch.cuda.synchronize(); t_serial = time.perf_counter() - t0
t0 = time.perf_counter()
th = [threading.Thread(target=run, args=(i, 6)) for i in range(2)]
[t.start() for t in th]; [t.join() for t in th]; torch.cuda.synchronize(); t_par = time.perf_counter() - t0
print(f"serial {12*n/t_serial/1e9:.1f} GB/s   two-stream {12*n/t_par/1e9:.1f} GB/s")
