#!/usr/bin/env python
"""Many tiny files in one batch (default 200,000 files of 0..20,000 bytes): every file hash against the oracle."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import numpy as np, torch
import ome_b200, xet_inputs, xet_oracle

nf = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
rng = np.random.default_rng(99)
sizes = rng.integers(0, 20_001, nf)
sizes[::1000] = 0
off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
data = xet_inputs.splitmix_bytes(123, int(off[-1]))
h = ome_b200.XetHasher(device=0, stage_timing=True)
d = torch.from_numpy(data).cuda()
for where, src in (("device", d), ("host", data)):
    h.hash_buffers(src, off)
    t0 = time.perf_counter(); cl = h.hash_buffers(src, off, want_chunks=True); dt = time.perf_counter() - t0
    bad = 0
    for f in range(nf):
        a, b = int(off[f]), int(off[f + 1])
        ref = xet_oracle.hash_file(data[a:b])
        fr = cl.files[f]
        if fr.file_hash != ref.file_hash or fr.n_chunks != ref.chunk_ends.size:
            bad += 1
    print({k: round(v, 2) if isinstance(v, float) else v for k, v in h.last_timings().items()})
    print(f"{where}: {nf} files, {off[-1]/1e9:.2f} GB, {dt*1e3:.1f} ms ({off[-1]/dt/1e9:.1f} GB/s, {nf/dt/1e6:.2f} M files/s), mismatches: {bad}")
    assert bad == 0
