#!/usr/bin/env python
"""One very large device-resident file (default 100 GB): internal consistency at offsets far beyond 2^32.

Checks: chunk sizes within [8192, 131072]; merkle of the returned chunk list reproduces the file hash; the
file hashed as two byte ranges with a carried cut reassembles to the same chunk list; the first 64 MiB
agree with the hf_xet golden prefix (splitmix seed 1).
"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import ome_b200

n = int(float(sys.argv[1]) * 1e9) if len(sys.argv) > 1 else 100_000_000_003
h = ome_b200.XetHasher(device=0, stage_timing=True)
buf = torch.empty(n + 256, dtype=torch.uint8, device="cuda")
h.fill_splitmix(buf.data_ptr(), 1, 0, n)
t0 = time.perf_counter()
cl = h.hash_buffers(buf[:n], [0, n], want_chunks=True)
dt = time.perf_counter() - t0
ends, hashes = cl.chunks_of(0)
sizes = np.diff(np.concatenate([[0], ends])).astype(np.uint64)
assert int(ends[-1]) == n and sizes.max() <= 131072 and sizes[:-1].min() >= 8192
root, fh = h.merkle_from_chunks(hashes, sizes)
assert root == cl.files[0].merkle_root and fh == cl.files[0].file_hash
mid = (n // 2) + 777
lead = 131072 + 64
r0 = h.hash_range(buf[:mid], n, 0, mid, 0)
r1 = h.hash_range(buf[mid - lead:n], n, mid, n, r0.exit_cut)
assert np.array_equal(np.concatenate([r0.chunk_ends, r1.chunk_ends]), ends)
assert np.array_equal(np.concatenate([r0.chunk_hashes, r1.chunk_hashes]), hashes)
gold = json.load(open(os.path.join(ROOT, "tests", "golden", "xet_golden.json")))
case = next(c for c in gold["cases"] if c["name"] == "splitmix_seed1_64MiB")
assert [int(x) for x in ends[:16]] == case["cuts_head"]
print(json.dumps({"bytes": n, "chunks": int(ends.size), "file_hash": cl.files[0].hex, "GBps_incl_chunk_readback": round(n / dt / 1e9, 1),
                  "timings": {k: (round(v, 2) if isinstance(v, float) else v) for k, v in h.last_timings().items()}, "checks": "ok"}))
