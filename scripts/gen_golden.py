#!/usr/bin/env python
"""Generate tests/golden/xet_golden.json from the reference arithmetic.

Ground truth = hf_xet.hash_files (hf_xet 1.5.0: the official compiled build of
huggingface/xet-core, the dependency /root/reference/pkg/xet/Cargo.lock:237 pins at an
older commit).  Run in the build container only (needs hf_xet and /dev/shm); the GPU
box never runs this.  For every input the file hash comes from hf_xet; the per-stage
intermediates (cuts, chunk-hash digest, merkle root) come from the oracle and are
recorded ONLY when the oracle's file hash equals hf_xet's (the file hash commits to
every (chunk_hash, chunk_len) pair in order).

usage: python scripts/gen_golden.py [--large]
"""
import hashlib
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import hf_xet  # noqa: E402
import xet_inputs  # noqa: E402
import xet_oracle  # noqa: E402
from xet_inputs import W_HEX  # noqa: E402

MiB = 1 << 20
GEAR = None


def gear_table():
    import re
    import struct
    txt = open(os.path.join(ROOT, "include", "xet_gear_table.h")).read()
    vals = [int(x[:-3], 16) for x in re.findall(r"0x[0-9a-f]{16}ULL", txt)]
    assert len(vals) == 256
    assert hashlib.sha256(b"".join(struct.pack("<Q", v) for v in vals)).hexdigest() == \
        "e1d3936666d7ae7a977c958e9afcc75f90aaca758ce5fbe4ece61dffefe1912c"
    return vals


def short_match_sequences(k, limit=4):
    """k-byte sequences whose *partially warm* rolling hash (from h=0) has top 16 bits 0."""
    T = gear_table()
    M = (1 << 64) - 1
    out = []
    import itertools
    for seq in itertools.product(range(256), repeat=k):
        h = 0
        for b in seq:
            h = ((h << 1) + T[b]) & M
        if h >> 48 == 0:
            out.append(bytes(seq).hex())
            if len(out) >= limit:
                break
    return out


def cases(large):
    c = []
    add = lambda name, spec, **kw: c.append(dict(name=name, spec=spec, **kw))
    add("empty", [])
    add("one_zero_byte", [["zeros", 1]])
    add("hello_world", [["ascii", "hello world"]])
    for n in (63, 64, 65, 1023, 1024, 1025, 2048, 2049, 3072, 4097, 8127, 8128, 8191, 8192, 8193,
              65536, 131071, 131072, 131073, 262144, 393216):
        add(f"zeros_{n}", [["zeros", n]])
    for k in (8, 9, 10, 11, 19):
        add(f"zeros_{k}x131072", [["zeros", k * 131072]])
    add("mod251_1MiB", [["mod251", MiB]])
    for n in (1, 64, 1000, 1024, 1025, 5000, 8192, 70000):
        add(f"splitmix_seed4_{n}", [["splitmix", 4, n]])
    add("splitmix_seed1_1MiB", [["splitmix", 1, MiB]])
    add("splitmix_seed1_1MiB_plus1", [["splitmix", 1, MiB + 1]])
    add("splitmix_seed2_10MiB", [["splitmix", 2, 10 * MiB]])
    add("splitmix_seed1_64MiB", [["splitmix", 1, 64 * MiB]])
    add("dense_W_x16384", [["repeat", W_HEX, 16384]])
    add("dense_W_x1048576", [["repeat", W_HEX, 1048576]])
    add("deadzone_after_forced", [["zeros", 131072], ["hex", W_HEX], ["splitmix", 9, 500000]])
    # window match landing exactly at chunk size 8191 / 8192 / 8193 (first chunk)
    for size in (8190, 8191, 8192, 8193, 8194):
        add(f"W_ends_at_{size}", [["splitmix", 21, size - 64], ["hex", W_HEX], ["splitmix", 22, 30000]])
    # the same, but for a chunk that starts after a forced cut at 131072
    for size in (8191, 8192, 8193):
        add(f"W_ends_at_forced_plus_{size}",
            [["zeros", 131072], ["zeros", size - 64], ["hex", W_HEX], ["splitmix", 23, 20000]])
    # partially-warm matches: xet-core starts rolling at chunk offset 8127 with h = 0
    for k in (2, 3):
        for j, seq in enumerate(short_match_sequences(k, 2)):
            add(f"partial_warm_{k}B_{j}", [["splitmix", 31 + j, 8127], ["hex", seq], ["splitmix", 41 + j, 40000]])
            add(f"partial_warm_zeros_{k}B_{j}", [["zeros", 8127], ["hex", seq], ["zeros", 140000]])
    # forced cut boundary: W ending exactly at 131071 / 131072 / 131073 after zeros
    for size in (131071, 131072, 131073):
        add(f"W_ends_at_{size}", [["zeros", size - 64], ["hex", W_HEX], ["zeros", 9000]])
    # safetensors-shaped shard: 8-byte LE header length, JSON header, payload
    hdr = b'{"__metadata__":{"format":"pt"},"t0":{"dtype":"BF16","shape":[1024,1024],"data_offsets":[0,2097152]}}'
    hdr = hdr + b" " * (-len(hdr) % 8)
    add("safetensors_shaped_2MiB",
        [["hex", len(hdr).to_bytes(8, "little").hex()], ["hex", hdr.hex()], ["splitmix", 10, 2 * MiB]])
    if large:
        # BASELINE config 1 size: 1 GiB (SURVEY.md Appendix B lists this one as an hf_xet-only vector)
        add("one_GiB_16x64MiB_splitmix_seeds_100_115", [["splitmix", 100 + i, 64 * MiB] for i in range(16)], large=True)
        add("gt_4GiB_splitmix3_then_zeros", [["splitmix", 3, MiB], ["zeros", 4831838208 - MiB]], large=True)
    return c


def main():
    large = "--large" in sys.argv
    out_path = os.path.join(ROOT, "tests", "golden", "xet_golden.json")
    prev = {}
    if os.path.exists(out_path):
        prev = {e["name"]: e for e in json.load(open(out_path))["cases"]}
    entries = []
    tmpdir = tempfile.mkdtemp(dir="/dev/shm")
    for case in cases(large):
        data = xet_inputs.build(case["spec"])
        path = os.path.join(tmpdir, "f.bin")
        data.tofile(path)
        ref = hf_xet.hash_files([path])[0]
        os.unlink(path)
        assert ref.file_size == data.size
        res = xet_oracle.hash_file(data)
        ours = xet_oracle.to_hex(res.file_hash)
        e = dict(name=case["name"], spec=case["spec"], n=int(data.size), file_hash=ref.hash,
                 source="hf_xet 1.5.0 hash_files")
        if case.get("large"):
            e["large"] = True
        if ours != ref.hash:
            print(f"ORACLE MISMATCH {case['name']}: oracle {ours} ref {ref.hash}")
            e["oracle_match"] = False
        else:
            e["oracle_match"] = True
            ends = res.chunk_ends
            e["n_chunks"] = int(ends.size)
            e["cuts_head"] = [int(x) for x in ends[:16]]
            e["cuts_sha256"] = hashlib.sha256(ends.astype("<u8").tobytes()).hexdigest()
            e["chunk_hashes_sha256"] = hashlib.sha256(res.chunk_hashes.tobytes()).hexdigest()
            e["merkle_root"] = xet_oracle.to_hex(res.merkle_root)
            if ends.size <= 16:
                e["chunk_hashes"] = [xet_oracle.to_hex(h.tobytes()) for h in res.chunk_hashes]
        print(e["name"], e["n"], e.get("n_chunks"), e["file_hash"], "OK" if e["oracle_match"] else "MISMATCH")
        entries.append(e)
    names = {e["name"] for e in entries}
    for name, e in prev.items():      # keep --large entries when regenerating without --large
        if name not in names and e.get("large"):
            entries.append(e)
    # reference repo fixtures: hashes only (the files stay in /root/reference)
    ref_files = {}
    base = "/root/reference/pkg/hfutil/modelconfig/testdata/tiny-random-PhiModel"
    for fn in ("model-1-of-2.safetensors", "model-2-of-2.safetensors", "config.json"):
        p = os.path.join(base, fn)
        if os.path.exists(p):
            ref_files[os.path.relpath(p, "/root/reference")] = dict(
                n=os.path.getsize(p), file_hash=hf_xet.hash_files([p])[0].hash)
    os.rmdir(tmpdir)
    json.dump(dict(generator="scripts/gen_golden.py", ground_truth="hf_xet 1.5.0 (compiled xet-core)",
                   cases=entries, reference_fixture_files=ref_files),
              open(out_path, "w"), indent=1)
    bad = [e["name"] for e in entries if not e["oracle_match"]]
    print(f"{len(entries)} cases, {len(bad)} oracle mismatches {bad}")


if __name__ == "__main__":
    main()
