"""CPU tests of the oracle: it must reproduce the reference arithmetic (hf_xet golden vectors)
before it may be used to judge the CUDA path.  No GPU needed."""
import hashlib
import os
import re
import struct

import numpy as np
import pytest

import xet_inputs
import xet_oracle
from conftest import ROOT, load_golden

GOLDEN = load_golden()
SMALL = [c for c in GOLDEN["cases"] if not c.get("large")]


def test_gear_table_digest():
    txt = open(os.path.join(ROOT, "include", "xet_gear_table.h")).read()
    vals = [int(x[:-3], 16) for x in re.findall(r"0x[0-9a-f]{16}ULL", txt)]
    assert len(vals) == 256
    digest = hashlib.sha256(b"".join(struct.pack("<Q", v) for v in vals)).hexdigest()
    assert digest == "e1d3936666d7ae7a977c958e9afcc75f90aaca758ce5fbe4ece61dffefe1912c"


def test_splitmix_generators_agree():
    assert xet_inputs.splitmix_bytes(1, 16).tobytes().hex() == "c15c0289ec2d0a9167ec8e65a18debbe"
    for seed, n, off in [(1, 1000, 0), (7, 4097, 13), (2**63 + 5, 100, 7)]:
        assert np.array_equal(xet_inputs.splitmix_bytes(seed, n, off), xet_oracle.splitmix_bytes(seed, n, off))


@pytest.mark.parametrize("case", SMALL, ids=[c["name"] for c in SMALL])
def test_oracle_matches_hf_xet_golden(case):
    assert case["oracle_match"]
    data = xet_inputs.build(case["spec"])
    assert data.size == case["n"]
    r = xet_oracle.hash_file(data)
    assert xet_oracle.to_hex(r.file_hash) == case["file_hash"]          # hf_xet ground truth
    assert r.chunk_ends.size == case["n_chunks"]
    assert [int(x) for x in r.chunk_ends[:16]] == case["cuts_head"]
    assert hashlib.sha256(r.chunk_ends.astype("<u8").tobytes()).hexdigest() == case["cuts_sha256"]
    assert hashlib.sha256(r.chunk_hashes.tobytes()).hexdigest() == case["chunk_hashes_sha256"]
    assert xet_oracle.to_hex(r.merkle_root) == case["merkle_root"]


def test_survey_stage_vector():
    """SURVEY.md Appendix B stage-by-stage intermediates for splitmix_seed1_1048576."""
    data = xet_inputs.build([["splitmix", 1, 1 << 20]])
    cand = xet_oracle.window_candidates(data)
    assert [int(x) for x in cand[:6]] == [25124, 104496, 330204, 378118, 380326, 401985]
    assert cand.size == 12
    r = xet_oracle.hash_file(data)
    assert [int(x) for x in r.chunk_ends] == [25125, 104497, 235569, 330205, 378119, 401986, 489058, 590648,
                                              717228, 777563, 834961, 966033, 1048576]
    assert xet_oracle.to_hex(r.chunk_hashes[0].tobytes()) == \
        "2520f307e75abcb6555d51e9019c4fb6eaa48883cb20c6fb5841450e5192f8b1"
    assert xet_oracle.to_hex(r.chunk_hashes[12].tobytes()) == \
        "327cc268a565a4718290ed133d0b028816306847f2c01763be5d668acb7269b9"
    assert xet_oracle.to_hex(r.merkle_root) == "5cb5e57306221297df4b92f4971739adf124cf6f7b505942a51b366d5d2ac29a"


def walk_candidates(cand, n):
    """SURVEY.md A.2 position-parallel form, in Python (what the CUDA path implements)."""
    cuts, start, i = [], 0, 0
    cand = [int(c) for c in cand]
    while start < n:
        hi = min(n, start + 131072)
        while i < len(cand) and cand[i] + 1 < start + 8192:
            i += 1
        cut = cand[i] + 1 if i < len(cand) and cand[i] + 1 <= hi else hi
        cuts.append(cut)
        start = cut
    return cuts


@pytest.mark.parametrize("name", ["splitmix_seed2_10MiB", "dense_W_x16384", "deadzone_after_forced",
                                  "W_ends_at_8191", "W_ends_at_8192", "partial_warm_3B_0",
                                  "W_ends_at_131072", "zeros_131073", "safetensors_shaped_2MiB"])
def test_position_parallel_form_equals_sequential_chunker(name):
    case = next(c for c in SMALL if c["name"] == name)
    data = xet_inputs.build(case["spec"])
    seq = [int(x) for x in xet_oracle.chunk_boundaries(data)]
    par = walk_candidates(xet_oracle.window_candidates(data), data.size)
    assert seq == par


def test_keyed_blake3_against_blake3_package():
    blake3 = pytest.importorskip("blake3")
    rng = np.random.default_rng(3)
    key = bytes(rng.integers(0, 256, 32, dtype=np.uint8))
    for n in [0, 1, 63, 64, 65, 1023, 1024, 1025, 2048, 2049, 3072, 3073, 4096, 5000, 8192, 31744, 65536, 131072]:
        d = bytes(rng.integers(0, 256, n, dtype=np.uint8))
        assert xet_oracle.blake3_keyed(key, d) == blake3.blake3(d, key=key).digest()


def test_hex_roundtrip():
    h = bytes(range(32))
    s = xet_oracle.to_hex(h)
    assert s.startswith("0706050403020100")
    assert xet_oracle.from_hex(s) == h
    with pytest.raises(ValueError):
        xet_oracle.from_hex("zz")


def test_reference_fixture_files():
    """The reference's own safetensors fixtures (pkg/hfutil/modelconfig/testdata) — only where
    /root/reference is mounted (never on the GPU box)."""
    base = "/root/reference"
    if not os.path.isdir(base):
        pytest.skip("reference tree not mounted")
    for rel, exp in GOLDEN["reference_fixture_files"].items():
        data = np.fromfile(os.path.join(base, rel), dtype=np.uint8)
        assert data.size == exp["n"]
        assert xet_oracle.to_hex(xet_oracle.hash_file(data).file_hash) == exp["file_hash"]
