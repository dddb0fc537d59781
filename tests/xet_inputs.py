"""Deterministic test inputs, described by small JSON-able specs.

A spec is a list of segments, concatenated in order:
  ["zeros", n]            n bytes of 0x00
  ["mod251", n]           byte i (segment-relative) = i % 251
  ["splitmix", seed, n]   SURVEY.md §8(d) counter-based splitmix64 stream, from byte 0
  ["hex", "aabb.."]       literal bytes
  ["repeat", "hex", k]    k copies of a literal pattern
  ["ascii", "text"]       literal ASCII
The same specs drive scripts/gen_golden.py (hf_xet ground truth), the oracle tests and
the GPU parity tests, so fixtures never need to store input bytes.
"""
from __future__ import annotations

import numpy as np

MASK64 = (1 << 64) - 1

# 64-byte pattern whose full gear window hash is a boundary candidate (SURVEY.md App. B:
# bytes 22289..22352 of splitmix seed 7, G = 0x0000c50053d0c6d6).
W_HEX = ("74be1013d0e2cff5e8fa34a94c1bd1d1298b20f3067306d3310f5375acff7cc8"
         "bfbf708db3248a9220d2430954e13930250e8204f872d15de36d3773defe6ded")


def splitmix_bytes(seed: int, n: int, byte_off: int = 0) -> np.ndarray:
    """numpy restatement of the SURVEY.md §8(d) generator (independent of the C oracle)."""
    if n == 0:
        return np.zeros(0, dtype=np.uint8)
    w0 = byte_off // 8
    w1 = (byte_off + n + 7) // 8
    i = np.arange(w0 + 1, w1 + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = np.uint64(seed & MASK64) + i * np.uint64(0x9E3779B97F4A7C15)
        z ^= z >> np.uint64(30)
        z *= np.uint64(0xBF58476D1CE4E5B9)
        z ^= z >> np.uint64(27)
        z *= np.uint64(0x94D049BB133111EB)
        z ^= z >> np.uint64(31)
    b = z.astype("<u8").view(np.uint8)
    s = byte_off - w0 * 8
    return b[s:s + n].copy()


def build(spec) -> np.ndarray:
    parts = []
    for seg in spec:
        kind = seg[0]
        if kind == "zeros":
            parts.append(np.zeros(seg[1], dtype=np.uint8))
        elif kind == "mod251":
            parts.append((np.arange(seg[1], dtype=np.uint64) % 251).astype(np.uint8))
        elif kind == "splitmix":
            parts.append(splitmix_bytes(seg[1], seg[2]))
        elif kind == "hex":
            parts.append(np.frombuffer(bytes.fromhex(seg[1]), dtype=np.uint8))
        elif kind == "repeat":
            parts.append(np.tile(np.frombuffer(bytes.fromhex(seg[1]), dtype=np.uint8), seg[2]))
        elif kind == "ascii":
            parts.append(np.frombuffer(seg[1].encode(), dtype=np.uint8))
        else:
            raise ValueError(f"unknown segment kind {kind!r}")
    if not parts:
        return np.zeros(0, dtype=np.uint8)
    return np.ascontiguousarray(np.concatenate(parts))


def spec_len(spec) -> int:
    n = 0
    for seg in spec:
        if seg[0] in ("zeros", "mod251"):
            n += seg[1]
        elif seg[0] == "splitmix":
            n += seg[2]
        elif seg[0] == "hex":
            n += len(seg[1]) // 2
        elif seg[0] == "repeat":
            n += len(seg[1]) // 2 * seg[2]
        elif seg[0] == "ascii":
            n += len(seg[1])
    return n
