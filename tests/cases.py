"""Deterministic test-case generators shared by the CPU and GPU parity tests."""
from __future__ import annotations

import numpy as np

ALL_DTYPES = [np.uint8, np.int8, np.uint16, np.int16, np.uint32, np.int32, np.uint64, np.int64]
EDGE_COUNTS = [1, 63, 64, 65, 2047, 2048, 2049]


def splitmix64(x: np.ndarray) -> np.ndarray:
    """numpy splitmix64 finaliser over uint64 (same sequence as the device generator)."""
    with np.errstate(over="ignore"):
        x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
        x = ((x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)).astype(np.uint64)
        x = ((x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)).astype(np.uint64)
        return x ^ (x >> np.uint64(31))


def width_values(dtype, w: int, n: int, base: int, seed: int) -> np.ndarray:
    """n values of `dtype` whose (max - min) needs exactly w bits: base + U[0, 2^w), wrapped into T.
    Element 0 is base + 2^w - 1 and (n > 1) element 1 is base, so the width is exact."""
    dt = np.dtype(dtype)
    bits = 8 * dt.itemsize
    idx = np.arange(n, dtype=np.uint64) + np.uint64(seed) * np.uint64(1000003)
    r = splitmix64(idx)
    mask = np.uint64((1 << w) - 1) if w < 64 else np.uint64(0xFFFFFFFFFFFFFFFF)
    r &= mask
    r[0] = mask
    if n > 1:
        r[1] = 0
    with np.errstate(over="ignore"):
        v = (r + np.uint64(base & 0xFFFFFFFFFFFFFFFF)).astype(np.uint64)
    if bits < 64:
        v &= np.uint64((1 << bits) - 1)
    return v.astype(np.dtype(f"u{dt.itemsize}")).view(dt)


def sweep_cases(dtype, counts=EDGE_COUNTS, bases=(0, 5, -7, 1000003)):
    """(w, n, base) triples for a dtype: every width 1..8S x edge counts x a few frames of reference."""
    dt = np.dtype(dtype)
    out = []
    for w in range(1, 8 * dt.itemsize + 1):
        for i, n in enumerate(counts):
            base = bases[(w + i) % len(bases)]
            if dt.kind == "u" and base < 0:
                base = 3
            out.append((w, n, base))
    return out
