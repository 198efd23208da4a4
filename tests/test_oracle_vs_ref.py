"""CPU: the SDSL-free restatement (L1) against the reference's own SDSL (L0, oracle/_ref) on randomised
inputs -- every primitive the codec uses, both encoders, scan, uncompress, NULL handling."""
import numpy as np
import pytest

import oracle
from tests.cases import ALL_DTYPES, width_values


def same(a, b):
    return a.meta() == b.meta() and np.array_equal(a.words(), b.words()) and a.serialize() == b.serialize()


@pytest.mark.parametrize("dtype", ALL_DTYPES, ids=lambda d: np.dtype(d).name)
def test_append_path_random(port_oracle, ref_oracle, dtype):
    rng = np.random.default_rng(5)
    bits = 8 * np.dtype(dtype).itemsize
    for trial in range(60):
        w = int(rng.integers(1, bits + 1))
        n = int(rng.integers(1, 5000))
        base = int(rng.integers(-1000, 1000)) if np.dtype(dtype).kind == "i" else int(rng.integers(0, 1000))
        vals = width_values(dtype, w, n, base, trial)
        for flags in (0, oracle.PADDED, oracle.LITERAL):
            a, b = port_oracle.encode(vals, flags), ref_oracle.encode(vals, flags)
            assert same(a, b), (w, n, base, flags)
            s = int(rng.integers(0, n))
            c = int(rng.integers(0, n - s + 1))
            assert np.array_equal(a.scan_partial(s, c, flags, result_offset=3), b.scan_partial(s, c, flags, result_offset=3))
            assert np.array_equal(a.uncompress(flags), b.uncompress(flags))


@pytest.mark.parametrize("dtype", [np.uint8, np.int16, np.uint32, np.int32, np.int64], ids=lambda d: np.dtype(d).name)
def test_adaptive_path_random(port_oracle, ref_oracle, dtype):
    rng = np.random.default_rng(6)
    bits = 8 * np.dtype(dtype).itemsize
    for trial in range(40):
        w = int(rng.integers(1, bits + 1))
        n = int(rng.integers(1, 4000))
        vals = width_values(dtype, w, n, int(rng.integers(0, 500)), trial)
        if trial % 5 == 0:
            vals[:] = vals[0]  # single distinct value: the adaptive width rule differs (max == min)
        for flags in (0, oracle.LITERAL, oracle.PADDED):
            a, b = port_oracle.encode_adaptive(vals, flags), ref_oracle.encode_adaptive(vals, flags)
            assert same(a, b), (w, n, flags)
            assert np.array_equal(a.decode_all(flags), b.decode_all(flags))
            assert np.array_equal(a.uncompress(flags), b.uncompress(flags))


def test_validity_and_selection(port_oracle, ref_oracle):
    rng = np.random.default_rng(8)
    for dtype in (np.int8, np.int32, np.uint64):
        n = 3000
        vals = width_values(dtype, 7, n, 11, 1)
        valid = rng.random(n) > 0.3
        words = np.zeros((n + 63) // 64, dtype=np.uint64)
        for i in np.nonzero(valid)[0]:
            words[i >> 6] |= np.uint64(1) << np.uint64(i & 63)
        sel = rng.permutation(n).astype(np.uint32)
        segs = []
        for o in (port_oracle, ref_oracle):
            seg = o.new_succinct(dtype)
            seg.append(vals, 0, 2048, sel=sel, validity=words)
            seg.append(vals, 2048, n - 2048, sel=sel, validity=words)
            seg.bit_compress(0)
            segs.append(seg)
        assert same(*segs)
        assert np.array_equal(segs[0].decode_all(), segs[1].decode_all())


def test_bits_hi(port_oracle, ref_oracle):
    rng = np.random.default_rng(9)
    for x in [0, 1, 2, 3, 2**63, 2**64 - 1] + [int(v) for v in rng.integers(0, 2**63, size=500)]:
        assert port_oracle.bits_hi(x) == ref_oracle.bits_hi(x)


def test_cpu_baseline_harness_agrees(port_oracle, ref_oracle):
    v = width_values(np.int64, 23, 200000, 77, 3)
    a = port_oracle.bench_column(v, 32767, 3)
    b = ref_oracle.bench_column(v, 32767, 2)
    assert a["sum"] == b["sum"] == int(v.sum()) % (1 << 128)
    assert a["packed_bytes"] == b["packed_bytes"]
