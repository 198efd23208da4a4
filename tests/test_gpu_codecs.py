"""SURVEY 8(f) row 4 -- DuckDB's bitpacking and RLE segments decoded on the GPU, against the REFERENCE's own code.

Inputs: the segment images the reference's BitpackingState / RLEState + fastpforlib produce (oracle/_ref/
libadacom_codecs_ref.so = oracle/l2/codecs_ref.cpp, which includes the reference's bitpacking.cpp and rle.cpp where they
lie).  Expected: that library's scan / fetch_row.  Bit-exact: decoded values for every width of every integer type and
every bitpacking mode, fused SUM, batched fetch_row, sub-ranges of segments, ragged tails."""
import numpy as np
import pytest

from oracle import codecs
from tests.cases import ALL_DTYPES, splitmix64

pytestmark = pytest.mark.gpu
MOD = 1 << 128


def _need_oracle():
    if not codecs.available():
        pytest.skip("oracle/_ref/libadacom_codecs_ref.so not built (needs /root/reference + oracle/l2/build_drivers.sh)")


def _width_sweep_values(dtype, n_per_width=2048 * 2):
    """groups whose FOR width sweeps 1 .. 8S (value = base + U[0, 2^w)), plus constant, constant-delta and sorted
    (delta-friendly) stretches and a ragged tail"""
    dt = np.dtype(dtype)
    bits = 8 * dt.itemsize
    parts = []
    # signed T: the reference only bit-packs a group when max - min fits T (TrySubtractOperator, bitpacking.cpp:147-149)
    for w in range(1, bits + (1 if dt.kind == "u" else 0)):
        r = splitmix64(np.arange(n_per_width, dtype=np.uint64) + np.uint64(w * 7919)) & np.uint64((1 << w) - 1 if w < 64 else 2**64 - 1)
        base = 0 if w >= bits - 1 else (5 if dt.kind == "u" else -(1 << (w - 1)) if w > 1 else -1)
        parts.append((r + np.uint64(base & (2**64 - 1))).astype(np.uint64))
    vals = np.concatenate(parts)
    if bits < 64:
        vals &= np.uint64((1 << bits) - 1)
    out = vals.astype(np.dtype(f"u{dt.itemsize}")).view(dt).copy()
    n = len(out)
    out[2048 * 3:2048 * 4] = 42                                                    # CONSTANT
    if dt.itemsize > 1:
        out[2048 * 5:2048 * 6] = (np.arange(2048) * 3 + 10).astype(dt)             # CONSTANT_DELTA
        out[2048 * 8:2048 * 12] = np.sort(out[2048 * 8:2048 * 12])                 # DELTA_FOR candidates
        out[2048 * 14:2048 * 15] = np.cumsum(splitmix64(np.arange(2048, dtype=np.uint64)) % np.uint64(5)).astype(dt)
    return out[: n - 777]                                                          # ragged last group


@pytest.mark.parametrize("dtype", ALL_DTYPES, ids=lambda d: np.dtype(d).name)
def test_bitpacking_decode_sum_fetch(gpu_ctx, dtype):
    _need_oracle()
    import duckdb_adaptive_compression_b200 as adacom
    v = _width_sweep_values(dtype)
    ref = codecs.RefCodecColumn(v)
    assert sum(ref.counts) == len(v)
    modes = {m for s in range(len(ref.counts)) for m in ref.modes(s)}
    assert codecs.MODE_FOR in modes and codecs.MODE_CONSTANT in modes
    col = adacom.CodecColumn(gpu_ctx, adacom.CODEC_BITPACKING, dtype, ref.images, ref.counts)
    want = np.concatenate([ref.scan(s, 0, c) for s, c in enumerate(ref.counts)])
    assert np.array_equal(want, v)                      # the reference round-trips its own encoding
    got = col.decode()
    assert np.array_equal(got, want)
    agg = col.scan_sum()
    assert agg.count == len(v) and agg.sum % MOD == int(v.astype(object).sum()) % MOD
    if len(ref.counts) > 1:
        assert np.array_equal(col.decode(1, len(ref.counts)), want[ref.counts[0]:])
        assert col.scan_sum(1, len(ref.counts)).sum % MOD == int(v[ref.counts[0]:].astype(object).sum()) % MOD
    rng = np.random.default_rng(4)
    q_seg = rng.integers(0, len(ref.counts), size=1500)
    q_row = np.array([rng.integers(0, ref.counts[s]) for s in q_seg])
    starts = np.concatenate([[0], np.cumsum(ref.counts)])
    fetched = col.fetch_rows(q_seg, q_row)
    assert np.array_equal(fetched, v[starts[q_seg] + q_row])
    for s, r in list(zip(q_seg, q_row))[:60]:
        assert ref.fetch_row(int(s), int(r)) == v[starts[s] + r]     # BitpackingFetchRow itself agrees
    col.close()


@pytest.mark.parametrize("mode", [codecs.MODE_FOR, codecs.MODE_DELTA_FOR], ids=["for", "delta_for"])
@pytest.mark.parametrize("dtype", [np.int16, np.uint32, np.int64], ids=lambda d: np.dtype(d).name)
def test_bitpacking_forced_modes(gpu_ctx, dtype, mode):
    """force_bitpacking_mode (bitpacking.cpp:129, 361): every group FOR, or DELTA_FOR wherever deltas are representable"""
    _need_oracle()
    import duckdb_adaptive_compression_b200 as adacom
    rng = np.random.default_rng(8)
    hi = min(np.iinfo(dtype).max // 4, 1 << 40)
    v = np.sort(rng.integers(0, hi, size=50_000)).astype(dtype)
    v[20_000:30_000] = rng.integers(0, 1000, size=10_000).astype(dtype)
    ref = codecs.RefCodecColumn(v, mode=mode)
    col = adacom.CodecColumn(gpu_ctx, adacom.CODEC_BITPACKING, dtype, ref.images, ref.counts)
    assert np.array_equal(col.decode(), v)
    assert col.scan_sum().sum % MOD == int(v.astype(object).sum()) % MOD
    rows = rng.integers(0, ref.counts[0], size=500)
    assert np.array_equal(col.fetch_rows(np.zeros(500, dtype=np.uint32), rows), v[rows])
    col.close()


def test_bitpacking_with_nulls(gpu_ctx):
    """NULL rows keep the stale buffer slot (bitpacking.cpp:279-287): whatever the reference decodes there, byte for byte"""
    _need_oracle()
    import duckdb_adaptive_compression_b200 as adacom
    rng = np.random.default_rng(2)
    v = (rng.integers(0, 1 << 11, size=30_000) - 300).astype(np.int32)
    valid = rng.random(len(v)) > 0.2
    valid[2048 * 2:2048 * 3] = False      # an all-NULL group -> CONSTANT
    ref = codecs.RefCodecColumn(v, validity=valid)
    col = adacom.CodecColumn(gpu_ctx, adacom.CODEC_BITPACKING, np.int32, ref.images, ref.counts)
    want = np.concatenate([ref.scan(s, 0, c) for s, c in enumerate(ref.counts)])
    assert np.array_equal(want[valid], v[valid])
    assert np.array_equal(col.decode(), want)
    col.close()


@pytest.mark.parametrize("dtype", ALL_DTYPES, ids=lambda d: np.dtype(d).name)
def test_rle_decode_sum_fetch(gpu_ctx, dtype):
    _need_oracle()
    import duckdb_adaptive_compression_b200 as adacom
    rng = np.random.default_rng(6)
    info = np.iinfo(dtype)
    n_runs = 60_000 if np.dtype(dtype).itemsize >= 4 else 150_000   # several segments (max_rle_count runs each)
    base = rng.integers(max(info.min, -(1 << 40)), min(info.max, 1 << 40), size=n_runs, endpoint=True).astype(dtype)
    lens = rng.integers(1, 9, size=n_runs)
    lens[100] = 70_000                                              # a run longer than the u16 count limit
    v = np.repeat(base, lens)
    ref = codecs.RefCodecColumn(v, codec=codecs.RLE)
    assert sum(ref.counts) == len(v) and len(ref.counts) > 1
    col = adacom.CodecColumn(gpu_ctx, adacom.CODEC_RLE, dtype, ref.images, ref.counts)
    want = np.concatenate([ref.scan(s, 0, c) for s, c in enumerate(ref.counts)])
    assert np.array_equal(want, v)
    assert np.array_equal(col.decode(), want)
    agg = col.scan_sum()
    assert agg.count == len(v) and agg.sum % MOD == int(v.astype(object).sum()) % MOD
    starts = np.concatenate([[0], np.cumsum(ref.counts)])
    assert np.array_equal(col.decode(1, 2), v[starts[1]:starts[2]])
    q_seg = rng.integers(0, len(ref.counts), size=1000)
    q_row = np.array([rng.integers(0, ref.counts[s]) for s in q_seg])
    assert np.array_equal(col.fetch_rows(q_seg, q_row), v[starts[q_seg] + q_row])
    col.close()
