"""Pull-model scan stream (adacom_scan_stream_*, the vtable path of ColumnData::ScanVector, column_data.cpp:92-139):
every segment served from the pinned window ring equals the oracle's SuccinctScanPartial output / the original values,
for ragged segments, mixed compacted / uncompacted / slot-width states, revisits, restarts and concurrent streams.
Bit-exact."""
from __future__ import annotations

import threading

import numpy as np
import pytest

from tests.cases import width_values

pytestmark = pytest.mark.gpu


def _column(gpu_ctx, dtype, n_seg):
    S = np.dtype(dtype).itemsize
    bits = 8 * S
    parts = []
    for g in range(n_seg):
        w = 1 + (g * 7) % bits
        n = [2049, 63, 1, 4097, 700, 0, 32767 if S == 8 else 5000][g % 7]
        parts.append(width_values(dtype, w, n, [0, 5, 1000003][g % 3], seed=g + 1) if n else np.zeros(0, dtype=dtype))
    return gpu_ctx.column_from_values(np.concatenate(parts), [len(p) for p in parts]), parts


@pytest.mark.parametrize("dtype", [np.uint8, np.int16, np.uint32, np.int64], ids=lambda d: np.dtype(d).name)
@pytest.mark.parametrize("window,depth", [(0, 0), (1, 2), (3, 3), (5, 4), (64, 2)])
def test_stream_serves_every_segment(gpu_ctx, port_oracle, dtype, window, depth):
    col, parts = _column(gpu_ctx, dtype, 41)
    col.compact(0, 30)              # segments 30.. stay in the raw arena
    col.uncompact(4, 9)             # and a few go back to slot width
    # the batched decode is pinned to the oracle elsewhere (tests/test_gpu_parity.py); a one-value segment holding -1
    # decodes to 1 in both (reference quirk Q7), so the expectation is the codec's own decode
    want = col.decode()
    starts = np.concatenate([[0], np.cumsum([len(p) for p in parts])])
    st = col.scan_stream(window_segments=window, depth=depth)
    for g, p in enumerate(parts):
        got = st.segment(g)
        assert got.dtype == np.dtype(dtype) and np.array_equal(got, want[starts[g]:starts[g + 1]]), g
        if len(p) > 1:
            assert np.array_equal(got, p), g
        if g % 9 == 0 and len(p):
            seg = port_oracle.encode(p, 0)
            assert np.array_equal(got[:min(2048, len(p))], seg.scan_partial(0, min(2048, len(p))))
    st.close()
    col.close()


def test_sub_range_revisit_and_restart(gpu_ctx):
    col, parts = _column(gpu_ctx, np.int64, 29)
    col.compact()
    st = col.scan_stream(5, 27, window_segments=4, depth=3)
    with pytest.raises(Exception):
        st.segment(4)
    with pytest.raises(Exception):
        st.segment(27)
    want = col.decode()
    starts = np.concatenate([[0], np.cumsum([len(p) for p in parts])])
    order = [5, 6, 9, 8, 7, 10, 13, 12, 26, 5, 6, 20, 21, 22, 23, 24, 25, 26]  # revisits, a jump ahead, restarts
    for g in order:
        assert np.array_equal(st.segment(g), want[starts[g]:starts[g + 1]]), g
    st.close()
    # a view handed out stays intact while the stream moves one window on (zero-copy vectors of the previous pull)
    st = col.scan_stream(window_segments=2, depth=3)
    first = st.segment(0)
    keep = first.copy()
    st.segment(2)
    st.segment(3)
    assert np.array_equal(first, keep)
    st.close()
    col.close()


def test_concurrent_streams_and_a_writer(gpu_ctx):
    """Four reader threads with a stream each while the main thread re-compacts / uncompacts the column: the decoded
    values never change (the policy only changes the representation)."""
    col, parts = _column(gpu_ctx, np.uint32, 64)
    col.compact()
    parts = np.split(col.decode(), np.cumsum([len(p) for p in parts])[:-1])
    errors = []

    def reader(k):
        try:
            for rep in range(3):
                st = col.scan_stream(window_segments=1 + k, depth=2 + k % 2)
                for g, p in enumerate(parts):
                    if not np.array_equal(st.segment(g), p):
                        errors.append((k, rep, g))
                        return
                st.close()
        except Exception as e:  # noqa: BLE001
            errors.append((k, repr(e)))

    threads = [threading.Thread(target=reader, args=(k,)) for k in range(4)]
    for t in threads:
        t.start()
    for rep in range(20):
        col.uncompact_list(list(range(rep % 5, 64, 5)))
        col.compact_list(list(range(rep % 5, 64, 5)))
    for t in threads:
        t.join()
    assert not errors, errors[:3]
    assert np.array_equal(col.decode(), np.concatenate(parts))
    col.close()
