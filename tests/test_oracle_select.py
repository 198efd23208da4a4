"""CPU: the oracle's FilterSelection restatement (orc_segment_select) against plain numpy set logic."""
import numpy as np
import pytest

from tests.cases import width_values

OPS = [np.equal, np.not_equal, np.less, np.greater, np.less_equal, np.greater_equal]


@pytest.mark.parametrize("dtype", [np.uint8, np.int16, np.uint32, np.int64], ids=lambda d: np.dtype(d).name)
def test_select_matches_numpy(port_oracle, dtype):
    rng = np.random.default_rng(3)
    S = np.dtype(dtype).itemsize
    for w, n, base in [(3, 100, 0), (min(11, 8 * S - 1), 5000, -7 if np.dtype(dtype).kind == "i" else 9), (8 * S, 2049, 0)]:
        v = width_values(dtype, w, n, base, seed=w)
        valid = rng.random(n) > 0.3
        words = np.packbits(np.pad(valid, (0, (-n) % 64)), bitorder="little").view(np.uint64).copy()
        seg = port_oracle.encode(v, 0, validity=words)
        dec = seg.decode_all()
        sel_in = rng.random(n) > 0.5
        in_words = np.packbits(np.pad(sel_in, (0, (-n) % 64)), bitorder="little").view(np.uint64).copy()
        c1, c2 = (dtype(x) for x in sorted(rng.choice(dec, size=2)))
        for conj in (0, 1):
            for preds, masks in [
                ([("cmp", 2, c1), ("cmp", 3, c2)], [valid & (dec < c1), valid & (dec > c2)]),
                ([("cmp", 5, c1), ("cmp", 4, c2)], [valid & (dec >= c1), valid & (dec <= c2)]),
                ([("is_null",), ("cmp", 0, c1)], [~valid, valid & (dec == c1)]),
                ([("is_not_null",)], [valid]),
            ]:
                want = np.logical_or.reduce(masks) if conj else np.logical_and.reduce(masks)
                want &= sel_in
                out = np.zeros((n + 63) // 64, dtype=np.uint64)
                sm, cnt, nv = seg.select(preds, conj, validity=words, bitmap_in=in_words, bitmap_out=out)
                got = np.unpackbits(out.view(np.uint8), bitorder="little")[:n].astype(bool)
                assert np.array_equal(got, want), (preds, conj)
                assert cnt == int(want.sum()) and nv == int((want & valid).sum())
                assert sm % (1 << 128) == int(dec[want & valid].astype(object).sum()) % (1 << 128)
        # single comparison without NULLs == the older orc_segment_filter
        for op in range(6):
            sm, cnt, bm = seg.filter(op, int(c1), want_bitmap=True)
            sm2, cnt2, _ = seg.select([("cmp", op, int(c1))])
            assert (sm, cnt) == (sm2, cnt2)
            assert cnt == int(OPS[op](dec, c1).sum())
