"""GPU parity against the committed golden digests (minted from the reference's own SDSL code, see
tests/golden/make_golden.py): payload words, serialize image, metadata, sizes, decoded values, SUM --
all through the C ABI, every entry of a dtype batched into ONE column / one launch per operation."""
import hashlib
import json
import os

import numpy as np
import pytest

from tests.cases import ALL_DTYPES, width_values

pytestmark = pytest.mark.gpu

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))
COLS = {c: i for i, c in enumerate(GOLDEN["columns"])}


def sha(b: bytes) -> str:
    return hashlib.sha256(b).hexdigest()[:16]


@pytest.mark.parametrize("dtype", ALL_DTYPES, ids=lambda d: np.dtype(d).name)
@pytest.mark.parametrize("flags", [0, 1])
def test_cuda_path_matches_golden(gpu_ctx, dtype, flags):
    name = np.dtype(dtype).name
    rows = [r for r in GOLDEN["entries"] if r[COLS["dtype"]] == name and r[COLS["flags"]] == flags]
    assert rows
    vals = [width_values(np.dtype(name), r[COLS["w"]], r[COLS["n"]], r[COLS["base"]], r[COLS["seed"]]) for r in rows]
    col = gpu_ctx.column_from_values(np.concatenate(vals), [len(v) for v in vals])
    col.compact(flags=flags)
    metas = col.meta()
    arena = col.download_packed()
    dec = col.decode(flags=flags)
    dec_lit = col.decode(flags=flags | 2)
    off = 0
    total = 0
    for g, (r, v) in enumerate(zip(rows, vals)):
        m = metas[g]
        assert m.width == r[COLS["width"]], r
        assert (str(m.min_factor), str(m.max_factor)) == (r[COLS["min_factor"]], r[COLS["max_factor"]]), r
        assert (m.state & 3) == r[COLS["state"]], r
        assert m.size_in_bytes == r[COLS["size_in_bytes"]], r
        if m.state & 2:
            assert m.words == r[COLS["words"]]
            assert sha(arena[m.word_offset:m.word_offset + m.words].tobytes()) == r[COLS["payload_sha"]], r
            if g % 9 == 0:
                assert sha(col.serialize(g)) == r[COLS["image_sha"]], r
        assert sha(dec[off:off + len(v)].tobytes()) == r[COLS["decoded_sha"]], r
        assert sha(dec_lit[off:off + len(v)].tobytes()) == r[COLS["decoded_literal_sha"]], r
        total += int(r[COLS["sum"]])
        off += len(v)
    agg = col.scan_sum(flags=flags)
    assert agg.count == off and agg.sum % (1 << 128) == total % (1 << 128)
    col.close()


def test_hand_checkable_vectors(gpu_ctx):
    for v in GOLDEN["vectors"]:
        vals = np.array(v["values"], dtype=np.dtype(v["dtype"]))
        col = gpu_ctx.column_from_values(vals, [len(vals)])
        col.compact()
        assert [f"{int(x):016x}" for x in col.segment_words(0)] == v["words_hex"]
        assert col.serialize(0).hex() == v["image_hex"]
        col.close()


def test_survey_sizes(gpu_ctx):
    """GetDataSize accounting of UINTEGER 0..199999 as measured on the running reference."""
    v = np.arange(200000, dtype=np.uint32)
    s = GOLDEN["survey_uinteger_0_199999"]
    # the first segment of a fresh collection has 2048 slots, the others the full block (column_data.cpp:343-351)
    first = gpu_ctx.column_from_values(v[:2048], [2048], seg_capacity=2048)
    first.compact()
    rest = gpu_ctx.column_from_values(v[2048:], [x["count"] for x in s[1:]])
    rest.compact()
    sizes = [first.data_size()] + [m.size_in_bytes for m in rest.meta()]
    assert sizes == [x["size_in_bytes"] for x in s] and sum(sizes) == 395885
    first.close()
    rest.close()
