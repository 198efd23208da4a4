"""CPU: the L1 port (and, where built, the L0 reference harness) against the committed golden digests that
tests/golden/make_golden.py minted from the reference's own SDSL code, plus the known answers the reference
tree itself records."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle
from tests.cases import width_values

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))
COLS = {c: i for i, c in enumerate(GOLDEN["columns"])}


def sha(b: bytes) -> str:
    return hashlib.sha256(b).hexdigest()[:16]


def check_entry(o, row):
    g = lambda k: row[COLS[k]]  # noqa: E731
    vals = width_values(np.dtype(g("dtype")), g("w"), g("n"), g("base"), g("seed"))
    flags = g("flags")
    seg = o.encode(vals, flags)
    m = seg.meta()
    assert m.width == g("width"), row
    assert (str(m.min_factor), str(m.max_factor)) == (g("min_factor"), g("max_factor")), row
    assert (int(m.compacted) | (int(m.repacked) << 1)) == g("state"), row
    assert m.data_size == g("size_in_bytes") and m.words == g("words"), row
    if m.repacked:
        assert sha(seg.words().tobytes()) == g("payload_sha"), row
        assert sha(seg.serialize()) == g("image_sha"), row
    dec = seg.decode_all(flags)
    assert sha(dec.tobytes()) == g("decoded_sha"), row
    assert sha(seg.decode_all(flags | oracle.LITERAL).tobytes()) == g("decoded_literal_sha"), row
    assert str(int(dec.astype(object).sum())) == g("sum"), row


@pytest.mark.parametrize("part", range(8))
def test_port_matches_golden(port_oracle, part):
    rows = GOLDEN["entries"][part::8]
    assert rows
    for row in rows:
        check_entry(port_oracle, row)


def test_reference_still_matches_golden(ref_oracle):
    for row in GOLDEN["entries"][::7]:
        check_entry(ref_oracle, row)


def test_hand_checkable_vectors(port_oracle):
    for v in GOLDEN["vectors"]:
        vals = np.array(v["values"], dtype=np.dtype(v["dtype"]))
        seg = port_oracle.encode(vals, 0)
        assert seg.meta().width == v["width"] and str(seg.meta().min_factor) == v["min_factor"]
        assert [f"{int(x):016x}" for x in seg.words()] == v["words_hex"]
        assert seg.serialize().hex() == v["image_hex"]
        # independent restatement of the layout: value i sits at bits [i*w, (i+1)*w) LSB-first
        w, mn = v["width"], int(v["min_factor"])
        big = int.from_bytes(seg.words().tobytes(), "little")
        bits = 8 * vals.dtype.itemsize
        for i, x in enumerate(v["values"]):
            field = (big >> (i * w)) & ((1 << w) - 1)
            assert (field + mn) % (1 << bits) == x % (1 << bits)


def test_survey_sizes_uinteger_0_199999(port_oracle):
    """SURVEY.md 6.3, measured on the running reference: 2825 + 131081 + 110609 + 131081 + 20289 = 395885 B."""
    v = np.arange(200000, dtype=np.uint32)
    total = 0
    for s in GOLDEN["survey_uinteger_0_199999"]:
        seg = port_oracle.encode(v[s["first"]:s["first"] + s["count"]], 0,
                                 segment_size=2048 * 4 if s["first"] == 0 else oracle.BLOCK_SIZE)
        assert seg.data_size() == s["size_in_bytes"] and seg.meta().width == s["width"]
        total += seg.data_size()
    assert total == 395885
    assert [s["size_in_bytes"] for s in GOLDEN["survey_uinteger_0_199999"]] == [2825, 131081, 110609, 131081, 20289]


def test_sdsl_known_answers(port_oracle):
    """Sizes printed by the authors' own SDSL experiments (experiments/data/sdsl_std/sdsl_fixed_benchmark.log:
    2..16 elements -> 17 B compressed / 25 B for two 64-bit slots; 1 048 576 elements at 20 bit -> 2 621 449 B,
    byte-aligned 24 bit -> 3 145 737 B) and bits::hi pins from sdsl-lite/test/bits_test.cpp."""
    assert 9 + 8 * ((2 * 64 + 63) // 64) == 25
    n = 1 << 20
    vals = (np.arange(n, dtype=np.uint64) * np.uint64(2654435761) % np.uint64(1 << 20)).astype(np.uint32)
    vals[0], vals[1] = (1 << 20) - 1, 0
    seg = port_oracle.encode(vals, 0, segment_size=n * 4)
    assert seg.meta().width == 20 and seg.data_size() == 2621449
    seg = port_oracle.encode(vals, oracle.PADDED, segment_size=n * 4)
    assert seg.meta().width == 24 and seg.data_size() == 3145737
    for cnt in (2, 4, 8, 16):
        seg = port_oracle.encode(np.arange(cnt, dtype=np.uint32), 0, segment_size=cnt * 4)
        assert seg.data_size() == 17
    assert port_oracle.bits_hi(0) == 0 and port_oracle.bits_hi(1) == 0
    for i in range(64):
        assert port_oracle.bits_hi(1 << i) == i
        assert port_oracle.bits_hi((1 << i) | ((1 << i) - 1)) == i


def test_quirk_domain_documented(port_oracle):
    """Q1: mixed-sign segments are not re-packed and the literal scan adds min anyway (corrupting values);
    Q7: an all -1 segment has min_factor == UINT64_MAX, which the reference treats as 'unset'."""
    seg = port_oracle.encode(np.array([-5, 8, 9], dtype=np.int32), 0)
    m = seg.meta()
    assert m.compacted and not m.repacked and m.width == 32
    assert list(seg.decode_all(0)) == [-5, 8, 9]
    assert list(seg.decode_all(oracle.LITERAL)) != [-5, 8, 9]
    seg = port_oracle.encode(np.array([-1, -1, -1], dtype=np.int16), 0)
    assert seg.meta().min_factor == 2**64 - 1 and seg.meta().width == 1
    assert list(seg.decode_all(0)) == [1, 1, 1]
