#!/usr/bin/env python
"""Mint the golden digests in tests/golden/golden.json from the REFERENCE'S OWN code path.

Runs only where /root/reference exists (this container): it drives oracle/_ref/libadacom_ref.so, i.e.
ref_harness.cpp linked against the reference's vendored SDSL (compiled from /root/reference by
oracle/Makefile), over the deterministic inputs of tests/cases.py, and records for every case

    width, min_factor, max_factor, compacted/repacked, size_in_bytes (== ColumnSegment::GetDataSize),
    sha256(packed payload words), sha256(serialize image), sha256(decoded values) in intended and literal
    mode, and the exact SUM of the decoded values.

The GPU box has no /root/reference; there the tests compare the L1 port, and the CUDA path, with this file.
Usage:  python tests/golden/make_golden.py          (rewrites golden.json)
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from tests.cases import ALL_DTYPES, EDGE_COUNTS, width_values  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden.json")


def sha(b: bytes) -> str:
    return hashlib.sha256(b).hexdigest()[:16]


def golden_cases():
    """(dtype name, w, n, base, flags) for the committed grid."""
    for dt in ALL_DTYPES:
        d = np.dtype(dt)
        bits = 8 * d.itemsize
        cap = oracle.BLOCK_SIZE // d.itemsize
        bases = [0, 5, 1000003] if d.kind == "u" else [0, 5, -7, -1000003]
        for w in range(1, bits + 1):
            for i, n in enumerate(EDGE_COUNTS):
                base = bases[(w + i) % len(bases)]
                for flags in (0, oracle.PADDED):
                    if flags and (w * 3 + i) % 4:
                        continue  # padded mode on a quarter of the grid
                    yield d.name, w, n, base, flags
        for w in sorted({1, bits // 2 + 1, bits - 1, bits}):
            for n in (cap - 1, cap):
                yield d.name, w, n, bases[w % len(bases)], 0


def record(o, name, w, n, base, flags, seed):
    vals = width_values(np.dtype(name), w, n, base, seed)
    seg = o.encode(vals, flags)
    m = seg.meta()
    dec = seg.decode_all(flags)
    dec_lit = seg.decode_all(flags | oracle.LITERAL)
    return [name, w, n, base, flags, seed, m.width, str(m.min_factor), str(m.max_factor),
            int(m.compacted) | (int(m.repacked) << 1), m.data_size, m.words,
            sha(seg.words().tobytes()) if m.repacked else "", sha(seg.serialize()) if m.repacked else "",
            sha(dec.tobytes()), sha(dec_lit.tobytes()), str(int(dec.astype(object).sum()))]


COLUMNS = ["dtype", "w", "n", "base", "flags", "seed", "width", "min_factor", "max_factor", "state", "size_in_bytes",
           "words", "payload_sha", "image_sha", "decoded_sha", "decoded_literal_sha", "sum"]


def main():
    o = oracle.Oracle("reference")
    entries = [record(o, *c, seed=i) for i, c in enumerate(golden_cases())]
    # a few complete vectors a reader can check by hand
    vectors = []
    for name, w, n, base in [("uint32", 11, 12, 100), ("int64", 40, 5, -1099511627776), ("uint8", 3, 70, 2), ("int16", 9, 9, -3000)]:
        vals = width_values(np.dtype(name), w, n, base, seed=99)
        seg = o.encode(vals, 0)
        vectors.append({"dtype": name, "values": [int(x) for x in vals], "width": seg.meta().width,
                        "min_factor": str(seg.meta().min_factor),
                        "words_hex": [f"{int(x):016x}" for x in seg.words()],
                        "image_hex": seg.serialize().hex()})
    # sizes the survey measured on the running reference: UINTEGER 0..199999
    survey = []
    v = np.arange(200000, dtype=np.uint32)
    for first, cnt in [(0, 2048), (2048, 65534), (67582, 55298), (122880, 65534), (188414, 11586)]:
        seg = o.encode(v[first:first + cnt], 0, segment_size=(2048 * 4 if first == 0 else oracle.BLOCK_SIZE))
        survey.append({"first": first, "count": cnt, "size_in_bytes": seg.data_size(), "width": seg.meta().width})
    doc = {"generator": "tests/golden/make_golden.py", "source": "oracle/_ref (reference SDSL, sdsl-lite c32874cb)",
           "columns": COLUMNS, "entries": entries, "vectors": vectors, "survey_uinteger_0_199999": survey}
    with open(OUT, "w") as f:
        json.dump(doc, f, separators=(",", ":"))
    print(f"{len(entries)} entries -> {OUT} ({os.path.getsize(OUT) >> 10} KiB)")
    print("survey sizes:", [s["size_in_bytes"] for s in survey], "sum", sum(s["size_in_bytes"] for s in survey))


if __name__ == "__main__":
    main()
