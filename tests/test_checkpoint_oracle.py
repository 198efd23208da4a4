"""CPU tests of the checkpoint-image restatement (oracle/checkpoint.py): the block checksum against the reference's
own duckdb::Checksum and known answers, the partial-block placement on hand-worked cases, image round trips."""
from __future__ import annotations

import ctypes as C
import struct

import numpy as np
import pytest

from oracle import checkpoint as ck
from tests.cases import width_values

M64 = (1 << 64) - 1
MUL = 0xBF58476D1CE4E5B9


def test_checksum_known_answers():
    # checksum.cpp:6-17: result = 5381; result ^= word * 0xbf58476d1ce4e5b9 (mod 2^64) for every word
    assert ck.checksum(bytes(8)) == 5381
    assert ck.checksum(bytes(ck.BLOCK_SIZE)) == 5381
    assert ck.checksum(struct.pack("<Q", 1)) == 5381 ^ MUL
    assert ck.checksum(struct.pack("<QQ", 2, 3)) == 5381 ^ ((2 * MUL) & M64) ^ ((3 * MUL) & M64)
    assert ck.checksum(struct.pack("<QQ", 7, 7)) == 5381  # XOR: equal words cancel


def test_checksum_against_reference(ref_oracle):
    """oracle/_ref links the reference's src/common/checksum.cpp as it lies."""
    fn = ref_oracle.lib.ref_checksum
    fn.restype = C.c_uint64
    fn.argtypes = [C.c_void_p, C.c_uint64]
    rng = np.random.default_rng(7)
    for n in (8, 16, 4096, ck.BLOCK_SIZE):
        buf = rng.integers(0, 256, size=n, dtype=np.uint8)
        assert ck.checksum(buf) == fn(buf.ctypes.data_as(C.c_void_p), n)


def test_block_placement_hand_worked():
    # partial_block_manager.cpp:14-77 with max_partial_block_size = 209708, BLOCK_SIZE = 262136
    p = ck.BlockPlacer()
    assert p.place(100000) == (0, 0)        # new block; 162136 left >= 52428 -> stays partial
    assert p.place(100000) == (0, 100000)   # lower_bound(100000) -> the 162136 entry; 62136 left -> still partial
    assert p.place(100000) == (1, 0)        # 62136 < 100000 -> new block
    assert p.place(60000) == (0, 200000)    # least sufficient space: 62136 (block 0), not 162136 (block 1)
    assert p.place(262136) == (2, 0)        # > 80 % of a block: always its own block
    assert p.place(209709) == (3, 0)        # just above the partial limit
    assert p.place(150000) == (1, 100000)   # block 1 still has 162136
    assert p.n_blocks == 4
    # offsets are 8-byte aligned (AlignValue, :58)
    q = ck.BlockPlacer()
    assert q.place(13) == (0, 0)
    assert q.place(5) == (0, 16)
    assert q.place(8) == (0, 24)
    assert q.place(1) == (0, 32)
    # a block that ends up less than 20 % free leaves the map (:62-65)
    r = ck.BlockPlacer()
    assert r.place(209708) == (0, 0)        # 209712 used, 52424 left < 52428
    assert r.place(8) == (1, 0)
    # equal keys: the earliest registered block is reused first (multimap order)
    s = ck.BlockPlacer()
    assert s.place(250000) == (0, 0) and s.place(250000) == (1, 0)  # 12136 left each: not partial
    assert s.place(131072) == (2, 0) and s.place(131072) == (3, 0)  # 131064 left each
    assert s.place(1000) == (2, 131072)
    assert s.place(131064) == (3, 131072)


def _segments(port_oracle, dtype, specs):
    segs = []
    for (w, n, base, compact) in specs:
        v = width_values(dtype, w, n, base, seed=w * 31 + n) if n else np.zeros(0, dtype)
        if n and compact:
            o = port_oracle.encode(v, 0)
            m = o.meta()
            payload = o.words().tobytes() if m.repacked else v.tobytes()
            state = 1 | (2 if m.repacked else 0) | 4
            segs.append(dict(count=n, capacity=ck.BLOCK_SIZE // np.dtype(dtype).itemsize, width=m.width, state=state,
                             min_factor=m.min_factor, max_factor=m.max_factor, payload=payload, values=v))
        else:
            segs.append(dict(count=n, capacity=ck.BLOCK_SIZE // np.dtype(dtype).itemsize,
                             width=8 * np.dtype(dtype).itemsize, state=0, min_factor=M64, max_factor=0,
                             payload=v.tobytes(), values=v))
    return segs


@pytest.mark.parametrize("dtype", [np.uint8, np.int16, np.uint32, np.int64])
def test_image_round_trip(port_oracle, dtype):
    S = np.dtype(dtype).itemsize
    cap = ck.BLOCK_SIZE // S
    specs = [(3, 2049, 5, True), (8 * S, 777, 0, True), (5, 0, 0, True), (min(13, 8 * S - 1), cap, 3, True),
             (2, 65, 0, False), (1, 1, 9, True)]
    segs = _segments(port_oracle, dtype, specs)
    img = ck.build_image(9, S, segs)
    hdr, ptrs = ck.parse_image(img)
    assert hdr["n_segments"] == len(segs) and hdr["n_values"] == sum(s["count"] for s in segs)
    assert hdr["blocks_offset"] == 4096 and hdr["total_bytes"] == img.size
    row = 0
    for s, p in zip(segs, ptrs):
        assert (p["row_start"], p["tuple_count"], p["width"], p["state"]) == (row, s["count"], s["width"], s["state"])
        assert ck.payload_of(img, hdr, p) == (bytes(s["payload"]) if s["count"] else b"")
        assert p["offset"] % 8 == 0 and (p["block_id"] == -1) == (s["count"] == 0)
        row += s["count"]
    # every byte outside header, pointers and payloads is zero: the image is deterministic
    used = np.zeros(img.size, dtype=bool)
    used[: 64 + 64 * len(segs)] = True
    for b in range(hdr["n_blocks"]):
        at = hdr["blocks_offset"] + b * ck.BLOCK_ALLOC_SIZE
        used[at: at + 8] = True
    for p in ptrs:
        if p["block_id"] >= 0:
            at = hdr["blocks_offset"] + p["block_id"] * ck.BLOCK_ALLOC_SIZE + 8 + p["offset"]
            used[at: at + p["payload_bytes"]] = True
    assert not img[~used].any()
    # corruption is detected: payload byte, checksum byte, pointer table
    for at in (hdr["blocks_offset"] + 8 + 100, hdr["blocks_offset"] + 3, 64 + 17):
        bad = img.copy()
        bad[at] ^= 0x40
        with pytest.raises(ValueError):
            ck.parse_image(bad)
