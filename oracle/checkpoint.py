"""CPU restatement of the checkpoint image of a succinct column -- TEST INFRASTRUCTURE ONLY.

Only tests/ may import this module; the product (duckdb_adaptive_compression_b200/csrc) assembles the image on
the GPU (kernels_checkpoint.cu) and never calls it.

What is restated, with the reference lines (paths relative to /root/reference):

* ``checksum``          duckdb::Checksum, src/common/checksum.cpp:6-25 (whole-word part; DuckDB blocks are a
                        multiple of 8 bytes).  PINNED against the reference's own function compiled into
                        oracle/_ref/libadacom_ref.so (``ref_checksum``) by tests/test_checkpoint_oracle.py.
* ``BlockPlacer``       ColumnCheckpointState::FlushSegment's allocation, src/storage/table/column_checkpoint_state.cpp
                        :107-139 + PartialBlockManager::{GetBlockAllocation, GetPartialBlock, RegisterPartialBlock},
                        src/storage/partial_block_manager.cpp:14-77 with the default limits
                        (partial_block_manager.hpp:69-74).  Parity UNPINNED by execution (the class needs the whole
                        DuckDB block manager to link); pinned by hand-worked cases in the test.
* ``build_image``       the image layout declared in include/adacom_gpu.h ("checkpoint image"): block = 8-byte
                        checksum + 262136 payload bytes (FileBuffer::ChecksumAndWrite, src/common/file_buffer.cpp:113-119).

Numpy integer arithmetic only (uint64 wrap-around multiply, XOR reduce); the placement loop is per segment.
"""
from __future__ import annotations

import bisect
import struct

import numpy as np

BLOCK_ALLOC_SIZE = 262144
BLOCK_SIZE = 262136
MAX_PARTIAL_BLOCK_SIZE = BLOCK_SIZE // 5 * 4      # partial_block_manager.hpp:69
INVALID_BLOCK = -1
MAGIC = b"ADACOMCK"
VERSION = 1
_MUL = np.uint64(0xBF58476D1CE4E5B9)


def checksum(buf) -> int:
    """duckdb::Checksum(buffer, size) for size % 8 == 0 (checksum.cpp:10-17)."""
    a = np.frombuffer(buf, dtype=np.uint8) if not isinstance(buf, np.ndarray) else buf.view(np.uint8).reshape(-1)
    assert a.size % 8 == 0
    w = a.view(np.uint64)
    with np.errstate(over="ignore"):
        prod = w * _MUL
    return int(np.bitwise_xor.reduce(prod, initial=np.uint64(5381)))


class BlockPlacer:
    """(block_id, offset_in_block) for each payload size, in FlushSegment order."""

    def __init__(self):
        self.keys = []     # space left, sorted ascending (std::multimap order: equal keys keep insertion order)
        self.vals = []     # (block_id, offset_in_block)
        self.n_blocks = 0

    def place(self, segment_size: int):
        assert 0 < segment_size <= BLOCK_SIZE
        found = None
        if segment_size <= MAX_PARTIAL_BLOCK_SIZE:             # partial_block_manager.cpp:22
            i = bisect.bisect_left(self.keys, segment_size)     # lower_bound, :42
            if i < len(self.keys):
                found = self.vals[i]
                del self.keys[i], self.vals[i]
        if found is None:
            block_id, offset = self.n_blocks, 0                 # AllocateBlock, :33-39
            self.n_blocks += 1
        else:
            block_id, offset = found
        new_size = (segment_size + offset + 7) & ~7             # AlignValue, :58
        left = BLOCK_SIZE - new_size
        if left >= BLOCK_SIZE - MAX_PARTIAL_BLOCK_SIZE:         # :62-65
            j = bisect.bisect_right(self.keys, left)
            self.keys.insert(j, left)
            self.vals.insert(j, (block_id, new_size))
        return block_id, offset


def build_image(ptype: int, type_size: int, segments) -> np.ndarray:
    """segments: list of dicts {count, capacity, width, state, min_factor, max_factor, payload (bytes)}; payload = the
    int_vector data words of a re-packed segment or its count slot-width values.  Returns the image bytes."""
    placer = BlockPlacer()
    ptrs = []
    placed = []
    row = 0
    for s in segments:
        payload = bytes(s["payload"]) if s["count"] else b""
        block_id, offset = INVALID_BLOCK, 0
        if s["count"]:
            block_id, offset = placer.place(len(payload))
            placed.append((block_id, offset, payload))
        comp = 1 if (s["state"] & 8) else 10
        ptrs.append(struct.pack("<QQqIIQQIBBBBQ", row, s["count"], block_id, offset, len(payload),
                                s["min_factor"] & (2**64 - 1), s["max_factor"] & (2**64 - 1), s["capacity"], comp,
                                s["width"], s["state"], type_size, 0))
        row += s["count"]
    n_blocks = placer.n_blocks
    blocks_offset = (64 + 64 * len(segments) + 4095) & ~4095
    total = blocks_offset + n_blocks * BLOCK_ALLOC_SIZE
    table = b"".join(ptrs)

    def header(meta_checksum):
        return struct.pack("<8sIIQQQQQQ", MAGIC, VERSION, ptype, len(segments), n_blocks, row, blocks_offset, total,
                           meta_checksum)

    meta = checksum(header(0) + table)
    img = np.zeros(total, dtype=np.uint8)
    head = header(meta) + table
    img[: len(head)] = np.frombuffer(head, dtype=np.uint8)
    for block_id, offset, payload in placed:
        at = blocks_offset + block_id * BLOCK_ALLOC_SIZE + 8 + offset
        img[at: at + len(payload)] = np.frombuffer(payload, dtype=np.uint8)
    for b in range(n_blocks):
        at = blocks_offset + b * BLOCK_ALLOC_SIZE
        img[at: at + 8] = np.frombuffer(struct.pack("<Q", checksum(img[at + 8: at + BLOCK_ALLOC_SIZE])), dtype=np.uint8)
    return img


def parse_image(img: np.ndarray):
    """(header dict, list of pointer dicts); verifies the metadata and block checksums (raises ValueError)."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    magic, version, ptype, n_seg, n_blocks, n_values, blocks_offset, total, meta = struct.unpack(
        "<8sIIQQQQQQ", img[:64].tobytes())
    if magic != MAGIC or version != VERSION or total != img.size:
        raise ValueError("bad header")
    table = img[64: 64 + 64 * n_seg].tobytes()
    if checksum(struct.pack("<8sIIQQQQQQ", magic, version, ptype, n_seg, n_blocks, n_values, blocks_offset, total, 0)
                + table) != meta:
        raise ValueError("metadata checksum mismatch")
    ptrs = []
    for g in range(n_seg):
        f = struct.unpack("<QQqIIQQIBBBBQ", table[64 * g: 64 * g + 64])
        ptrs.append(dict(zip(("row_start", "tuple_count", "block_id", "offset", "payload_bytes", "min_factor",
                              "max_factor", "capacity", "compression_type", "width", "state", "type_size",
                              "reserved"), f)))
    for b in range(n_blocks):
        at = blocks_offset + b * BLOCK_ALLOC_SIZE
        stored = struct.unpack("<Q", img[at: at + 8].tobytes())[0]
        if checksum(img[at + 8: at + BLOCK_ALLOC_SIZE]) != stored:
            raise ValueError(f"block {b}: checksum mismatch")
    hdr = dict(ptype=ptype, n_segments=n_seg, n_blocks=n_blocks, n_values=n_values, blocks_offset=blocks_offset,
               total_bytes=total)
    return hdr, ptrs


def payload_of(img: np.ndarray, hdr, p) -> bytes:
    if p["block_id"] < 0:
        return b""
    at = hdr["blocks_offset"] + p["block_id"] * BLOCK_ALLOC_SIZE + 8 + p["offset"]
    return img[at: at + p["payload_bytes"]].tobytes()
