"""Checkpoint image of a column through the C ABI (adacom_column_checkpoint / adacom_column_restore) against the
CPU restatement: the image bytes are identical, restore gives back a column that scans identically, corruption is
refused.  Bit-exact."""
from __future__ import annotations

import numpy as np
import pytest

import duckdb_adaptive_compression_b200 as adacom
from duckdb_adaptive_compression_b200 import _capi
from oracle import checkpoint as ck
from tests.cases import ALL_DTYPES, width_values

pytestmark = pytest.mark.gpu
M64 = (1 << 64) - 1


def _mixed_column(gpu_ctx, port_oracle, dtype, cap=0):
    """re-packed, slot-width (compacted, width did not shrink), empty and never-compacted segments, ragged counts
    so that narrow types start segments off 8-byte boundaries"""
    S = np.dtype(dtype).itemsize
    bits = 8 * S
    full = ck.BLOCK_SIZE // S
    specs = [(3, 2049, 5), (bits, 777, 0), (5, 0, 0), (bits - 1, full, 0), (2, 65, 0), (1, 1, 9), (bits, 1001, 0),
             (min(9, bits - 1), 63, 3), (bits, 5, 0), (min(17, bits - 1), min(40001, full) if S > 1 else 4001, 11)]
    parts = [width_values(dtype, w, n, b, seed=w * 31 + n) if n else np.zeros(0, dtype) for (w, n, b) in specs]
    counts = [len(p) for p in parts]
    col = gpu_ctx.column_from_values(np.concatenate(parts), counts, cap)
    never = {4, 8}  # left un-compacted (fresh slot-width segments)
    for g in range(len(parts)):
        if g not in never:
            col.compact(g, g + 1)
    segs = []
    for g, p in enumerate(parts):
        if len(p) and g not in never:
            o = port_oracle.encode(p, 0)
            m = o.meta()
            segs.append(dict(count=len(p), capacity=full, width=m.width, state=1 | (2 if m.repacked else 0) | 4,
                             min_factor=m.min_factor, max_factor=m.max_factor,
                             payload=o.words().tobytes() if m.repacked else p.tobytes()))
        else:
            # never compacted: fresh slot-width segment; an empty segment that went through compact() was analyzed
            # (ADACOM_SEG_ANALYZED) and left alone (Compact() returns early, column_segment.cpp:278)
            segs.append(dict(count=len(p), capacity=full, width=bits, state=0 if g in never else 4, min_factor=M64,
                             max_factor=0, payload=p.tobytes()))
    return col, parts, segs


def _diff_report(img, want):
    hdr, ptrs = ck.parse_image(img)
    _, wptrs = ck.parse_image(want)
    lines = ["checkpoint image differs from the CPU restatement"]
    for g, (a, b) in enumerate(zip(ptrs, wptrs)):
        if a != b:
            lines.append(f"pointer {g}: got {a} want {b}")
    at = int(np.flatnonzero(img != want)[0])
    lines.append(f"first differing byte {at} (blocks_offset {hdr['blocks_offset']}): got {img[at]} want {want[at]}")
    return "\n".join(lines)


@pytest.mark.parametrize("dtype", ALL_DTYPES)
def test_image_matches_oracle_and_restores(gpu_ctx, port_oracle, dtype):
    col, parts, segs = _mixed_column(gpu_ctx, port_oracle, dtype)
    S = np.dtype(dtype).itemsize
    img = col.checkpoint()
    want = ck.build_image(_capi.__dict__[np.dtype(dtype).name.upper()], S, segs)
    assert img.size == col.checkpoint_size() == want.size
    hdr, ptrs = ck.parse_image(img)  # all checksums hold
    assert np.array_equal(img, want), _diff_report(img, want)
    values = np.concatenate(parts)
    back = gpu_ctx.restore(img)
    assert back.dtype == np.dtype(dtype)
    for a, b in zip(col.meta(), back.meta()):
        assert (a.row_start, a.count, a.words, a.min_factor, a.max_factor, a.size_in_bytes, a.width, a.state,
                a.capacity) == (b.row_start, b.count, b.words, b.min_factor, b.max_factor, b.size_in_bytes, b.width,
                                b.state, b.capacity)
    assert np.array_equal(back.decode(), values)
    assert np.array_equal(back.decode(), col.decode())
    g0, g1 = col.scan_sum(), back.scan_sum()
    assert (g0.sum, g0.count) == (g1.sum, g1.count)
    assert back.data_size() == col.data_size()
    # the restored column checkpoints to the same bytes, and keeps working through an uncompact / re-compact cycle
    assert np.array_equal(back.checkpoint(), img)
    back.uncompact()
    back.compact()
    assert np.array_equal(back.decode(), values)
    back.close()
    col.close()


def test_device_image_round_trip(gpu_ctx, port_oracle):
    import torch
    col, parts, _ = _mixed_column(gpu_ctx, port_oracle, np.int32)
    n = col.checkpoint_size()
    buf = torch.zeros(n + 64, dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    wrote = col.checkpoint(device_out=buf.data_ptr(), capacity=n)
    assert wrote == n
    host = col.checkpoint()
    assert np.array_equal(buf[:n].cpu().numpy(), host)
    back = gpu_ctx.restore(buf.data_ptr(), dtype=np.int32, device_bytes=n)
    assert np.array_equal(back.decode(), np.concatenate(parts))
    back.close()
    col.close()


def test_corrupt_images_are_refused(gpu_ctx, port_oracle):
    col, parts, _ = _mixed_column(gpu_ctx, port_oracle, np.uint16)
    img = col.checkpoint()
    hdr, ptrs = ck.parse_image(img)
    first_block = hdr["blocks_offset"]
    last_block = hdr["blocks_offset"] + (hdr["n_blocks"] - 1) * ck.BLOCK_ALLOC_SIZE
    for at in (first_block + 8 + 1000,       # payload byte
               first_block + 2,              # stored checksum
               last_block + ck.BLOCK_ALLOC_SIZE - 1,  # zero tail of the last block
               64 + 8,                       # pointer table
               20):                          # header
        bad = img.copy()
        bad[at] ^= 0x01
        with pytest.raises(adacom.AdacomError) as e:
            gpu_ctx.restore(bad)
        assert e.value.status == _capi.ERR_CORRUPT, (at, str(e.value))
    with pytest.raises(adacom.AdacomError) as e:
        gpu_ctx.restore(img[:-8].copy())
    assert e.value.status == _capi.ERR_CORRUPT
    with pytest.raises(adacom.AdacomError) as e:
        gpu_ctx.restore(img[:32].copy())
    assert e.value.status == _capi.ERR_CORRUPT
    # capacity too small
    with pytest.raises(adacom.AdacomError) as e:
        import ctypes as C
        out = np.empty(img.size - 1, dtype=np.uint8)
        gpu_ctx.check(gpu_ctx.lib.adacom_column_checkpoint(col.h, out.ctypes.data_as(C.c_void_p), out.nbytes, 0, None))
    assert e.value.status == _capi.ERR_INVALID
    # the context is still usable
    back = gpu_ctx.restore(img)
    assert np.array_equal(back.decode(), np.concatenate(parts))
    back.close()
    col.close()


def test_empty_column_and_block_sharing(gpu_ctx, port_oracle):
    col = gpu_ctx.column(np.int64, [])
    img = col.checkpoint()
    assert img.size == 4096
    back = gpu_ctx.restore(img)
    assert back.n_values == 0 and back.n_segments == 0
    back.close()
    col.close()
    # 40 small segments share blocks under the partial-block policy: far fewer blocks than segments
    parts = [width_values(np.uint32, 1 + g % 20, 3000 + 7 * g, g, seed=g) for g in range(40)]
    col = gpu_ctx.column_from_values(np.concatenate(parts), [len(p) for p in parts])
    col.compact()
    img = col.checkpoint()
    hdr, ptrs = ck.parse_image(img)
    placer = ck.BlockPlacer()
    for p in ptrs:
        assert (p["block_id"], p["offset"]) == placer.place(p["payload_bytes"])
    assert hdr["n_blocks"] == placer.n_blocks < 5
    back = gpu_ctx.restore(img)
    assert np.array_equal(back.decode(), np.concatenate(parts))
    back.close()
    col.close()


def test_full_size_round_trip(gpu_ctx):
    """2^27 INT64 values (BASELINE config 2 shape) at 17 bits: image on the device, restore, exact SUM and samples."""
    import torch
    n = 1 << 27
    counts = adacom.row_group_segment_counts(n, np.int64)
    col = gpu_ctx.column(np.int64, counts)
    col.fill_splitmix(0xADAC0000 + 17, 0, (1 << 17) - 1)
    col.compact()
    want = col.scan_sum()
    size = col.checkpoint_size()
    buf = torch.empty(size, dtype=torch.uint8, device="cuda:0")
    torch.cuda.synchronize()
    assert col.checkpoint(device_out=buf.data_ptr(), capacity=size) == size
    back = gpu_ctx.restore(buf.data_ptr(), dtype=np.int64, device_bytes=size)
    got = back.scan_sum()
    assert (got.sum, got.count) == (want.sum, want.count)
    for g in (0, 1, len(counts) // 2, len(counts) - 1):
        assert np.array_equal(back.decode(g, g + 1), col.decode(g, g + 1))
    # a flipped bit deep inside the device image is caught by the block checksum
    buf[size // 2 + 12345] ^= 0x10
    torch.cuda.synchronize()
    with pytest.raises(adacom.AdacomError) as e:
        gpu_ctx.restore(buf.data_ptr(), dtype=np.int64, device_bytes=size)
    assert e.value.status == _capi.ERR_CORRUPT
    back.close()
    col.close()
