"""GPU parity at BASELINE.json's full size (configs[1]: 2^27 INT64 values per column, DuckDB segmentation).

The oracle cannot chew 1 GiB columns in seconds, so these tests use size-independent properties plus oracle checks
on a sample of segments: the generator is re-stated in numpy (checksum of the input), encode -> decode and
encode -> uncompact are round trips, the fused SUM equals the sum of the decoded vectors and of the generator,
re-compacting is idempotent, and the payload words / metadata of sampled segments equal the oracle's bit for bit."""
import numpy as np
import pytest

import duckdb_adaptive_compression_b200 as adacom
from tests.cases import splitmix64

pytestmark = pytest.mark.gpu

LOG2_N = 27
SEED0 = 0xADAC0000


def generator(b: int, n: int) -> np.ndarray:
    v = splitmix64(np.arange(n, dtype=np.uint64) + np.uint64(SEED0 + b))
    if b < 64:
        v &= np.uint64((1 << b) - 1)
    return v


@pytest.mark.parametrize("b", [1, 13, 32, 33, 63, 64])
def test_full_column_round_trip_and_sampled_payload(gpu_ctx, port_oracle, b):
    import torch
    n = 1 << LOG2_N
    counts = adacom.row_group_segment_counts(n, np.int64)
    assert len(counts) == 4370
    col = gpu_ctx.column(np.int64, counts)
    col.fill_splitmix(SEED0 + b, 0, (1 << b) - 1)
    want = generator(b, n)
    if b == 64:  # signed values: split into 32-bit halves so that numpy sums exactly
        sv = want.view(np.int64)
        want_sum = (int((sv >> 32).sum()) << 32) + int((sv & 0xFFFFFFFF).sum())
    else:
        want_sum = (int((want >> np.uint64(32)).sum(dtype=np.uint64)) << 32) + int((want & np.uint64(0xFFFFFFFF)).sum(dtype=np.uint64))
    flat = torch.empty(n, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()

    col.compact()
    metas = col.meta()
    if b < 64:
        # every segment of 24k+ uniform b-bit values needs exactly b bits (the column ends with a 1-value segment)
        assert all(m.width == b for m in metas if m.count > 1000), sorted({(m.width, m.count) for m in metas})
        assert all(m.state & adacom.SEG_REPACKED for m in metas)
        assert col.data_size() == sum(9 + 8 * ((m.count * m.width + 63) // 64) for m in metas)
    else:
        # Q1 domain: mixed signs -> range >= 2^63 -> width 64 -> stays at slot width (the 1-value tail segment packs)
        assert not any(m.state & adacom.SEG_REPACKED for m in metas if m.count > 1000)

    # encode -> decode round trip (all 65 552 vectors) == generator, checked on the device
    col.decode(device_out=flat.data_ptr())
    gpu_ctx.synchronize()
    ref = torch.from_numpy(want.view(np.int64)).cuda()
    assert torch.equal(flat, ref), "decoded vectors differ from the generator"

    # fused SUM == sum of the generator (exact 128-bit)
    agg = col.scan_sum()
    assert agg.count == n and agg.sum == want_sum

    # sampled segments: payload words and metadata equal the oracle's, bit for bit
    starts = [int(x) for x in np.concatenate(([0], np.cumsum(counts)))]
    for g in (0, 3, 2184, 4368, 4369):
        seg = port_oracle.encode(want.view(np.int64)[starts[g]:starts[g + 1]], 0)
        om = seg.meta()
        m = metas[g]
        assert (m.count, m.width, m.min_factor, m.max_factor, m.size_in_bytes) == \
               (om.count, om.width, om.min_factor, om.max_factor, om.data_size)
        if om.repacked:
            assert np.array_equal(col.segment_words(g), seg.words())

    # idempotence: compacting again changes nothing; uncompact restores the raw arena
    words_before = col.packed_words
    col.compact()
    assert col.packed_words in (words_before, 0) or b == 64
    col.decode(device_out=flat.data_ptr())
    gpu_ctx.synchronize()
    assert torch.equal(flat, ref)
    col.uncompact()
    assert not any(m.state & adacom.SEG_COMPACTED for m in col.meta())
    col.decode(device_out=flat.data_ptr())
    gpu_ctx.synchronize()
    assert torch.equal(flat, ref), "uncompact did not restore the raw values"
    col.close()
