"""CPU checks of the workload generators behind the secondary benchmarks (Zipf look-ups, TPC-H lineitem)."""
import numpy as np

from duckdb_adaptive_compression_b200 import tpch
from duckdb_adaptive_compression_b200.zipf import Zipf, zipf_keys


def test_mt19937_engine_is_the_std_one():
    # std::mt19937 seeded with 42 yields 1608637542 first (init_genrand seeding), 10000th output of the default
    # seed 5489 is 4123659995 (the C++ standard's check value)
    assert int(np.random.RandomState(42).randint(0, 1 << 32, dtype=np.uint64)) == 1608637542
    assert int(np.random.RandomState(5489).randint(0, 1 << 32, size=10000, dtype=np.uint64)[-1]) == 4123659995


def test_zipf_keys_are_deterministic_and_skewed():
    a = zipf_keys(1_000_000, 1.0, 2000, seed=42)
    b = zipf_keys(1_000_000, 1.0, 2000, seed=42)
    assert np.array_equal(a, b) and a.min() >= 1 and a.max() <= 1_000_000
    # q = 1: P(k = 1) = 1 / H_n ~ 1 / 14.4; P(k <= 10) ~ 2.93 / 14.4
    assert 0.04 < np.mean(a == 1) < 0.10 and 0.15 < np.mean(a <= 10) < 0.26
    u = zipf_keys(1000, 0.0, 4000, seed=7)  # q = 0 degenerates to the uniform distribution
    assert 450 < u.mean() < 550
    steep = zipf_keys(1_000_000, 2.0, 2000, seed=3)
    assert np.mean(steep == 1) > 0.55  # 1 / zeta(2) = 0.608
    z = Zipf(10, 1.0, seed=1)
    assert all(1 <= z() <= 10 for _ in range(200))


def test_lineitem_generator_shapes_and_widths():
    cols = tpch.lineitem(2, scale_factor=0.1, seed=42)
    n = 2 * 122880
    assert all(len(v) == n and v.dtype == tpch.INT_COLUMNS[k] for k, v in cols.items())
    assert cols["l_quantity"].min() == 1 and cols["l_quantity"].max() == 50          # 6 bits
    assert cols["l_discount"].min() == 0 and cols["l_discount"].max() == 10          # 4 bits
    assert cols["l_tax"].min() == 0 and cols["l_tax"].max() == 8                     # 4 bits
    assert cols["l_shipdate"].min() > tpch.D_1992_01_01 and cols["l_shipdate"].max() <= tpch.D_1998_08_02 + 121
    assert (cols["l_receiptdate"] > cols["l_shipdate"]).all()
    assert (np.diff(cols["l_orderkey"].astype(np.int64)) >= 0).all()                 # clustered on the order key
    assert set(np.unique(cols["l_rfls"])) <= {0, 1, 2, 3}
    price = cols["l_extendedprice"] // cols["l_quantity"]
    assert price.min() >= 90000 and price.max() <= 90000 + 20000 + 99900             # retail price in cents
    # deterministic per (seed, row group): any rank can mint its own shard
    again = tpch.lineitem_row_group(1, 0.1, 42)
    assert all(np.array_equal(again[k], cols[k][122880:]) for k in cols)
    rev, rows = tpch.q6_expected(cols)
    assert 0.012 < rows / n < 0.028 and rev > 0                                      # TPC-H Q6 selects ~1.9 %
    q1 = tpch.q1_expected(cols)
    assert sum(g["count"] for g in q1.values()) == int((cols["l_shipdate"] <= tpch.D_1998_09_02).sum())


def test_zipf_keys_match_the_reference_sampler():
    """golden: keys drawn by the reference's own zipf.cpp + std::mt19937 (l2_driver zipfkeys, tests/golden/zipf_keys.json)"""
    import json
    import os
    from duckdb_adaptive_compression_b200.zipf import zipf_keys_scalar
    gold = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "zipf_keys.json")))["keys"]
    for spec, want in gold.items():
        n, q, c, seed = spec.split(",")
        got = zipf_keys(int(n), float(q), int(c), seed=int(seed))
        assert got.tolist() == want, spec
        assert zipf_keys_scalar(int(n), float(q), int(c), seed=int(seed)).tolist() == want
