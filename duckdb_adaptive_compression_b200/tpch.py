"""TPC-H lineitem on the succinct codec: synthetic integer columns + the Q6 / Q1 plans over the fused scan.

BASELINE.json configs 4 and 5 ("TPC-H SF10 lineitem integer columns succinct-compressed, Q1/Q6 scans";
"SF100 row groups sharded across 8 GPUs, Q6 SUM reduced over NCCL").  The reference fills lineitem with
dbgen (extension/tpch/dbgen/dbgen.cpp:395-412: l_orderkey/l_partkey/l_suppkey/l_linenumber/l_quantity INTEGER,
l_extendedprice/l_discount/l_tax DECIMAL(15,2) -> INT64 scaled by 100, dates -> INT32 days since 1970).  dbgen
itself is not part of the hot path; `lineitem_row_group` draws the same value distributions (TPC-H spec 4.2.3)
from a counter-based generator keyed by (seed, row group), so that any rank can mint its own row groups without
a 14 GB file.  Queries: extension/tpch/dbgen/queries/q06.sql and q01.sql.
"""
from __future__ import annotations

import numpy as np

from .codec import ROW_GROUP_SIZE, Context, row_group_segment_counts, scan_columns

# days since 1970-01-01
D_1992_01_01, D_1994_01_01, D_1995_01_01, D_1995_06_17, D_1998_08_02, D_1998_09_02 = 8035, 8766, 9131, 9298, 10440, 10471

INT_COLUMNS = {"l_orderkey": np.int32, "l_partkey": np.int32, "l_suppkey": np.int32, "l_linenumber": np.int32,
               "l_quantity": np.int32, "l_extendedprice": np.int64, "l_discount": np.int64, "l_tax": np.int64,
               "l_shipdate": np.int32, "l_commitdate": np.int32, "l_receiptdate": np.int32,
               # VARCHAR in the reference (not succinct-eligible); supplied as one UTINYINT group code
               # (returnflag, linestatus) -> 0:A/F 1:N/F 2:N/O 3:R/F  (documented deviation)
               "l_rfls": np.uint8}


def lineitem_row_group(rg: int, scale_factor: float = 1.0, seed: int = 42, rows: int = ROW_GROUP_SIZE) -> dict:
    """The integer columns of row group `rg` (`rows` lineitems) as numpy arrays."""
    rng = np.random.default_rng([seed, rg])
    n_orders = rows  # upper bound: every order has >= 1 line
    lines = rng.integers(1, 8, size=n_orders)
    first = np.concatenate(([0], np.cumsum(lines)))
    n_orders = int(np.searchsorted(first, rows, side="left"))
    order_of = np.repeat(np.arange(n_orders), lines[:n_orders])[:rows]
    linenumber = (np.arange(rows) - first[order_of] + 1).astype(np.int32)
    # sparse order keys: 8 used out of every 32 (TPC-H 4.2.3), monotone over the row groups
    o_idx = rg * rows + np.arange(n_orders, dtype=np.int64)
    orderkey = ((o_idx // 8) * 32 + (o_idx % 8) + 1)[order_of]
    orderdate = rng.integers(D_1992_01_01, D_1998_08_02 + 1, size=n_orders)[order_of]
    n_part = max(int(200000 * scale_factor), 1)
    n_supp = max(int(10000 * scale_factor), 1)
    partkey = rng.integers(1, n_part + 1, size=rows)
    supp_i = rng.integers(0, 4, size=rows)
    suppkey = (partkey + supp_i * (n_supp // 4 + (partkey - 1) // n_supp)) % n_supp + 1
    quantity = rng.integers(1, 51, size=rows)
    retail = 90000 + ((partkey // 10) % 20001) + 100 * (partkey % 1000)  # cents
    discount = rng.integers(0, 11, size=rows)
    tax = rng.integers(0, 9, size=rows)
    shipdate = orderdate + rng.integers(1, 122, size=rows)
    commitdate = orderdate + rng.integers(30, 91, size=rows)
    receiptdate = shipdate + rng.integers(1, 31, size=rows)
    returned = rng.integers(0, 2, size=rows)  # 'R' or 'A' when already received
    code = np.where(receiptdate <= D_1995_06_17, np.where(returned == 1, 3, 0), np.where(shipdate > D_1995_06_17, 2, 1))
    cols = {"l_orderkey": orderkey, "l_partkey": partkey, "l_suppkey": suppkey, "l_linenumber": linenumber,
            "l_quantity": quantity, "l_extendedprice": quantity * retail, "l_discount": discount, "l_tax": tax,
            "l_shipdate": shipdate, "l_commitdate": commitdate, "l_receiptdate": receiptdate, "l_rfls": code}
    return {k: np.ascontiguousarray(v).astype(INT_COLUMNS[k]) for k, v in cols.items()}


def lineitem(n_row_groups: int, scale_factor: float = 1.0, seed: int = 42, row_groups=None, names=None) -> dict:
    """Concatenated columns of the given row groups (default 0 .. n_row_groups-1)."""
    ids = list(range(n_row_groups)) if row_groups is None else list(row_groups)
    names = list(INT_COLUMNS) if names is None else list(names)
    parts = [lineitem_row_group(g, scale_factor, seed) for g in ids]
    return {k: np.concatenate([p[k] for p in parts]) if parts else np.zeros(0, INT_COLUMNS[k]) for k in names}


class LineitemTable:
    """lineitem columns resident on one GPU, succinct-compressed segment by segment like DuckDB would hold them."""

    def __init__(self, ctx: Context, columns: dict, flags: int = 0):
        self.ctx = ctx
        self.n_rows = len(next(iter(columns.values()))) if columns else 0
        self.cols = {}
        for name, values in columns.items():
            counts = row_group_segment_counts(self.n_rows, values.dtype)
            col = ctx.column_from_values(values, counts)
            col.compact(flags=flags)
            self.cols[name] = col

    def close(self):
        for c in self.cols.values():
            c.close()
        self.cols = {}

    def packed_bytes(self, names) -> int:
        return sum(sum(m.words for m in self.cols[n].meta()) * 8 for n in names)

    # ---- Q6: sum(l_extendedprice * l_discount) where shipdate in [1994, 1995), discount in [5, 7], quantity < 24
    Q6_COLUMNS = ("l_shipdate", "l_discount", "l_quantity", "l_extendedprice")

    def q6(self, device_out: int | None = None):
        c = self.cols
        terms = [(c["l_shipdate"], (D_1994_01_01, D_1995_01_01 - 1)), (c["l_discount"], (5, 7)),
                 (c["l_quantity"], (-(1 << 31), 23)), (c["l_extendedprice"], None)]
        res = scan_columns(self.ctx, terms, [[(3, 0, 1), (1, 0, 1)]], 0, self.n_rows, device_out=device_out)
        return None if res is None else res[0]

    # ---- Q1: per (returnflag, linestatus) group, shipdate <= 1998-09-02
    Q1_COLUMNS = ("l_shipdate", "l_rfls", "l_quantity", "l_extendedprice", "l_discount", "l_tax")
    Q1_SUMS = ("sum_qty", "sum_base_price", "sum_disc_price", "sum_charge", "sum_disc")

    Q1_SUM_EXPRS = [[(2, 0, 1)], [(3, 0, 1)], [(3, 0, 1), (4, 100, -1)], [(3, 0, 1), (4, 100, -1), (5, 100, 1)], [(4, 0, 1)]]

    def _q1_terms(self, code=None):
        c = self.cols
        return [(c["l_shipdate"], (-(1 << 31), D_1998_09_02)), (c["l_rfls"], None if code is None else (code, code)),
                (c["l_quantity"], None), (c["l_extendedprice"], None), (c["l_discount"], None), (c["l_tax"], None)]

    def q1(self, device_out: int | None = None) -> dict:
        """{group code: {sum_qty, sum_base_price, sum_disc_price (x100^1 scale 4), sum_charge (scale 6), sum_disc,
        count}} -- integer sums at DuckDB's DECIMAL scales; the averages are sum / count.  ONE fused pass: the
        (returnflag, linestatus) code is the GROUP BY key of the scan (4 groups x 5 sums)."""
        res = scan_columns(self.ctx, self._q1_terms(), self.Q1_SUM_EXPRS, 0, self.n_rows, group_by=(1, 4),
                           device_out=device_out)
        if res is None:
            return None
        return {code: dict(zip(self.Q1_SUMS, (r.sum for r in grp)), count=grp[0].count)
                for code, grp in enumerate(res) if grp[0].count}

    def q1_per_group_passes(self) -> dict:
        """The same answer from one filtered pass per group (the group code as a fourth predicate)."""
        out = {}
        for code in range(4):
            res = scan_columns(self.ctx, self._q1_terms(code), self.Q1_SUM_EXPRS, 0, self.n_rows)
            if res[0].count:
                out[code] = dict(zip(self.Q1_SUMS, (r.sum for r in res)), count=res[0].count)
        return out


def q6_expected(cols: dict) -> tuple[int, int]:
    """(revenue at scale 4, matching rows) straight from the uncompressed columns -- what DuckDB's filter + SUM
    computes on the decoded vectors (column_segment.cpp:575-658 + SUM over HUGEINT)."""
    sel = ((cols["l_shipdate"] >= D_1994_01_01) & (cols["l_shipdate"] < D_1995_01_01) & (cols["l_discount"] >= 5) &
           (cols["l_discount"] <= 7) & (cols["l_quantity"] < 24))
    p = cols["l_extendedprice"][sel].astype(object)
    d = cols["l_discount"][sel].astype(object)
    return int((p * d).sum()) if sel.any() else 0, int(sel.sum())


def q1_expected(cols: dict) -> dict:
    out = {}
    base = cols["l_shipdate"] <= D_1998_09_02
    for code in range(4):
        sel = base & (cols["l_rfls"] == code)
        if not sel.any():
            continue
        q = cols["l_quantity"][sel].astype(object)
        p = cols["l_extendedprice"][sel].astype(object)
        d = cols["l_discount"][sel].astype(object)
        t = cols["l_tax"][sel].astype(object)
        out[code] = dict(sum_qty=int(q.sum()), sum_base_price=int(p.sum()), sum_disc_price=int((p * (100 - d)).sum()),
                         sum_charge=int((p * (100 - d) * (100 + t)).sum()), sum_disc=int(d.sum()), count=int(sel.sum()))
    return out
