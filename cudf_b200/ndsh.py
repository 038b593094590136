"""NDS-H (TPC-H derived) query 1 over the hash groupby: the caller of SURVEY.md section 8(f) row 4.

Mirrors the reference's driver /root/reference/cpp/benchmarks/ndsh/q01.cpp:101-152 (filter l_shipdate <= 1998-09-02, disc_price and
charge projections, `apply_groupby` on (l_returnflag, l_linestatus) with the eight aggregates of the query,
/root/reference/cpp/benchmarks/ndsh/utilities.cpp:233-274, then order by the two keys). What is NOT this library's path stays thin
and is the caller's business: the two elementwise projections are torch expressions, the final ORDER BY sorts four rows on the host.

One thing is done differently on purpose. The reference filters while reading parquet and then groups the surviving rows; here the
predicate never materialises a filtered table: it becomes the NULL MASK of the first key column (bgb_bytes_to_bitmask) and the
groupby runs with null_policy::EXCLUDE, which drops rows whose key is null (/root/reference/cpp/src/groupby/hash/groupby.cu:41-43).
The 98 % of lineitem that pass the predicate are therefore read once, by the aggregation kernel, instead of being copied first.
(Key columns that carry nulls of their own would need the two masks ANDed; lineitem's do not.)
"""
from __future__ import annotations

import datetime

import numpy as np
import torch

from . import aggregation as agg
from .column import Column, Table, TypeId
from .distributed import column_from_wire
from .groupby import GroupBy, GroupByRequest
from .aggregation import NullPolicy

SHIPDATE_CUTOFF_DAYS = (datetime.date(1998, 9, 2) - datetime.date(1970, 1, 1)).days

RESULT_COLUMNS = ["l_returnflag", "l_linestatus", "sum_base_price", "avg_price", "sum_qty", "avg_qty", "avg_disc", "sum_disc_price",
                  "sum_charge", "count_order"]


def q1(lineitem: dict, shipdate_cutoff_days: int = SHIPDATE_CUTOFF_DAYS, options: dict | None = None):
    """lineitem: {name: Column} with l_returnflag, l_linestatus (int8 codes; the reference's one-character strings are dictionary
    codes to a hash groupby), l_quantity, l_extendedprice, l_discount, l_tax (float64) and l_shipdate (int32 days since epoch:
    TIMESTAMP_DAYS). Returns ({result column name: numpy array}, path) ordered by (l_returnflag, l_linestatus)."""
    rf, ls = lineitem["l_returnflag"], lineitem["l_linestatus"]
    qty, price, disc, tax, ship = (lineitem[k] for k in ("l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_shipdate"))
    for c in (rf, ls, qty, price, disc, tax, ship):
        if c.mask is not None or c.offset != 0:
            raise ValueError("the Q1 driver expects lineitem columns without nulls and with offset 0")
    # projections (q01.cpp:55-98): disc_price = extendedprice * (1 - discount); charge = disc_price * (1 + tax)
    disc_price = price.data * (1.0 - disc.data)
    charge = disc_price * (1.0 + tax.data)
    # predicate -> null mask of the first key (one byte per row -> cudf bitmask words)
    keep = (ship.data <= shipdate_cutoff_days).to(torch.uint8)
    rf_filtered = column_from_wire(rf.data, keep, rf.type_id)
    f64 = TypeId.FLOAT64
    reqs = [
        GroupByRequest(price, [agg.sum(), agg.mean()]),
        GroupByRequest(qty, [agg.sum(), agg.mean()]),
        GroupByRequest(disc, [agg.mean()]),
        GroupByRequest(Column(disc_price, None, 0, f64), [agg.sum()]),
        GroupByRequest(Column(charge, None, 0, f64), [agg.sum(), agg.count(NullPolicy.INCLUDE)]),
    ]
    gb = GroupBy(Table([rf_filtered, ls]), NullPolicy.EXCLUDE, options=options)
    keys, results = gb.aggregate(reqs)
    cols = [c.to_numpy()[0] for c in keys.columns()]
    for t in results:
        cols += [c.to_numpy()[0] for c in t.columns()]
    order = np.lexsort((cols[1], cols[0]))  # ORDER BY l_returnflag, l_linestatus
    return {name: c[order] for name, c in zip(RESULT_COLUMNS, cols)}, gb.last_path


def synthetic_lineitem(n: int, seed: int = 42, device="cuda"):
    """lineitem-shaped columns with the TPC-H value domains of the Q1 inputs: returnflag/linestatus codes
    with the joint weights (A,F) 24.7 %, (N,F) 0.6 %, (N,O) 50 %, (R,F) 24.7 %; quantity 1..50; extendedprice 900..104999; discount
    0.00..0.10; tax 0.00..0.08; shipdate uniform over 1992-01-02 .. 1998-12-01."""
    rng = np.random.default_rng(seed)
    u = rng.random(n)
    combo = np.select([u < 0.247, u < 0.253, u < 0.753], [0, 1, 2], 3)  # (A,F) (N,F) (N,O) (R,F)
    rf = np.array([0, 1, 1, 2], np.int8)[combo]
    ls = np.array([0, 0, 1, 0], np.int8)[combo]
    lo = (datetime.date(1992, 1, 2) - datetime.date(1970, 1, 1)).days
    hi = (datetime.date(1998, 12, 1) - datetime.date(1970, 1, 1)).days
    cols = {
        "l_returnflag": rf, "l_linestatus": ls,
        "l_quantity": rng.integers(1, 51, n).astype(np.float64),
        "l_extendedprice": rng.integers(900, 105000, n).astype(np.float64),
        "l_discount": rng.integers(0, 11, n).astype(np.float64) / 100.0,
        "l_tax": rng.integers(0, 9, n).astype(np.float64) / 100.0,
        "l_shipdate": rng.integers(lo, hi + 1, n).astype(np.int32),
    }
    return {k: Column.from_numpy(v, device=device) for k, v in cols.items()}, cols  # (device columns, the same data as numpy)
