"""NDS-H query 1 (cudf_b200/ndsh.py; reference driver cpp/benchmarks/ndsh/q01.cpp) on a synthetic lineitem of N rows generated on the
device: time of the whole query (projections + predicate mask + groupby + host-side ORDER BY of 4 rows) and of the groupby alone.
    python tools/ndsh_q1_bench.py [rows, default 6e8 = scale factor 100]"""
import datetime
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cudf_b200 as cb  # noqa: E402
from cudf_b200 import ndsh  # noqa: E402
from cudf_b200.column import Column  # noqa: E402


def device_lineitem(n, seed=42):
    g = torch.Generator(device="cuda").manual_seed(seed)
    u = torch.rand(n, device="cuda", generator=g)
    combo = (u >= 0.247).to(torch.int8) + (u >= 0.253).to(torch.int8) + (u >= 0.753).to(torch.int8)
    rf = torch.tensor([0, 1, 1, 2], dtype=torch.int8, device="cuda")[combo.long()]
    ls = torch.tensor([0, 0, 1, 0], dtype=torch.int8, device="cuda")[combo.long()]
    del u, combo
    lo = (datetime.date(1992, 1, 2) - datetime.date(1970, 1, 1)).days
    hi = (datetime.date(1998, 12, 1) - datetime.date(1970, 1, 1)).days
    r = lambda a, b: torch.randint(a, b, (n,), device="cuda", generator=g)  # noqa: E731
    cols = {"l_returnflag": rf, "l_linestatus": ls, "l_quantity": r(1, 51).double(), "l_extendedprice": r(900, 105000).double(),
            "l_discount": r(0, 11).double() / 100.0, "l_tax": r(0, 9).double() / 100.0, "l_shipdate": r(lo, hi + 1).int()}
    return {k: Column(v) for k, v in cols.items()}


def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 600_000_000
    li = device_lineitem(n)
    torch.cuda.synchronize()
    for _ in range(3):
        res, path = ndsh.q1(li)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    steps = 5
    for _ in range(steps):
        res, path = ndsh.q1(li)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / steps * 1e3
    # the groupby alone, on prepared columns (profile on: per-kernel times)
    disc_price = Column(li["l_extendedprice"].data * (1.0 - li["l_discount"].data))
    charge = Column(disc_price.data * (1.0 + li["l_tax"].data))
    from cudf_b200 import aggregation as agg
    from cudf_b200.distributed import column_from_wire
    keep = (li["l_shipdate"].data <= ndsh.SHIPDATE_CUTOFF_DAYS).to(torch.uint8)
    rf = column_from_wire(li["l_returnflag"].data, keep, li["l_returnflag"].type_id)
    reqs = [cb.GroupByRequest(li["l_extendedprice"], [agg.sum(), agg.mean()]), cb.GroupByRequest(li["l_quantity"], [agg.sum(), agg.mean()]),
            cb.GroupByRequest(li["l_discount"], [agg.mean()]), cb.GroupByRequest(disc_price, [agg.sum()]),
            cb.GroupByRequest(charge, [agg.sum(), agg.count(cb.NullPolicy.INCLUDE)])]
    gb = cb.GroupBy(cb.Table([rf, li["l_linestatus"]]), cb.NullPolicy.EXCLUDE, options={"profile": 1})
    for _ in range(3):
        gb.aggregate(reqs)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        gb.aggregate(reqs)
    e1.record()
    torch.cuda.synchronize()
    gms = e0.elapsed_time(e1) / steps
    bytes_gb = n * (1 + 1 + 0.125 + 5 * 8)
    print(f"NDS-H Q1, {n:.3g} rows: whole query {ms:.2f} ms/run ({n / ms / 1e6:.1f} G rows/s... {n / (ms * 1e-3):.3g} rows/s), path={path}")
    print(f"  groupby alone {gms:.2f} ms ({bytes_gb / (gms * 1e-3) / 1e9:.0f} GB/s of the {bytes_gb / 1e9:.1f} GB it reads once), kernels: "
          + ", ".join(f"{k}={v:.3f}" for k, v in gb.last_profile))
    for k in ndsh.RESULT_COLUMNS:
        print(f"  {k:16s} {res[k]}")


if __name__ == "__main__":
    main()
