"""Small calls through every kernel family, for compute-sanitizer (memcheck / racecheck / initcheck):
    compute-sanitizer --tool memcheck python tools/sanitize_smoke.py
Sizes are small on purpose (the tools slow kernels down 10-100x); every call's counts and sums are checked against numpy."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cudf_b200 as cb  # noqa: E402
from cudf_b200 import aggregation as agg  # noqa: E402


def run(name, keys, vals, aggs, options=None, include=False):
    dk = [cb.Column.from_numpy(k, m) for k, m in keys]
    dv = cb.Column.from_numpy(*vals)
    gb = cb.GroupBy(cb.Table(dk), cb.NullPolicy.INCLUDE if include else cb.NullPolicy.EXCLUDE, options=options)
    kt, rts = gb.aggregate([cb.GroupByRequest(dv, aggs)])
    torch.cuda.synchronize()
    cnt = rts[0].columns()[-1].to_numpy()[0]
    keep = np.ones(len(vals[0]), bool)
    if not include:
        for k, m in keys:
            if m is not None:
                keep &= m
    assert int(cnt.sum()) == int(keep.sum()), (name, int(cnt.sum()), int(keep.sum()))
    print(f"{name:28s} path={gb.last_path:24s} groups={kt.num_rows()}", flush=True)


def main():
    rng = np.random.default_rng(1)
    n = 200_000
    v64 = rng.integers(-100, 100, n).astype(np.int64)
    vm = rng.random(n) > 0.2
    sc = [agg.sum(), agg.count(cb.NullPolicy.INCLUDE)]
    full = [agg.sum(), agg.min(), agg.max(), agg.mean(), agg.variance(1), agg.count(cb.NullPolicy.INCLUDE)]
    run("tiny-reg", [(rng.integers(0, 4, n).astype(np.int32), None)], (v64, None), sc)
    run("tiny", [(rng.integers(0, 20, n).astype(np.int32), None)], (rng.standard_normal(n), vm), full)
    run("smem-direct (TMA ring)", [(rng.integers(0, 10_000, n).astype(np.int64), None)], (v64, None), sc)
    run("smem hashed", [(rng.integers(0, 1 << 40, 5000).astype(np.int64)[rng.integers(0, 5000, n)], None)], (v64, None), sc, {"force_path": 2})
    run("sorted", [(rng.integers(0, 2000, n).astype(np.int32), None)], (rng.standard_normal(n), vm), full)
    run("sorted-hashed", [(rng.integers(0, 1 << 40, 900).astype(np.int64)[rng.integers(0, 900, n)], None)], (rng.standard_normal(n), vm), full, {"force_path": 8})
    run("global", [(rng.integers(0, 50_000, n).astype(np.int64), rng.random(n) > 0.1)], (v64, vm), full, {"force_path": 1}, include=True)
    run("global 128-bit key", [(rng.integers(-(1 << 62), 1 << 62, 3000).astype(np.int64)[rng.integers(0, 3000, n)], None),
                               (rng.integers(0, 1 << 50, n).astype(np.int64) >> 40, None)], (v64, vm), full, {"force_path": 1})
    wide = [rng.integers(-(1 << 62), 1 << 62, 700).astype(np.int64) for _ in range(3)]
    pick = rng.integers(0, 700, n)
    run("global row keys", [(w[pick], None) for w in wide], (v64, vm), full)
    run("records form (hc), lean", [(rng.integers(0, 60_000, n).astype(np.int64), None)], (v64, None), sc, {"force_path": 7})
    run("records form (hc), general", [(rng.integers(0, 60_000, n).astype(np.int64), None)], (rng.standard_normal(n), vm), full, {"force_path": 7})
    run("records form, 2 levels", [(rng.integers(0, 150_000, n).astype(np.int64), None)], (v64, None), sc, {"force_path": 7, "bucket_slice_log2": 6})
    run("column-form buckets", [(rng.integers(0, 60_000, n).astype(np.int64), rng.random(n) > 0.05)], (v64, vm), full, {"force_path": 7}, include=True)
    run("partitioned", [(rng.integers(0, 60_000, n).astype(np.int64), None)], (v64, vm), full, {"force_path": 3})
    run("repack", [(rng.integers(0, 300, n).astype(np.int64), None), (rng.integers(-50, 50, n).astype(np.int64), None)], (v64, None), sc)
    # the allocation-free entry point and the bitmask helper
    L = cb._lib.lib()
    vb = torch.from_numpy((rng.random(100_000) > 0.3).astype(np.uint8)).cuda()
    words = torch.zeros((100_000 + 31) // 32, dtype=torch.int32, device="cuda")
    import ctypes
    nc = ctypes.c_int32(0)
    rc = L.bgb_bytes_to_bitmask_noalloc(ctypes.c_void_p(vb.data_ptr()), ctypes.c_int64(vb.numel()), ctypes.c_void_p(words.data_ptr()), ctypes.byref(nc), None)
    assert rc == 0 and nc.value == int((vb == 0).sum().item()), (rc, nc.value)
    print("noalloc bitmask ok", flush=True)
    print("SMOKE OK")


if __name__ == "__main__":
    main()
