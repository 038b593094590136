"""Two int64 key columns with small ranges (a 128-bit image, 14 bits of range), int64 SUM + COUNT: key repacking on (automatic)
against the 128-bit generic kernel (force_path=1). python tools/repack_bench.py [rows]"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cudf_b200 as cb  # noqa: E402
from cudf_b200 import _lib, aggregation as A  # noqa: E402

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000_000
dev = torch.device("cuda:0")
L = _lib.lib()
s = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def gen(seed, mod, add=0):
    t = torch.empty(n, dtype=torch.int64, device=dev)
    _lib.check(L.bgb_gen_uniform_int(C.c_void_p(t.data_ptr()), cb.TypeId.INT64, n, seed, 0, mod, add, s))
    return cb.Column(t)


for name, m1, m2 in (("1e4 groups", 100, 100), ("6 groups", 3, 2), ("1e8 groups", 10_000, 10_000)):
    keys = cb.Table([gen(1, m1, -50), gen(2, m2)])
    reqs = [cb.GroupByRequest(gen(3, 2001, -1000), [A.sum(), A.count()])]
    for fp in (0, 1):
        if fp == 1 and m1 * m2 > 1e6:
            continue  # the generic kernel on 1e8 groups of 128-bit keys takes ~0.2 s per step; not the comparison of interest
        gb = cb.GroupBy(keys, options={"force_path": fp, "profile": 1})
        for _ in range(2):
            gb.aggregate(reqs)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            kt, _ = gb.aggregate(reqs)
        e1.record(); torch.cuda.synchronize()
        print(f"{name:11s} force={fp} {e0.elapsed_time(e1) / 3:8.2f} ms/step  groups={kt.num_rows()}  path={gb.last_path}  "
              + " ".join(f"{k}={v:.2f}" for k, v in gb.last_profile if v > 0.2))
    del keys, reqs
